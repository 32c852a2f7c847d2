"""Per-kernel totals of an `ncu --metrics gpu__time_duration.sum --csv` launch list.  usage: python tools/launch_summary.py LIST.csv "header line\""""
import collections
import csv
import re
import sys

tot = collections.defaultdict(float)
cnt = collections.Counter()
with open(sys.argv[1]) as f:
    lines = [l for l in f if l.startswith('"')]
for r in csv.DictReader(lines):
    if r.get("Metric Name") != "gpu__time_duration.sum":
        continue
    name = re.sub(r"\(.*$", "", r["Kernel Name"])
    ns = float(r["Metric Value"].replace(",", ""))
    if r.get("Metric Unit") == "us":
        ns *= 1e3
    elif r.get("Metric Unit") == "ms":
        ns *= 1e6
    tot[name] += ns
    cnt[name] += 1
allns = sum(tot.values())
print(sys.argv[2])
for name, ns in sorted(tot.items(), key=lambda kv: -kv[1]):
    print("%-70s n=%4d %14.1f us %6.1f%%" % (name[:70], cnt[name], ns / 1e3, 100 * ns / allns))
