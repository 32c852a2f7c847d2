"""Distil an `ncu --set full --import-source on` capture of the leaf kernel (exported with `ncu -i X.ncu-rep --page raw --csv`
and `--page source --csv`) into the text summary kept under profiles/.
usage: python tools/ncu_leaf_summary.py RAW.csv SOURCE.csv "header line" LOG_ROWS PERMS_PER_ROW [MIX.json] > profiles/...txt
With MIX.json the per-class counts per permutation (the classes of tools/sass_cost.py) are also written there for bench.py."""
import collections
import csv
import re
import sys

raw, src, header, log_rows, perms = sys.argv[1], sys.argv[2], sys.argv[3], int(sys.argv[4]), int(sys.argv[5])
rows = list(csv.reader(open(raw)))
d = {h: (v, u) for h, u, v in zip(rows[0], rows[1], rows[2])}
pat = (r"^(Kernel Name|dram__bytes_(read|write)\.sum(\.pct_of_peak_sustained_elapsed|\.per_second)?|gpu__dram_throughput\.avg\.pct_of_peak_sustained_elapsed|"
       r"gpu__time_duration\.sum|launch__(block_size|grid_size|registers_per_thread|registers_per_thread_allocated|occupancy_limit_\w+)|"
       r"sm__inst_executed_pipe_(alu|fma|fp64|xu)\.avg\.pct_of_peak_sustained_active|"
       r"sm__pipe_(alu|fma|fmaheavy|fp64)_cycles_active\.avg\.pct_of_peak_sustained_(active|elapsed)|sm__throughput\.avg\.pct_of_peak_sustained_elapsed|"
       r"sm__warps_active\.avg\.pct_of_peak_sustained_active|smsp__average_warps_issue_stalled_\w+_per_issue_active\.ratio|smsp__inst_executed\.sum|"
       r"smsp__issue_active\.avg\.per_cycle_active)$")
print(header)
print("----")
for k in sorted(d):
    if re.match(pat, k) and d[k][0] not in ("0", ""):
        print(k, "[%s]" % d[k][1], d[k][0])
print("----")
rows = list(csv.reader(open(src)))
hdr = rows[1]
isrc, iex = hdr.index("Source"), hdr.index("Instructions Executed")
c = collections.Counter()
tot = 0
for r in rows[2:]:
    if len(r) <= iex:
        continue
    ins = re.sub(r"^@!?U?P\d+\s+", "", r[isrc].strip()).split()
    n = int(r[iex] or 0)
    c[ins[0].rstrip(";") if ins else "?"] += n
    tot += n
wp = (1 << log_rows) * perms / 32
print("per-opcode warp instructions per permutation (source page, 2^%d rows x %d permutations): total %.0f" % (log_rows, perms, tot / wp))
for k, n in c.most_common(20):
    print("  %-22s %8.1f" % (k, n / wp))

if len(sys.argv) > 6:
    import json
    cls = dict(imad_wide=0.0, imad_other=0.0, fp64=0.0, alu=0.0, i2f=0.0)
    for k, n in c.items():
        b = k.split(".")[0]
        if k.startswith(("IMAD.WIDE", "IMAD.HI")):
            cls["imad_wide"] += n / wp
        elif k.startswith("IMAD"):
            cls["imad_other"] += n / wp
        elif b in ("DFMA", "DADD", "DMUL"):
            cls["fp64"] += n / wp
        elif b in ("IADD3", "LOP3", "SHF", "SEL", "LEA", "ISETP", "PRMT", "MOV", "VIADD", "PLOP3", "IABS", "FSEL", "SGXT"):
            cls["alu"] += n / wp
        elif b in ("I2F", "F2I"):
            cls["i2f"] += n / wp
    rd, wr = float(d["dram__bytes_read.sum"][0].replace(",", "")), float(d["dram__bytes_write.sum"][0].replace(",", ""))
    unit = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    rd *= unit.get(d["dram__bytes_read.sum"][1], 1.0)
    wr *= unit.get(d["dram__bytes_write.sum"][1], 1.0)
    algo = (1 << log_rows) * (perms * 8 * 8 + 32)
    out = dict(source=header, warp_instr_per_perm=tot / wp, dram_bytes_per_algorithmic_byte=(rd + wr) / algo, **cls)
    json.dump(out, open(sys.argv[6], "w"), indent=1)
