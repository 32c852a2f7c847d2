#!/usr/bin/env python3
"""LDE time of a 2^23 -> 2^24 matrix of NCOLS columns as a function of the scratch column block (GLB_SCRATCH_MIB sets the limit).
usage: python tools/lde_sweep.py NCOLS CB [CB ...]"""
import os
import subprocess
import sys

ncols = int(sys.argv[1])
per_col = ((1 << 23) + (1 << 24)) * 8
for cb in [int(x) for x in sys.argv[2:]]:
    env = dict(os.environ, GLB_SCRATCH_MIB=str((per_col * cb >> 20) + 1))
    r = subprocess.run([sys.executable, os.path.join(os.path.dirname(os.path.abspath(__file__)), "prof_lde.py"), "23", str(ncols), "4"], capture_output=True, text=True, env=env)
    ms = [float(l.split("passes): ")[1].split(" ms")[0]) for l in r.stdout.splitlines() if "passes): " in l]
    print("ncols %d  column block %3d  LDE %s ms (best %.2f)" % (ncols, cb, " ".join("%.2f" % m for m in ms), min(ms) if ms else -1), flush=True)
