"""Profiling driver: LDE (extendPol) of a 2^LOGN x NCOLS matrix resident in HBM, timed with CUDA events by the
library; run under ncu to capture the NTT pass kernels.  usage: python tools/prof_lde.py [logn] [ncols] [reps]"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load

glb = load()
glb.init(0)
logn = int(sys.argv[1]) if len(sys.argv) > 1 else 20
ncols = int(sys.argv[2]) if len(sys.argv) > 2 else 64
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
n, e = 1 << logn, 1 << (logn + 1)
src = glb.DeviceBuffer(n * ncols)
dst = glb.DeviceBuffer(e * ncols)
glb.fill_synthetic_dev(src.ptr, n * ncols, seed=3)
smem = os.environ.get("GLB_NTT_SMEM") == "1"
def passes(k):
    return (k + 7) // 8
ep = (n * passes(logn) + e * passes(logn + 1)) * ncols  # element-passes of one LDE
for r in range(reps):
    glb.extend_pol_dev(dst.ptr, src.ptr, e, n, ncols)
    ms = glb.last_timings()["lde_ms"]
    print("LDE 2^%d -> 2^%d x %d (%s passes): %.3f ms, %.1f G element-passes/s, %.0f GB/s of pass traffic" % (
        logn, logn + 1, ncols, "shared-memory radix-256" if smem else "library default", ms, ep / ms / 1e6, ep * 16 / ms / 1e6))
