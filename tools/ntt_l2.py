#!/usr/bin/env python3
"""Compute-side ceiling of the NTT pass kernels: the same transform on an L2-resident matrix (no HBM wait) against an
HBM-resident one, in ns per element-pass (one element through one <= 8-bit pass).  Run on the GPU box."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as ge

glb = ge.load()
glb.init(0)
for logn, ncols in [(16, 32), (16, 64), (16, 128), (17, 32), (20, 64), (24, 64), (23, 64), (24, 128)]:
    n = 1 << logn
    d = glb.DeviceBuffer(n * ncols)
    o = glb.DeviceBuffer(n * ncols)
    glb.fill_synthetic_dev(d.ptr, n * ncols, 3)
    npass = 1 if logn <= 8 else (logn + 7) // 8
    best = 1e9
    for rep in range(6):
        glb.ntt_dev(o.ptr, d.ptr, n, ncols)
        ms = glb.last_timings()["lde_ms"]
        if rep:
            best = min(best, ms)
    ep = n * ncols * npass
    print("ntt 2^%d x %3d cols (%7.1f MB, %d passes): %8.3f ms  %6.3f ns/element-pass  %7.1f GB/s of pass traffic" %
          (logn, ncols, n * ncols * 8 / 1e6, npass, best, best * 1e6 / ep, ep * 16 / best / 1e6))
    d.free(); o.free()
