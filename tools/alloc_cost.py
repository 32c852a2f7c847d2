"""Cost of the device allocations the host-pointer commit makes per call (cudaMalloc / cudaFree of the 80 GB extended matrix, the
1 GB tree and the staging block), and of one gl_commit_host with and without them in the call.  Usage: python tools/alloc_cost.py"""
import ctypes as C
import sys
import time
import os

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as ge

glb = ge.load()
glb.init(0)
L = glb.lib()
for gb in (1, 8, 40, 80.5):
    b = int(gb * (1 << 30))
    ts = []
    for rep in range(3):
        p = C.c_void_p()
        t0 = time.perf_counter()
        rc = L.gl_dev_alloc(C.byref(p), b)
        t1 = time.perf_counter()
        assert rc == 0, L.gl_last_error()
        L.gl_dev_free(p)
        t2 = time.perf_counter()
        ts.append(((t1 - t0) * 1e3, (t2 - t1) * 1e3))
    print("cudaMalloc / cudaFree of %5.1f GiB: " % gb + "  ".join("%.1f / %.1f ms" % t for t in ts), flush=True)
