"""A/B of the host-pointer commit's column-block schedule on one box: gl_commit_host on c2 with pinned buffers, per-block timeline
(GLB_HOST_TRACE=1).  Usage: GLB_HOST_BLOCKS=0|1|32 python tools/e2e_ab.py [reps]"""
import ctypes as C
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as ge

glb = ge.load()
glb.init(0)
L = glb.lib()
logn, loge, ncols = 23, 24, 600
n, e = 1 << logn, 1 << loge
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 3
hin, hlde, htree = C.c_void_p(), C.c_void_p(), C.c_void_p()
in_b, lde_b, tree_b = n * ncols * 8, e * ncols * 8, glb.tree_num_elements(e) * 8
for hp, sz in ((hin, in_b), (hlde, lde_b), (htree, tree_b)):
    assert L.gl_host_alloc(C.byref(hp), sz) == 0
tmp = glb.DeviceBuffer(n * ncols)
glb.fill_synthetic_dev(tmp.ptr, n * ncols, 1)
L.gl_d2h(hin, C.c_void_p(tmp.ptr), in_b)
# raw link rates on this box: contiguous 20 GB each way
for name, fn, a, b in (("D2H", L.gl_d2h, hlde, C.c_void_p(tmp.ptr)), ("H2D", L.gl_h2d, C.c_void_p(tmp.ptr), hlde)):
    t0 = time.perf_counter()
    fn(a, b, 20 << 30)
    dt = time.perf_counter() - t0
    print("raw %s 20 GiB: %.1f GB/s" % (name, (20 << 30) / dt / 1e9), flush=True)
del tmp
u64p = C.POINTER(C.c_uint64)
root = np.zeros(4, dtype=np.uint64)
for i in range(reps + 1):
    t0 = time.perf_counter()
    rc = L.gl_commit_host(C.cast(hlde, u64p), C.cast(htree, u64p), C.cast(hin, u64p), e, n, ncols, root.ctypes.data_as(u64p))
    assert rc == 0, L.gl_last_error()
    print("GLB_HOST_BLOCKS=%s gl_commit_host %d: %.1f ms  root %016X" % (os.environ.get("GLB_HOST_BLOCKS", "default"), i, (time.perf_counter() - t0) * 1e3, int(root[0])), flush=True)
