"""Cost of the panel reader of the leaf kernel (k_linear_hash_rows<true>, the multi-GPU path) on ONE GPU: gl_commit_mgpu with world = 1
(one panel, chunked launches) against gl_commit_dev on the same input.  usage: python tools/leaf_panels_ab.py [logn] [ncols]"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as ge

glb = ge.load()
glb.init(0)
logn = int(sys.argv[1]) if len(sys.argv) > 1 else 22
ncols = int(sys.argv[2]) if len(sys.argv) > 2 else 600
n, e = 1 << logn, 1 << (logn + 1)
din = glb.DeviceBuffer(n * ncols)
glb.fill_synthetic_dev(din.ptr, n * ncols, 1)
dlde = glb.DeviceBuffer(e * ncols)
dtree = glb.DeviceBuffer(glb.tree_num_elements(e))
for _ in range(3):
    root = glb.commit_dev(dlde.ptr, dtree.ptr, din.ptr, e, n, ncols)
    print("gl_commit_dev      ", {k: round(v, 2) for k, v in glb.last_timings().items()}, flush=True)
glb.mgpu_init(glb.mgpu_unique_id(), 0, 1)
dpan = glb.DeviceBuffer(e * ncols)
for _ in range(3):
    root2 = glb.commit_mgpu(dlde.ptr, dpan.ptr, dtree.ptr, din.ptr, e, n, ncols)
    print("gl_commit_mgpu (1) ", {k: round(v, 2) for k, v in glb.last_timings4().items()}, flush=True)
assert np.array_equal(root, root2)
glb.mgpu_shutdown()
