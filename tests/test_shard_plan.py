"""Multi-GPU host logic on CPU: world_size-2 (and 4) gloo groups run the column-split -> row-block exchange
on oracle data and assemble the tree with the shard plan; the result must equal the single-process oracle tree."""
import os
import subprocess
import sys

import numpy as np
import pytest

import __graft_entry__ as ge
from conftest import ROOT


def _sp():
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import shard_plan
    return shard_plan


def test_column_and_level_arithmetic():
    sp = _sp()
    for ncols in (0, 1, 3, 37, 600, 665):
        for world in (1, 2, 4, 8):
            edges = [sp.col_start(ncols, world, r) for r in range(world + 1)]
            assert edges[0] == 0 and edges[-1] == ncols and edges == sorted(edges)
            widths = [b - a for a, b in zip(edges, edges[1:])]
            if ncols // 8 >= world:  # balanced in 8-column chunks, every boundary on a chunk, the remainder on the last rank
                assert all(e % 8 == 0 for e in edges[:-1]) and max(widths[:-1] + [widths[-1] - ncols % 8]) - min(widths[:-1] + [widths[-1] - ncols % 8]) <= 8
            else:
                assert max(widths) - min(widths) <= 1
    assert sp.level_offset(8, 0) == 0 and sp.level_offset(8, 1) == 32 and sp.level_offset(8, 3) == 56
    mod = ge.load()
    assert sp.level_offset(1 << 10, 10) == mod.tree_num_elements(1 << 10) - 4
    # the C side uses the same split
    L = mod.lib()
    for ncols in (3, 37, 600, 665):
        for world in (2, 4, 8):
            for r in range(world + 1):
                assert int(L.gl_mgpu_col_start(ncols, world, r)) == sp.col_start(ncols, world, r)


@pytest.mark.parametrize("world", [2, 4])
def test_gloo_exchange_and_tree_assembly(world):
    env = dict(os.environ, OMP_NUM_THREADS="2")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world), "--master-addr", "127.0.0.1",
           "--master-port", str(29611 + world), os.path.join(ROOT, "tests", "mgpu_worker.py"), "gloo"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    assert r.stdout.count("ok") == world
