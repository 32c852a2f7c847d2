"""Multi-GPU parity (needs >= 2 GPUs on the box; run with `gpurun --gpus 2 -- python -m pytest tests/test_gpu_mgpu.py -m gpu`)."""
import os
import subprocess
import sys

import pytest

from conftest import ROOT


@pytest.mark.gpu
def test_commit_mgpu_matches_oracle(glb):
    ngpu = glb.device_count()
    if ngpu < 2:
        pytest.skip("needs >= 2 GPUs (single-GPU parity is covered by test_gpu_parity.py)")
    world = 8 if ngpu >= 8 else (4 if ngpu >= 4 else 2)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world), "--master-addr", "127.0.0.1",
           "--master-port", "29655", os.path.join(ROOT, "tests", "mgpu_worker.py"), "nccl"]
    # GLB_MGPU_FULL=1: after the small shapes (every word against the oracle) also the c2 commit (BASELINE configs[2]) at full size:
    # sampled LDE columns against the CPU reference, the row panels, and the root pinned by the single-GPU full-size test
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=1500, cwd=ROOT, env=dict(os.environ, GLB_MGPU_FULL="1"))
    print(r.stdout[-1500:])
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    assert r.stdout.count("ok") == world
