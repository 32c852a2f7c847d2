#!/bin/bash
# Round 2, GPU call 14 (N GPUs): multi-GPU parity + strong-scaling bench after the wave-aligned exchange chunks and the __grid_constant__ panel table.
N=${1:-2}
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_mgpu.py -m gpu -q -s > gpurun_out/pytest_mgpu14_n$N.txt 2>&1; echo "pytest mgpu rc=$?"; tail -5 gpurun_out/pytest_mgpu14_n$N.txt
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29701 bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/bench14_c2_n$N.json 2> gpurun_out/bench14_c2_n$N.err; echo "bench rc=$?"; tail -2 gpurun_out/bench14_c2_n$N.err
python -c "import json;d=json.load(open('gpurun_out/bench14_c2_n$N.json'));print(d['n_gpus'],d['ms_per_step'],d['phases'],d['roots_agree'],d['root']);print(d['ms_per_rank']['rows']);print(d['e2e'])"
