#!/bin/bash
# Round 2, GPU call 16 (N GPUs): copy-engine exchange (CUDA IPC + peer copies) -- parity incl. the full c2 commit, then the bench, A/B against the NCCL transport.
N=${1:-2}
mkdir -p gpurun_out
timeout 500 python -m pytest tests/test_gpu_mgpu.py -m gpu -q -s > gpurun_out/pytest_mgpu16_n$N.txt 2>&1; echo "pytest mgpu rc=$?"; tail -8 gpurun_out/pytest_mgpu16_n$N.txt
for mode in ce nccl; do
GLB_MGPU_EXCHANGE=$mode timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29701 bench.py --gpus $N --steps 3 --warmup 3 --no-e2e > gpurun_out/bench16_c2_n${N}_$mode.json 2> gpurun_out/bench16_c2_n${N}_$mode.err; echo "bench $mode rc=$?"; tail -2 gpurun_out/bench16_c2_n${N}_$mode.err | cut -c1-300
python -c "import json;d=json.load(open('gpurun_out/bench16_c2_n${N}_$mode.json'));print(d['n_gpus'],d['ms_per_step'],{k:round(v['ms'],2) for k,v in d['phases'].items()},d['phases']['exchange'].get('transport','')[:40],d['roots_agree']);print(d['ms_per_rank']['rows'])"
done
