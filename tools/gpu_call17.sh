#!/bin/bash
# Round 2, GPU call 17 (8 GPUs): copy-engine exchange at 8 ranks -- parity incl. the full c2 commit, then the c2 line with the per-rank phase table.
N=${1:-8}
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_gpu_mgpu.py -m gpu -q -s > gpurun_out/pytest_mgpu17_n$N.txt 2>&1; echo "pytest mgpu rc=$?"; tail -12 gpurun_out/pytest_mgpu17_n$N.txt
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29701 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/bench17_c2_n$N.json 2> gpurun_out/bench17_c2_n$N.err; echo "bench rc=$?"; tail -2 gpurun_out/bench17_c2_n$N.err | cut -c1-300
python -c "import json;d=json.load(open('gpurun_out/bench17_c2_n$N.json'));print(d['n_gpus'],d['ms_per_step'],{k:round(v['ms'],2) for k,v in d['phases'].items()},d['phases']['exchange'].get('transport','')[:40],d['roots_agree']);print(d['ms_per_rank']['rows']);print(d['e2e'])"
