#!/bin/bash
# Multi-GPU call: parity test (small shapes every word + the c2 commit at full size), strong-scaling bench and the reference arm under torchrun.
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo_n$N.txt 2>&1
timeout 1700 python -m pytest tests/test_gpu_mgpu.py -m gpu -q -s > gpurun_out/pytest_mgpu_n$N.txt 2>&1; echo "pytest mgpu rc=$?"; tail -12 gpurun_out/pytest_mgpu_n$N.txt
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29701 bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/bench_c2_n$N.json 2> gpurun_out/bench_c2_n$N.err; echo "bench rc=$?"; tail -2 gpurun_out/bench_c2_n$N.err
python -c "import json;d=json.load(open('gpurun_out/bench_c2_n$N.json'));print(d['n_gpus'],d['ms_per_step'],d['phases'],d['roots_agree'],d['root']);print(d['e2e'])"
( time timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29702 bench.py --impl reference --gpus $N --steps 3 --warmup 1 > gpurun_out/bench_ref_n$N.json 2> gpurun_out/bench_ref_n$N.err ) 2>&1 | tail -3; echo "ref rc=$?"; cut -c1-600 gpurun_out/bench_ref_n$N.json
