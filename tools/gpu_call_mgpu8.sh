#!/bin/bash
N=8
mkdir -p gpurun_out
{ nproc; free -g | head -2; lscpu | grep -E "Model name|Socket|NUMA node\(s\)"; nvidia-smi topo -m | head -12; } > gpurun_out/box_n$N.txt 2>&1
timeout 300 tools/pcie_all > gpurun_out/pcie_all_n$N.txt 2>&1; echo "pcie rc=$?"; cat gpurun_out/pcie_all_n$N.txt
timeout 1700 python -m pytest tests/test_gpu_mgpu.py -m gpu -q -s > gpurun_out/pytest_mgpu_n$N.txt 2>&1; echo "pytest mgpu rc=$?"; tail -14 gpurun_out/pytest_mgpu_n$N.txt
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29701 bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/bench_c2_n$N.json 2> gpurun_out/bench_c2_n$N.err; echo "bench rc=$?"; tail -2 gpurun_out/bench_c2_n$N.err
python -c "import json;d=json.load(open('gpurun_out/bench_c2_n$N.json'));print(d['n_gpus'],d['ms_per_step'],d['phases'],d['roots_agree'],d['root']);print(d['ms_per_rank']);print(d['e2e'])"
