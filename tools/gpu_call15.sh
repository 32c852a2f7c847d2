#!/bin/bash
# Round 2, GPU call 15 (2 GPUs): does the NCCL exchange that runs under the leaf hashing cost SM time?  Same bench with the transport's SM footprint varied.
N=${1:-2}
mkdir -p gpurun_out; rm -f gpurun_out/nccl_ab.txt
run() { # label, env...
  label=$1; shift
  env "$@" timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29701 bench.py --gpus $N --steps 3 --warmup 2 --no-e2e > gpurun_out/nccl_ab_tmp.json 2> gpurun_out/nccl_ab_tmp.err
  echo "== $label rc=$?" >> gpurun_out/nccl_ab.txt
  python -c "import json;d=json.load(open('gpurun_out/nccl_ab_tmp.json'));print(d['ms_per_step'],{k:round(v['ms'],2) for k,v in d['phases'].items()},d['roots_agree'])" >> gpurun_out/nccl_ab.txt 2>&1 || tail -3 gpurun_out/nccl_ab_tmp.err >> gpurun_out/nccl_ab.txt
}
run default X=1
run max_ctas_2 NCCL_MAX_CTAS=2
run max_ctas_1 NCCL_MAX_CTAS=1
run cuda_memcpy NCCL_P2P_USE_CUDA_MEMCPY=1
run nthreads_128_ctas_2 NCCL_MAX_CTAS=2 NCCL_NTHREADS=128
cat gpurun_out/nccl_ab.txt
