#!/bin/bash
# Round-2 A/B of the Poseidon leaf kernel switches (all default-off).
#   tools/ab_round2.sh build   (in the build container: compiles tools/ubench_<variant> for sm_100a)
#   tools/ab_round2.sh run     (on the GPU box via gpurun: M perm/s of every variant -> gpurun_out/ab_round2.txt,
#                               then tools/ubench_fp64 with the real-IMAD.WIDE lines -> gpurun_out/ubench_fp64.txt)
# A variant that wins still has to pass `pytest -m gpu` with the library rebuilt with the same -D flags.
cd "$(dirname "$0")/.."
VARIANTS=("base:" "red_alu1:-DGLB_POSEIDON_RED_ALU=1" "red_alu2:-DGLB_POSEIDON_RED_ALU=2" "stagger200:-DGLB_LEAF_STAGGER=200"
          "stagger400:-DGLB_LEAF_STAGGER=400" "stagger800:-DGLB_LEAF_STAGGER=800" "split1:-DGLB_MDS_SPLIT=1" "split0:-DGLB_MDS_SPLIT=0")
case "$1" in
build)
  for v in "${VARIANTS[@]}"; do
    name=${v%%:*}; flags=${v#*:}
    nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a --use_fast_math $flags -I merlin-zkevm-prover_b200/csrc -o tools/ubench_$name tools/ubench.cu &
  done
  nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tools/ubench_fp64 tools/ubench_fp64.cu &
  wait ;;
run)
  mkdir -p gpurun_out; : > gpurun_out/ab_round2.txt
  for v in "${VARIANTS[@]}"; do
    name=${v%%:*}
    echo "$name $(timeout 40 tools/ubench_$name leaf 2>&1 | grep '^leaf' | awk '{print $(NF-2)}' | sort -n | tail -1) M perm/s" >> gpurun_out/ab_round2.txt
  done
  timeout 30 tools/ubench_fp64 > gpurun_out/ubench_fp64.txt 2>&1
  cat gpurun_out/ab_round2.txt; tail -12 gpurun_out/ubench_fp64.txt ;;
*) echo "usage: $0 build|run"; exit 2 ;;
esac
