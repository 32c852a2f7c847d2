#!/bin/bash
# A/B of the Poseidon leaf kernel variants (u32->f64 conversion x resident CTAs per SM) on the GPU box:
#   1. tools/ubench_cvtX_mbY leaf  -> M perm/s of each variant
#   2. the full GPU suite on the default library
#   3. the fastest variant's library swapped in: Poseidon / Merkle / commit parity subset + a device-only c2 bench line
# Everything lands in gpurun_out/r1_21_*.
O=gpurun_out; mkdir -p $O
R=$O/r1_21_leaf_variants.txt; : > $R
best=cvt0_mb4; bestv=0
for b in cvt0_mb4 cvt1_mb4 cvt2_mb4 cvt0_mb5 cvt1_mb5 cvt2_mb5 cvt0_mb6 cvt1_mb6; do
  v=$(timeout 40 tools/ubench_$b leaf 2>&1 | grep "^leaf" | awk '{print $(NF-2)}' | sort -n | tail -1)
  echo "$b $v M perm/s (best of 3)" >> $R
  if [ -n "$v" ] && awk "BEGIN{exit !($v > $bestv)}"; then best=$b; bestv=$v; fi
done
echo "best $best $bestv" >> $R
(timeout 240 python -m pytest tests -m gpu -x -q > $O/r1_21_pytest_gpu.txt 2>&1; echo "pytest rc=$?" >> $O/r1_21_pytest_gpu.txt)
if [ "$best" != "cvt0_mb4" ]; then
  cp merlin-zkevm-prover_b200/libglb200.so /tmp/libglb200_default.so
  cp merlin-zkevm-prover_b200/libglb200_$best.so merlin-zkevm-prover_b200/libglb200.so
  (timeout 200 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "poseidon or linear_hash or merkletree or config1 or production or group_proofs or consttree or fri or q_poly" > $O/r1_21_pytest_best_variant.txt 2>&1; echo "pytest rc=$? variant=$best" >> $O/r1_21_pytest_best_variant.txt)
  timeout 120 python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu > $O/r1_21_bench_c2_best_variant.json 2> $O/r1_21_bench_err.txt
  cp /tmp/libglb200_default.so merlin-zkevm-prover_b200/libglb200.so
fi
cat $R; tail -2 $O/r1_21_pytest_gpu.txt; tail -2 $O/r1_21_pytest_best_variant.txt 2>/dev/null; head -c 300 $O/r1_21_bench_c2_best_variant.json 2>/dev/null
