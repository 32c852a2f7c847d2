#!/bin/bash
# Round 2, GPU call 3: pipe-balanced butterflies (csel) vs the round-1 form (nocsel), both pass kernels; squaring with 3 products in the S-box.
mkdir -p gpurun_out; rm -f gpurun_out/ntt_ab3.txt
for lib in "" "$PWD/tools/libglb200_nocsel.so"; do
 for v in 0 2; do
  echo "== lib=${lib:-csel(default)} GLB_NTT=$v" >> gpurun_out/ntt_ab3.txt
  GLB_LIB=$lib GLB_NTT=$v timeout 120 python tools/prof_lde.py 23 128 5 >> gpurun_out/ntt_ab3.txt 2>&1; echo "rc=$?" >> gpurun_out/ntt_ab3.txt
  GLB_LIB=$lib GLB_NTT=$v timeout 120 python tools/prof_lde.py 20 96 5 >> gpurun_out/ntt_ab3.txt 2>&1
 done
done
cat gpurun_out/ntt_ab3.txt
for n in base sqr3a sqr3b; do echo "$n $(timeout 60 tools/ubench_$n leaf 2>&1 | grep '^leaf' | awk '{print $(NF-2)}' | sort -n | tail -1) M perm/s"; done | tee gpurun_out/leaf_ab3.txt
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu3.txt 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/pytest_gpu3.txt
for v in 0 2; do
  GLB_NTT=$v timeout 300 python bench.py --steps 2 --warmup 2 --no-e2e --no-cpu > gpurun_out/bench3_ntt$v.json 2> gpurun_out/bench3_ntt$v.err; echo "bench v=$v rc=$?"
  python -c "import json;d=json.load(open('gpurun_out/bench3_ntt$v.json'));print(d['ms_per_step'],d['phases']['lde']['ms'],d['phases']['lde']['frac'],d['root'])"
done
