#!/bin/bash
# Round 2, GPU call 2: correctness of the TMA-fed k_pass2 passes (full GPU suite) and A/B timing against the round-1 k_pass.
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active --format=csv -lms 500 > gpurun_out/clocks2.csv &
SMI=$!
for v in 0 2 3; do
  echo "== GLB_NTT=$v" >> gpurun_out/ntt_ab.txt
  GLB_NTT=$v timeout 120 python tools/prof_lde.py 23 128 3 >> gpurun_out/ntt_ab.txt 2>&1; echo "rc=$?" >> gpurun_out/ntt_ab.txt
  GLB_NTT=$v timeout 120 python tools/prof_lde.py 20 96 3 >> gpurun_out/ntt_ab.txt 2>&1
  GLB_NTT=$v timeout 120 python tools/ntt_l2.py >> gpurun_out/ntt_ab.txt 2>&1
done
for v in 2 3; do
  echo "== nocsel GLB_NTT=$v" >> gpurun_out/ntt_ab.txt
  GLB_LIB=$PWD/tools/libglb200_nocsel.so GLB_NTT=$v timeout 120 python tools/prof_lde.py 23 128 3 >> gpurun_out/ntt_ab.txt 2>&1
done
cat gpurun_out/ntt_ab.txt | grep -E "==|LDE|rc=|2\^24 x 128"
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu2.txt 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_gpu2.txt
for v in 0 2 3; do
  GLB_NTT=$v timeout 300 python bench.py --steps 2 --warmup 2 --no-e2e --no-cpu > gpurun_out/bench_c2_ntt$v.json 2> gpurun_out/bench_c2_ntt$v.err; echo "bench v=$v rc=$?"
  python -c "import json;d=json.load(open('gpurun_out/bench_c2_ntt$v.json'));print(d['ms_per_step'],d['phases']['lde']['ms'],d['phases']['lde']['frac'],d['root'])"
done
kill $SMI
