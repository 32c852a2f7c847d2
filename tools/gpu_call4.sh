#!/bin/bash
# Round 2, GPU call 4: corrected pipe-balanced butterflies -- full GPU suite, A/B timing, c2 bench with e2e + shim sequence.
mkdir -p gpurun_out; rm -f gpurun_out/ntt_ab4.txt
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu4.txt 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/pytest_gpu4.txt
for lib in "" "$PWD/tools/libglb200_nocsel.so"; do
 for v in 0 2; do
  echo "== lib=${lib:-csel(default)} GLB_NTT=$v" >> gpurun_out/ntt_ab4.txt
  GLB_LIB=$lib GLB_NTT=$v timeout 120 python tools/prof_lde.py 23 128 4 >> gpurun_out/ntt_ab4.txt 2>&1; echo "rc=$?" >> gpurun_out/ntt_ab4.txt
 done
done
grep -E "==|LDE" gpurun_out/ntt_ab4.txt
GLB_NTT=0 timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/bench4_c2.json 2> gpurun_out/bench4_c2.err; echo "bench rc=$?"; tail -3 gpurun_out/bench4_c2.err
python -c "import json;d=json.load(open('gpurun_out/bench4_c2.json'));print(d['ms_per_step'],d['phases']['lde']['ms'],d['phases']['lde']['frac'],d['root'],d['parity_checked']);print(d['e2e']);print(d['roofline']['frac'],d['roofline']['peak_source'])"
