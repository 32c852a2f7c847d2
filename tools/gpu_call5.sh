#!/bin/bash
# Round 2, GPU call 5: full GPU suite on the new defaults, c2 bench, ncu --set full of the shipped leaf kernel (mid shape).
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu5.txt 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/pytest_gpu5.txt
timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/bench5_c2.json 2> gpurun_out/bench5_c2.err; echo "bench rc=$?"; tail -3 gpurun_out/bench5_c2.err
python -c "import json;d=json.load(open('gpurun_out/bench5_c2.json'));print(d['ms_per_step'],d['phases']);print(d['e2e'])"
CMD="python bench.py --workload mid --steps 1 --warmup 1 --no-e2e --no-cpu"
$CMD > gpurun_out/plain_mid.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_linear_hash_rows -c 1 -f -o gpurun_out/leaf_r2 $CMD > gpurun_out/ncu_leaf.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu_leaf.log
if [ -f gpurun_out/leaf_r2.ncu-rep ]; then
  ncu -i gpurun_out/leaf_r2.ncu-rep --page raw --csv > gpurun_out/leaf_r2_raw.csv 2>/dev/null
  ncu -i gpurun_out/leaf_r2.ncu-rep --page source --csv > gpurun_out/leaf_r2_source.csv 2>/dev/null
  ls -la gpurun_out/leaf_r2*; rm -f gpurun_out/leaf_r2.ncu-rep
fi
