#!/bin/bash
# Round 2, GPU call 18: GPU suite and c2 line on the leaf kernel with the three IMAD.WIDE savings, ncu --set full of that kernel (mid shape) for the instruction mix.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu18.txt 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu18.txt
CMD="python bench.py --workload mid --steps 1 --warmup 1 --no-e2e --no-cpu"
$CMD > gpurun_out/plain_mid18.log 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_linear_hash_rows -c 1 -f -o gpurun_out/leaf_r2b $CMD > gpurun_out/ncu_leaf18.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/ncu_leaf18.log
if [ -f gpurun_out/leaf_r2b.ncu-rep ]; then
  ncu -i gpurun_out/leaf_r2b.ncu-rep --page raw --csv > gpurun_out/leaf_r2b_raw.csv 2>/dev/null
  ncu -i gpurun_out/leaf_r2b.ncu-rep --page source --csv > gpurun_out/leaf_r2b_source.csv 2>/dev/null
  ls -la gpurun_out/leaf_r2b*; rm -f gpurun_out/leaf_r2b.ncu-rep
fi
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench18_c2.json 2> gpurun_out/bench18_c2.err; echo "bench rc=$?"; tail -3 gpurun_out/bench18_c2.err
python -c "import json;d=json.load(open('gpurun_out/bench18_c2.json'));print(d['ms_per_step'],{k:round(v['ms'],1) for k,v in d['phases'].items()},d['e2e']['ms_per_step'],d['e2e']['shim_sequence']['ms_per_step'],d['parity_checked'],d['roofline']['frac'])"
