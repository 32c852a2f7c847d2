#!/bin/bash
# Round 2, GPU call 20: GPU suite, smoke and c2 line on the final kernels (NTT passes without validity tests where nothing is padded, canon_sel with one carry
# chain, gl::sqr3 as one PTX block), ncu --set full of the shipped leaf kernel (mid shape) for the instruction mix.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu20.txt 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu20.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke20.txt 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke20.txt
CMD="python bench.py --workload mid --steps 1 --warmup 1 --no-e2e --no-cpu"
$CMD > gpurun_out/plain_mid20.log 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_linear_hash_rows -c 1 -f -o gpurun_out/leaf_r2c $CMD > gpurun_out/ncu_leaf20.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/ncu_leaf20.log
if [ -f gpurun_out/leaf_r2c.ncu-rep ]; then
  ncu -i gpurun_out/leaf_r2c.ncu-rep --page raw --csv > gpurun_out/leaf_r2c_raw.csv 2>/dev/null
  ncu -i gpurun_out/leaf_r2c.ncu-rep --page source --csv > gpurun_out/leaf_r2c_source.csv 2>/dev/null
  python tools/ncu_leaf_summary.py gpurun_out/leaf_r2c_raw.csv gpurun_out/leaf_r2c_source.csv "profiles/r2_42_leaf_ncu_shipped.txt (ncu --set full --clock-control none --import-source on of poseidon::k_linear_hash_rows<false> as shipped after the sqr3 / fold / canon changes; bench.py --workload mid, 2^21 rows x 32 permutations; tools/gpu_call20.sh)" 21 32 profiles/r2_leaf_instr_mix.json > gpurun_out/r2_42_leaf_ncu_shipped.txt
  cp profiles/r2_leaf_instr_mix.json gpurun_out/r2_leaf_instr_mix.json
  rm -f gpurun_out/leaf_r2c.ncu-rep
fi
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench20_c2.json 2> gpurun_out/bench20_c2.err; echo "bench rc=$?"; tail -3 gpurun_out/bench20_c2.err
python -c "import json;d=json.load(open('gpurun_out/bench20_c2.json'));print(d['ms_per_step'],{k:round(v['ms'],1) for k,v in d['phases'].items()},d['e2e']['ms_per_step'],d['e2e']['shim_sequence']['ms_per_step'],d['parity_checked'],d['roofline']['frac'],d['roofline']['pipe_model']['frac_of_pipe_bound'],d['roofline']['instr_per_permutation'])"
