#!/bin/bash
# Round 2, GPU call 7: full suite on the final library, ncu of the NTT pass (balanced butterflies), launch list of a c2 step, c5 line.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu7.txt 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/pytest_gpu7.txt
CMD1="python tools/prof_lde.py 22 32 2"
$CMD1 > gpurun_out/plain_lde.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_pass -c 12 -f -o gpurun_out/ntt_r2 $CMD1 > gpurun_out/ncu_ntt.log 2>&1
echo "ncu ntt rc=$?"; tail -2 gpurun_out/ncu_ntt.log
if [ -f gpurun_out/ntt_r2.ncu-rep ]; then ncu -i gpurun_out/ntt_r2.ncu-rep --page raw --csv > gpurun_out/ntt_r2_raw.csv 2>/dev/null; ls -la gpurun_out/ntt_r2*; rm -f gpurun_out/ntt_r2.ncu-rep; fi
CMD2="python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu"
$CMD2 > gpurun_out/plain_c2.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_c2.csv $CMD2 > gpurun_out/ncu_launch.log 2>&1
echo "ncu launches rc=$?"; wc -l gpurun_out/launches_c2.csv
timeout 600 python bench.py --workload c5 --steps 2 --warmup 1 > gpurun_out/bench_c5.json 2> gpurun_out/bench_c5.err; echo "c5 rc=$?"; python -c "import json;d=json.load(open('gpurun_out/bench_c5.json'));print(d['ms_per_step'],d['device_ms_per_step'],d['roofline']['frac'])"
timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/bench7_c2.json 2> gpurun_out/bench7_c2.err; echo "bench rc=$?"
python -c "import json;d=json.load(open('gpurun_out/bench7_c2.json'));print(d['ms_per_step'],{k:v['ms'] for k,v in d['phases'].items()},d['parity_checked']);print(d['e2e']['ms_per_step'],d['e2e']['shim_sequence']);r=d['roofline'];print(r['frac'],r['pipe_model']['frac_of_pipe_bound'],r['imad']['frac'])"
