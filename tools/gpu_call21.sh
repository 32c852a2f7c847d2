#!/bin/bash
# Round 2, GPU call 21: c4 (recursive shape, blowup 16) and c5 (full commit sequence) lines on the final kernels.
mkdir -p gpurun_out
timeout 600 python bench.py --workload c4 --steps 5 --warmup 3 > gpurun_out/bench21_c4.json 2> gpurun_out/bench21_c4.err; echo "c4 rc=$?"; tail -2 gpurun_out/bench21_c4.err
python -c "import json;d=json.load(open('gpurun_out/bench21_c4.json'));print(d['ms_per_step'],{k:round(v['ms'],1) for k,v in d['phases'].items()},d['e2e']['ms_per_step'],d.get('parity_checked'))"
timeout 900 python bench.py --workload c5 --steps 3 --warmup 2 > gpurun_out/bench21_c5.json 2> gpurun_out/bench21_c5.err; echo "c5 rc=$?"; tail -2 gpurun_out/bench21_c5.err
python -c "import json;d=json.load(open('gpurun_out/bench21_c5.json'));print(d['ms_per_step'],d['e2e'])"
