#!/bin/bash
# Round 2, GPU call 25: ragged column tiles of 17..24 columns as 16 + <= 8: GPU suite, c2 line (device only).
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu25.txt 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu25.txt
timeout 600 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu > gpurun_out/bench25_c2.json 2> gpurun_out/bench25_c2.err; echo "bench rc=$?"; tail -2 gpurun_out/bench25_c2.err
python -c "import json;d=json.load(open('gpurun_out/bench25_c2.json'));print(d['ms_per_step'],{k:round(v['ms'],1) for k,v in d['phases'].items()},d['root'][:16])"
