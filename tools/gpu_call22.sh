#!/bin/bash
# Round 2, GPU call 22: GPU suite and the c2 line (default flags) at HEAD, with the secondary end-to-end figure (extended matrix kept in HBM).
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu22.txt 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu22.txt
( time timeout 900 python bench.py > gpurun_out/bench22_c2.json 2> gpurun_out/bench22_c2.err ) 2>&1 | grep real; echo "bench rc=$?"; tail -3 gpurun_out/bench22_c2.err
python -c "import json;d=json.load(open('gpurun_out/bench22_c2.json'));print(d['ms_per_step'],{k:round(v['ms'],1) for k,v in d['phases'].items()},d['e2e']['ms_per_step'],d['e2e']['shim_sequence']['ms_per_step'],d['e2e']['matrix_kept_on_device'],d['parity_checked'],d['roofline']['frac'])"
