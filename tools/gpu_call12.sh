#!/bin/bash
# Round 2, GPU call 12: host-pointer commit with graded column blocks (16, 24, 40, 64, 112, 128 ...) and cached device buffers: GPU suite, c2 line (e2e).
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu12.txt 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/pytest_gpu12.txt
timeout 900 python bench.py > gpurun_out/bench12_c2.json 2> gpurun_out/bench12_c2.err; echo "bench rc=$?"; tail -3 gpurun_out/bench12_c2.err
python -c "import json;d=json.load(open('gpurun_out/bench12_c2.json'));print(d['ms_per_step'],d['e2e']['ms_per_step'],d['e2e']['shim_sequence'],d['parity_checked'])"
