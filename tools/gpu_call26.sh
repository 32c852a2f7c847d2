#!/bin/bash
# Round 2, GPU call 26: first pass of the extended forward transform without its trivial butterflies: GPU suite, c2 and c4 lines (device only).
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu26.txt 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu26.txt
for w in c2 c4; do
timeout 600 python bench.py --workload $w --steps 5 --warmup 3 --no-e2e --no-cpu > gpurun_out/bench26_$w.json 2> gpurun_out/bench26_$w.err; echo "bench $w rc=$?"; tail -2 gpurun_out/bench26_$w.err
python -c "import json;d=json.load(open('gpurun_out/bench26_$w.json'));print(d['ms_per_step'],{k:round(v['ms'],1) for k,v in d['phases'].items()},d['root'][:16])"
done
