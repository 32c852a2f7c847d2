#!/bin/bash
# Round 2, GPU call 8: big-domain path as its own instantiation (the run-time branch cost 6 % of the LDE), L2 prefetch A/B, host Q / FRI entry points, c5 e2e.
mkdir -p gpurun_out; rm -f gpurun_out/ntt_ab8.txt
for pf in 0 1 2 3; do
  echo "== GLB_NTT_PREFETCH=$pf" >> gpurun_out/ntt_ab8.txt
  GLB_NTT_PREFETCH=$pf timeout 120 python tools/prof_lde.py 23 128 4 >> gpurun_out/ntt_ab8.txt 2>&1
  GLB_NTT_PREFETCH=$pf timeout 120 python tools/prof_lde.py 20 96 4 >> gpurun_out/ntt_ab8.txt 2>&1
done
grep -E "==|LDE" gpurun_out/ntt_ab8.txt
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu8.txt 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/pytest_gpu8.txt
for pf in 0 1; do
GLB_NTT_PREFETCH=$pf timeout 600 python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu > gpurun_out/bench8_c2_pf$pf.json 2> gpurun_out/bench8_c2_pf$pf.err; echo "bench pf=$pf rc=$?"
python -c "import json;d=json.load(open('gpurun_out/bench8_c2_pf$pf.json'));print(d['ms_per_step'],{k:v['ms'] for k,v in d['phases'].items()},d['root'])"
done
timeout 900 python bench.py --workload c5 --steps 2 --warmup 1 > gpurun_out/bench8_c5.json 2> gpurun_out/bench8_c5.err; echo "c5 rc=$?"; tail -2 gpurun_out/bench8_c5.err; python -c "import json;d=json.load(open('gpurun_out/bench8_c5.json'));print(d['ms_per_step'],d['e2e'])"
