#!/bin/bash
# Round 2, GPU call 19: canon_sel without its second carry chain in the NTT butterflies -- parity of the field ops / NTT tests, LDE A/B against the previous form.
mkdir -p gpurun_out; rm -f gpurun_out/lde_canon_ab.txt
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "field_ops or ntt or extend_pol" > gpurun_out/pytest_gpu19.txt 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu19.txt
for rep in 1 2; do
for lib in tools/ubench_libglb_sel0.so merlin-zkevm-prover_b200/libglb200.so; do
  echo "== $lib" >> gpurun_out/lde_canon_ab.txt
  GLB_LIB=$PWD/$lib timeout 120 python tools/prof_lde.py 23 128 4 >> gpurun_out/lde_canon_ab.txt 2>&1
done; done
grep -E "==|LDE" gpurun_out/lde_canon_ab.txt
