#!/bin/bash
# Round 2, GPU call 10: whole GPU suite after the big-domain dispatch fix (ragged tiles / one-pass transforms after a 2^28 transform), no-alloc test alone and last.
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/pytest_noalloc_alone.txt 2>&1; echo "noalloc alone rc=$?"; tail -3 gpurun_out/pytest_noalloc_alone.txt
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu10.txt 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/pytest_gpu10.txt
