#!/bin/bash
# Round 2, GPU call 9: allocation-free extendPol (static device scratch) -- alone in a fresh process, under memcheck, then the whole suite; c2 bench line at HEAD.
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/pytest_noalloc_alone.txt 2>&1; echo "noalloc alone rc=$?"; tail -4 gpurun_out/pytest_noalloc_alone.txt
timeout 600 compute-sanitizer --tool memcheck --error-exitcode 9 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -k twice > gpurun_out/memcheck_noalloc.txt 2>&1; echo "memcheck rc=$?"; tail -5 gpurun_out/memcheck_noalloc.txt
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu9.txt 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/pytest_gpu9.txt
timeout 900 python bench.py > gpurun_out/bench9_c2.json 2> gpurun_out/bench9_c2.err; echo "bench rc=$?"; tail -3 gpurun_out/bench9_c2.err; cat gpurun_out/bench9_c2.json | head -c 3000
