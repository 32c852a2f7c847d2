#!/bin/bash
mkdir -p gpurun_out
python tools/lde_sweep.py 300 32 64 96 128 160 300 > gpurun_out/lde_sweep.txt 2>&1
python tools/lde_sweep.py 600 32 64 128 256 >> gpurun_out/lde_sweep.txt 2>&1
python tools/lde_sweep.py 75 32 64 75 >> gpurun_out/lde_sweep.txt 2>&1
cat gpurun_out/lde_sweep.txt
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu6.txt 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/pytest_gpu6.txt
