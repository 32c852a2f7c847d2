#!/bin/bash
# Round 2, GPU call 11: what the per-call device allocations of the host-pointer commit cost.
mkdir -p gpurun_out
timeout 300 python tools/alloc_cost.py > gpurun_out/alloc_cost.txt 2>&1; echo "rc=$?"; cat gpurun_out/alloc_cost.txt
