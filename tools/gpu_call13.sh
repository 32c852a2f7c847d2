#!/bin/bash
# Round 2, GPU call 13: column-block schedule of the host-pointer commit, A/B on one box with the per-block timeline.
mkdir -p gpurun_out; rm -f gpurun_out/e2e_ab.txt
for s in 0 1 32 0 1; do
  GLB_HOST_BLOCKS=$s GLB_HOST_TRACE=1 timeout 300 python tools/e2e_ab.py 2 >> gpurun_out/e2e_ab.txt 2>&1
done
grep -E "raw|gl_commit_host" gpurun_out/e2e_ab.txt
