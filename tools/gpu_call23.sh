#!/bin/bash
# Round 2, GPU call 23: ncu --set full of the shipped NTT passes (prof_lde 2^22 x 32) and the launch list of a c2 step at HEAD.
mkdir -p gpurun_out
CMD1="python tools/prof_lde.py 22 32 2"
$CMD1 > gpurun_out/plain_lde23.log 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_pass -c 12 -f -o gpurun_out/ntt_r2b $CMD1 > gpurun_out/ncu_ntt23.log 2>&1
echo "ncu ntt rc=$?"; tail -2 gpurun_out/ncu_ntt23.log
if [ -f gpurun_out/ntt_r2b.ncu-rep ]; then ncu -i gpurun_out/ntt_r2b.ncu-rep --page raw --csv > gpurun_out/ntt_r2b_raw.csv 2>/dev/null; ls -la gpurun_out/ntt_r2b*; rm -f gpurun_out/ntt_r2b.ncu-rep; fi
CMD2="python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu"
$CMD2 > gpurun_out/plain_c2_23.log 2>&1 && timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches23_c2.csv $CMD2 > gpurun_out/ncu_launch23.log 2>&1
echo "ncu launches rc=$?"; wc -l gpurun_out/launches23_c2.csv
