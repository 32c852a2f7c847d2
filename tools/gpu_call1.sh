#!/bin/bash
# Round 2, GPU call 1: full GPU test-suite (with the new c2/c4-size parity tests), the issue-cost and TMA tile micro-benchmarks
# that size the NTT rebuild, the prepared leaf-kernel A/B, and a baseline bench line on this round's box.
mkdir -p gpurun_out
{ nproc; free -g | head -2; lscpu | grep -E "Model name|Socket|NUMA node\(s\)|Thread"; nvidia-smi --query-gpu=name,clocks.max.sm,clocks.max.mem,power.limit --format=csv; } > gpurun_out/box.txt 2>&1
nvidia-smi --query-gpu=index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap --format=csv -lms 500 > gpurun_out/clocks.csv &
SMI=$!
timeout 120 tools/ubench_pipes > gpurun_out/ubench_pipes.txt 2>&1; echo "ubench_pipes rc=$?"
timeout 120 tools/ubench_tile > gpurun_out/ubench_tile.txt 2>&1; echo "ubench_tile rc=$?"
timeout 120 python tools/ntt_l2.py > gpurun_out/ntt_l2.txt 2>&1; echo "ntt_l2 rc=$?"
timeout 300 bash tools/ab_round2.sh run > gpurun_out/ab_round2.log 2>&1; echo "ab rc=$?"
timeout 1500 python -m pytest tests -m gpu -q --durations=8 > gpurun_out/pytest_gpu.txt 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_gpu.txt
timeout 600 python bench.py --steps 3 --warmup 3 > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err; echo "bench rc=$?"; tail -c 1500 gpurun_out/bench_c2.json; tail -5 gpurun_out/bench_c2.err
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"; cat gpurun_out/bench_ref.json
kill $SMI
head -40 gpurun_out/ubench_pipes.txt
