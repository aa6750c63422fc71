#!/bin/bash
# One GPU-box pass: parity tests, bench lines, ncu launch list + one full capture of the dominant kernel.
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/smi.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 300 python bench.py --steps 200 --warmup 10 > gpurun_out/bench_config3.json 2> gpurun_out/bench_config3.err; echo rc=$?
timeout 300 python bench.py --steps 50 --warmup 5 --workload config5 > gpurun_out/bench_config5.json 2> gpurun_out/bench_config5.err; echo rc=$?
timeout 300 python bench.py --impl reference --steps 100 --warmup 5 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo rc=$?
MOCOB200_KERNEL=thread timeout 300 python bench.py --steps 50 --warmup 5 --cpu-seconds 1 > gpurun_out/bench_config3_thread.json 2>&1
cat gpurun_out/bench_config3.json gpurun_out/bench_config5.json gpurun_out/bench_reference.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --cpu-seconds 0.5 > gpurun_out/ncu_launch.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_wpoints_fd1 -s 4 -c 2 -o gpurun_out/wfd1_full -f python bench.py --steps 2 --warmup 3 --cpu-seconds 0.5 > gpurun_out/ncu_full.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_wpoints_central -s 2 -c 1 -o gpurun_out/wcentral_full -f python bench.py --steps 2 --warmup 3 --cpu-seconds 0.5 --workload config5 > gpurun_out/ncu_full5.log 2>&1
ls -la gpurun_out
