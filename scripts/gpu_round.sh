#!/bin/bash
# One GPU-box pass: parity tests, bench lines, ncu launch list + full captures of the dominant kernels.
# usage: bash scripts/gpu_round.sh <tag>   (outputs under gpurun_out/<tag>_*)
TAG=${1:-r01}
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/${TAG}_smi.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${TAG}_pytest_gpu.log
tail -3 gpurun_out/${TAG}_pytest_gpu.log
timeout 300 python bench.py --impl reference --steps 100 --warmup 5 > gpurun_out/${TAG}_bench_reference.json 2> gpurun_out/${TAG}_bench_reference.err; echo rc=$?
timeout 300 python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo rc=$?
timeout 300 python bench.py --l2 flush --cpu-seconds 1 > gpurun_out/${TAG}_bench_flush.json 2> gpurun_out/${TAG}_bench_flush.err; echo rc=$?
timeout 300 python bench.py --steps 50 --warmup 5 --workload config5 > gpurun_out/${TAG}_bench_config5.json 2> gpurun_out/${TAG}_bench_config5.err; echo rc=$?
timeout 300 python bench.py --impl reference --steps 5 --warmup 1 --workload config5 > gpurun_out/${TAG}_bench_config5_reference.json 2>&1; echo rc=$?
cat gpurun_out/${TAG}_bench.json gpurun_out/${TAG}_bench_config5.json
# launch list (per-launch durations are cold-cache and serialised under ncu: shares, not absolutes)
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv python bench.py --steps 3 --warmup 3 --cpu-seconds 0.3 > gpurun_out/${TAG}_ncu_launch.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_wpoints_fd1 -s 16 -c 4 -o gpurun_out/${TAG}_wfd1_full -f python bench.py --steps 3 --warmup 3 --cpu-seconds 0.3 > gpurun_out/${TAG}_ncu_full.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_points_fd_flat|k_wlocal_central|k_wpoints_nominal" -s 9 -c 3 -o gpurun_out/${TAG}_tfd_full -f python bench.py --steps 2 --warmup 3 --cpu-seconds 0.3 --workload config5 > gpurun_out/${TAG}_ncu_full5.log 2>&1
ls -la gpurun_out | tail -30
