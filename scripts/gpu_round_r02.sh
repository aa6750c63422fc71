#!/bin/bash
# One GPU-box pass for round 2: GPU test log, bench lines of both workloads and both arms, ncu launch lists, then the
# `ncu --set full` captures (scripts/gpu_ncu_r02.sh; exported as CSV on the box).  usage: bash scripts/gpu_round_r02.sh <tag>
TAG=${1:-r02}
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > $O/${TAG}_smi.txt
timeout 2400 python -m pytest tests -m gpu -x -q --durations=8 > $O/pytest_gpu_${TAG}.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu_${TAG}.log
tail -3 $O/pytest_gpu_${TAG}.log
timeout 300 python bench.py --impl reference --steps 20 --warmup 5 > $O/bench_${TAG}_reference.json 2> $O/bench_${TAG}_reference.err; echo rc=$?
timeout 300 python bench.py --steps 20 --warmup 5 > $O/bench_${TAG}.json 2> $O/bench_${TAG}.err; echo rc=$?
timeout 300 python bench.py --l2 flush --cpu-seconds 1 > $O/bench_${TAG}_flush.json 2> $O/bench_${TAG}_flush.err; echo rc=$?
timeout 300 python bench.py --sparsity random --cpu-seconds 1 > $O/bench_${TAG}_sparsity_random.json 2>/dev/null; echo rc=$?
timeout 300 python bench.py --steps 50 --warmup 5 --workload config5 > $O/bench_${TAG}_config5.json 2> $O/bench_${TAG}_config5.err; echo rc=$?
timeout 300 python bench.py --impl reference --steps 5 --warmup 1 --workload config5 > $O/bench_${TAG}_config5_reference.json 2>/dev/null; echo rc=$?
timeout 400 python bench.py --steps 20 --warmup 5 --workload config4 > $O/bench_${TAG}_config4.json 2> $O/bench_${TAG}_config4.err; echo rc=$?
timeout 400 python bench.py --steps 20 --warmup 5 --workload config4 --sparsity random --cpu-seconds 1 > $O/bench_${TAG}_config4_sparsity_random.json 2>/dev/null; echo rc=$?
timeout 400 python bench.py --impl reference --steps 3 --warmup 1 --workload config4 > $O/bench_${TAG}_config4_reference.json 2>/dev/null; echo rc=$?
cat $O/bench_${TAG}.json $O/bench_${TAG}_config5.json $O/bench_${TAG}_config4.json | cut -c1-600
# launch lists (per-launch durations under ncu are cold-cache and serialised: shares, not absolutes)
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_${TAG}_config3.csv \
    python bench.py --steps 3 --warmup 3 --cpu-seconds 0.3 > $O/ncu_launch3.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_${TAG}_config5.csv \
    python bench.py --steps 3 --warmup 3 --cpu-seconds 0.3 --workload config5 > $O/ncu_launch5.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_${TAG}_config4.csv \
    python bench.py --steps 3 --warmup 3 --cpu-seconds 0.3 --workload config4 > $O/ncu_launch4.log 2>&1
bash scripts/gpu_ncu_r02.sh
ls -la $O | tail -30
