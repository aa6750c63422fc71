#!/bin/bash
# Box pass at the round's final commit (tag r02i): GPU tests, smoke(), the default bench line, the CPU arm, and the
# end-to-end call with mocob200_set_host_delivery(1) beside the host CPU's name (what that option gains depends on the host).
TAG=${1:-r02i}
O=gpurun_out
mkdir -p $O
(lscpu | grep -E "Model name|^CPU\(s\)|L2 cache|L3 cache"; nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv,noheader) > $O/host_${TAG}.txt 2>&1
cat $O/host_${TAG}.txt
timeout 900 python -m pytest tests -m gpu -x -q --durations=5 > $O/pytest_gpu_${TAG}.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu_${TAG}.log
tail -3 $O/pytest_gpu_${TAG}.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke_${TAG}.log 2>&1; echo smoke rc=$?; tail -1 $O/smoke_${TAG}.log
timeout 300 python bench.py --impl reference --steps 20 --warmup 5 > $O/bench_${TAG}_reference.json 2>/dev/null; echo rc=$?
timeout 300 python bench.py > $O/bench_${TAG}.json 2> $O/bench_${TAG}.err; echo rc=$?
timeout 300 python bench.py --host-delivery masked --solve-iters 0 --cpu-seconds 0.5 > $O/bench_${TAG}_masked.json 2>/dev/null; echo rc=$?
for f in $O/bench_${TAG}.json $O/bench_${TAG}_masked.json $O/bench_${TAG}_reference.json; do
python - <<P
import json
b=json.load(open("$f")); print("$f", round(b["value"],1), "e2e", round(b["e2e"]["value"],1), b["e2e"].get("ms_per_step"), b["e2e"].get("d2h_bytes_per_step"))
P
done
