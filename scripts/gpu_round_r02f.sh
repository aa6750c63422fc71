#!/bin/bash
# Final box pass of round 2 (tag r02f) at HEAD: GPU tests, smoke(), and the bench lines of the three workloads with their CPU arms.
TAG=${1:-r02f}
O=gpurun_out
mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q --durations=5 > $O/pytest_gpu_${TAG}.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu_${TAG}.log
tail -3 $O/pytest_gpu_${TAG}.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke_${TAG}.log 2>&1; echo smoke rc=$?; tail -1 $O/smoke_${TAG}.log
timeout 300 python bench.py --impl reference --steps 20 --warmup 5 > $O/bench_${TAG}_reference.json 2>/dev/null; echo rc=$?
timeout 300 python bench.py --steps 20 --warmup 5 > $O/bench_${TAG}.json 2> $O/bench_${TAG}.err; echo rc=$?
timeout 300 python bench.py --steps 50 --warmup 5 --workload config5 > $O/bench_${TAG}_config5.json 2>/dev/null; echo rc=$?
timeout 400 python bench.py --steps 20 --warmup 5 --workload config4 > $O/bench_${TAG}_config4.json 2>/dev/null; echo rc=$?
for f in $O/bench_${TAG}.json $O/bench_${TAG}_config5.json $O/bench_${TAG}_config4.json $O/bench_${TAG}_reference.json; do
python - <<P
import json
b=json.load(open("$f")); print("$f", round(b["value"],1), "e2e", round(b["e2e"]["value"],1), b.get("roofline",{}).get("frac"), b.get("cpu_baseline",{}).get("value"))
P
done
