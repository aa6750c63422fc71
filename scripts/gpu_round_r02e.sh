#!/bin/bash
# Last box pass of round 2 (tag r02e): GPU tests at HEAD, the solve stand-in with both callback arms (scripts/solve_standin.py),
# a WARM ncu capture of the config-3 point-function kernels (--cache-control none: the default capture flushes the caches before
# every replay pass, which charges the 237 KB straight-line routine a cold instruction fetch each time), and the headline bench line.
TAG=${1:-r02e}
O=gpurun_out
mkdir -p $O
export MOCOB200_DEVELOPER_SWITCHES=1
timeout 900 python -m pytest tests -m gpu -x -q --durations=5 > $O/pytest_gpu_${TAG}.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu_${TAG}.log
tail -3 $O/pytest_gpu_${TAG}.log
timeout 300 python bench.py --steps 20 --warmup 5 > $O/bench_${TAG}.json 2> $O/bench_${TAG}.err; echo bench rc=$?
cut -c1-400 $O/bench_${TAG}.json
timeout 420 python scripts/solve_standin.py --intervals 10 --maxiter ${MAXITER:-150} --out $O/solve_standin_${TAG}.json > /dev/null 2> $O/solve_standin_${TAG}.err; echo solve rc=$?
cat $O/solve_standin_${TAG}.err | cut -c1-400
MOCOB200_GRAPH=0 timeout 600 ncu --set full --clock-control none --cache-control none -k regex:"k_wpoints_fd1" -c 8 -o /tmp/c3w \
    python bench.py --steps 1 --warmup 3 --cpu-seconds 0.3 > $O/ncu_c3w.log 2>&1; echo ncu rc=$?
ncu -i /tmp/c3w.ncu-rep --page raw --csv > $O/ncu_${TAG}_config3_warm_raw.csv 2>/dev/null
python scripts/ncu_summary.py $O/ncu_${TAG}_config3_warm $O/ncu_${TAG}_config3_warm_raw.csv > /dev/null 2>&1
cat $O/ncu_${TAG}_config3_warm.md 2>/dev/null | cut -c1-300
ls -la $O | tail -12
