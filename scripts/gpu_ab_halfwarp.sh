#!/bin/bash
# A-B on one box: two evaluations per warp (16 lanes each) in the one-sided FD kernels of small models (default for them) against
# one evaluation per warp (MOCOB200_HALFWARP=0); then the GPU tests at the default. Development switches, same binary.
export MOCOB200_DEVELOPER_SWITCHES=1
O=gpurun_out
mkdir -p $O
run() { # name, env...
    local name=$1; shift
    for rep in 1 2; do
        env "$@" timeout 200 python bench.py --steps 200 --warmup 10 --cpu-seconds 0.2 > $O/ab_half_${name}_${rep}.json 2> $O/ab_half_${name}_${rep}.err
        python - <<P
import json
try:
    b=json.load(open("$O/ab_half_${name}_${rep}.json"))
    print("$name", "$rep", "device %.0f" % b["value"], "e2e %.0f" % b["e2e"]["value"], "fd stage ms %.4f" % b["roofline"]["kernel_ms"])
except Exception as e:
    print("$name", "$rep", "failed", e, open("$O/ab_half_${name}_${rep}.err").read()[-300:])
P
    done
}
run full MOCOB200_HALFWARP=0
run half MOCOB200_HALFWARP=1
run half_wpb8 MOCOB200_HALFWARP=1 MOCOB200_WPB=8
run half_wpb6 MOCOB200_HALFWARP=1 MOCOB200_WPB=6
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu_half.log 2>&1; tail -3 $O/pytest_gpu_half.log
