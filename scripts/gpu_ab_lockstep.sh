#!/bin/bash
# A-B on one box: do the warps of a block of the config-3 warp kernels share instruction fetches when they meet at the phase
# boundaries (MOCOB200_LOCKSTEP=1), with 4 and with 8 warps per block (MOCOB200_WPB)? Development switches, same binary.
export MOCOB200_DEVELOPER_SWITCHES=1
O=gpurun_out
mkdir -p $O
run() { # name, env...
    local name=$1; shift
    for rep in 1 2; do
        env "$@" timeout 200 python bench.py --steps 200 --warmup 10 --cpu-seconds 0.2 > $O/ab_lockstep_${name}_${rep}.json 2>/dev/null
        python - <<P
import json
try:
    b=json.load(open("$O/ab_lockstep_${name}_${rep}.json"))
    print("$name", "$rep", "device %.0f" % b["value"], "e2e %.0f" % b["e2e"]["value"], "fd stage ms %.4f" % b["roofline"]["kernel_ms"])
except Exception as e:
    print("$name", "$rep", "failed", e)
P
    done
}
run default MOCOB200_NONE=0
run lockstep MOCOB200_LOCKSTEP=1
run wpb8 MOCOB200_WPB=8
run wpb8_lockstep MOCOB200_WPB=8 MOCOB200_LOCKSTEP=1
run wpb2_lockstep MOCOB200_WPB=2 MOCOB200_LOCKSTEP=1
# parity of the lockstep variant (the barriers change no arithmetic): the config-3 parity tests under the switch
MOCOB200_LOCKSTEP=1 timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "gait2d_tracking or config3" 2>&1 | tail -2
