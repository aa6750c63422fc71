#!/bin/bash
# A-B of the warp-kernel development switches on the config-3 bench workload (run under gpurun).
# prints: switches, device Jac/s, e2e Jac/s, FD-stage ms
export MOCOB200_DEVELOPER_SWITCHES=1  # the library ignores its A-B switches without it
run() {
    env "$@" python bench.py --steps 200 --warmup 10 --cpu-seconds 0.5 2>/dev/null | tail -1 | python -c "
import sys, json
d = json.loads(sys.stdin.read())
print('$*', round(d['value'], 1), round(d['e2e']['value'], 1), round(d['roofline']['kernel_ms'], 4))"
}
run A=0
run MOCOB200_LOCKSTEP=1
run MOCOB200_WPB=8
run MOCOB200_WPB=8 MOCOB200_LOCKSTEP=1
run MOCOB200_WPB=12 MOCOB200_LOCKSTEP=1
run MOCOB200_WPB=2 MOCOB200_LOCKSTEP=1
run MOCOB200_GRAPH=0
