#!/bin/bash
# A-B of the development switches of the thread-per-evaluation kernels on the config-5 bench workload (run under gpurun).
# prints: switches, device Jac/s, e2e Jac/s, FD-stage ms, assembly ms
export MOCOB200_DEVELOPER_SWITCHES=1  # the library ignores its A-B switches without it
run() {
    env "$@" python bench.py --steps 40 --warmup 5 --workload config5 --cpu-seconds 1 2>/dev/null | tail -1 | python -c "
import sys, json
d = json.loads(sys.stdin.read())
print('$*', round(d['value'], 1), round(d['e2e']['value'], 1), round(d['roofline']['kernel_ms'], 4), round(d['roofline']['assemble_kernel_ms'], 4))"
}
run A=0
run MOCOB200_FLAT_MINBLOCKS=4
run MOCOB200_VEL_MINBLOCKS=4
run MOCOB200_FLAT_MINBLOCKS=4 MOCOB200_VEL_MINBLOCKS=4
run MOCOB200_VEL_KERNEL=0
run MOCOB200_GRAPH=0
