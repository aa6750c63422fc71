#!/bin/bash
# A-B of the pinned-buffer delivery of mocob200_eval_jac_g (config 3, bench.py --steps 200): plain copy against
# mocob200_set_host_delivery(1) (bench.py --host-delivery masked); MOCOB200_HOST_PUSH=2 = masked WITHOUT the host zero-fill
# (wrong results; shows what the fill costs); MOCOB200_PUSH_SHORT_RUN = length below which a zero run is stored by the kernel
# instead of filled by the host.
export MOCOB200_DEVELOPER_SWITCHES=1
mkdir -p gpurun_out
one() { # tag, bench flags, env...
    local tag=$1 flags=$2; shift; shift
    for i in 1 2; do env "$@" python bench.py --steps 200 --warmup 10 --cpu-seconds 0.2 --solve-iters 0 $flags > gpurun_out/bench_push_${tag}_$i.json 2>/dev/null; python -c "
import json; b=json.load(open('gpurun_out/bench_push_${tag}_$i.json')); print('$tag', round(b['value']), round(b['e2e']['value']), round(b['e2e']['ms_per_step']*1e3,1), b['e2e']['d2h_bytes_per_step'])"; done
}
one plain "" MOCOB200_NONE=0
one masked "--host-delivery masked" MOCOB200_NONE=0
one masked_nofill "--host-delivery masked" MOCOB200_HOST_PUSH=2
