timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
# A-B of the local-column work split (development switch): default = one warp per 32 owners of local inputs;
# "1" = one warp per grid point (MOCOB200_LOCAL_PARTS=1)
for wl in config3 config4; do
  for parts in default 1; do
    unset MOCOB200_DEVELOPER_SWITCHES MOCOB200_LOCAL_PARTS
    if [ $parts = 1 ]; then export MOCOB200_DEVELOPER_SWITCHES=1 MOCOB200_LOCAL_PARTS=1; fi
    for rep in 1 2; do
    python bench.py --steps 40 --warmup 5 --workload $wl --cpu-seconds 0.2 2>/dev/null | tail -1 | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('$wl parts=$parts: value %.1f e2e %.1f ms %.4f stage %.4f'%(d['value'],d['e2e']['value'],d['ms_per_step'],d['roofline']['kernel_ms']), d['roofline']['evals_executed_per_step'])"
    done
  done
done
