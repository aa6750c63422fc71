#!/bin/bash
# usage: bash scripts/gpu_scaling_r02.sh N tag  (under gpurun --gpus N) -- what the driver's scaling step runs at N GPUs:
# multi-rank bit-identity check, then `bench.py --gpus N --steps 20 --warmup 5` under torchrun (default workload of a
# distributed run = config 5, the configuration that shards), then the reference arm launched the same way.
N=$1; TAG=${2:-r02}; O=gpurun_out
mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
$TR --master-port 29511 scripts/gpu_multirank_check.py 2>&1 | grep -E "sharded|MULTIRANK" | tee $O/multirank_${TAG}_n$N.txt
$TR --master-port 29512 bench.py --gpus $N --steps 20 --warmup 5 2>$O/scale_${TAG}_n$N.err | tail -1 > $O/scale_${TAG}_n$N.json
$TR --master-port 29513 bench.py --gpus $N --steps 20 --warmup 5 --impl reference 2>/dev/null | tail -1 > $O/scale_${TAG}_reference_n$N.json
python - <<PY
import json
d = json.load(open('$O/scale_${TAG}_n$N.json')); r = json.load(open('$O/scale_${TAG}_reference_n$N.json'))
print('n=$N', d['config'].get('workload'), 'value %.1f e2e %.1f ms %.4f stage_ms %.4f working ranks %s | reference %.2f (%s cores)' % (
    d['value'], d['e2e']['value'], d['ms_per_step'], d['roofline']['kernel_ms'], d.get('run', {}).get('working_ranks'), r['value'], r['cpu_baseline']['cores']))
PY
