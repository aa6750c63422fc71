#!/bin/bash
# usage: bash scripts/gpu_scaling.sh N tag  -- multi-rank parity check + bench lines at N GPUs
N=$1; TAG=${2:-r01}
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 scripts/gpu_multirank_check.py 2>&1 | grep -E "sharded|MULTIRANK" | tee gpurun_out/${TAG}_multirank_n$N.txt
for w in config3 config5; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 50 --warmup 5 --workload $w 2>/dev/null | tail -1 > gpurun_out/${TAG}_scale_${w}_n$N.json
  python -c "
import json;d=json.load(open('gpurun_out/${TAG}_scale_${w}_n$N.json'));print('$w n=$N value %.1f e2e %.1f ms %.4f kernel_ms %.4f'%(d['value'],d['e2e']['value'],d['ms_per_step'],d['roofline']['kernel_ms']))"
done
