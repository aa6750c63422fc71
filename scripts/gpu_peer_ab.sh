#!/bin/bash
# usage: bash scripts/gpu_peer_ab.sh N tag (under gpurun --gpus N): the sharded bench with the peer-memory gather (default) and
# with the NCCL send/recv gather (development switch MOCOB200_P2P=0), alternating, on the same box.
N=$1; TAG=${2:-r02}; O=gpurun_out
mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
for rep in 1 2; do
  for mode in peer nccl; do
    if [ $mode = nccl ]; then export MOCOB200_DEVELOPER_SWITCHES=1 MOCOB200_P2P=0; else unset MOCOB200_DEVELOPER_SWITCHES MOCOB200_P2P; fi
    $TR --master-port 295$rep$N bench.py --gpus $N --steps 40 --warmup 5 --cpu-seconds 0.2 2>/dev/null | tail -1 > $O/peer_ab_${TAG}_n${N}_${mode}_$rep.json
    python - <<PY
import json
d = json.load(open('$O/peer_ab_${TAG}_n${N}_${mode}_$rep.json'))
print('n=$N $mode rep $rep: value %.1f e2e %.1f ms %.4f stage %.4f tail(rank 0) %.4f | %s' % (d['value'], d['e2e']['value'], d['ms_per_step'], d['roofline']['kernel_ms'], d['roofline']['assemble_kernel_ms'], d['run']['sharding'][-40:]))
PY
  done
done
