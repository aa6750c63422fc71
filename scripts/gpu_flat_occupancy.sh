#!/bin/bash
# config 5: flat thread kernel variants — central sides paired on adjacent lanes or not (MOCOB200_FLAT_PAIR), compiled for
# 4 / 8 resident 64-thread blocks per SM (MOCOB200_FLAT_MINBLOCKS)
export MOCOB200_DEVELOPER_SWITCHES=1  # the library ignores its A-B switches without it
mkdir -p gpurun_out
for pair in 0 1; do for mb in 4 8; do
  tag=p${pair}_mb$mb
  MOCOB200_FLAT_PAIR=$pair MOCOB200_FLAT_MINBLOCKS=$mb timeout 300 python bench.py --steps 30 --warmup 5 --workload config5 --cpu-seconds 0.2 > gpurun_out/flat_$tag.json 2> gpurun_out/flat_$tag.err
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/flat_$tag.json").read().strip().split("\n")[-1])
    print("$tag value %.1f e2e %.1f kernel_ms %.4f" % (d["value"], d["e2e"]["value"], d["roofline"]["kernel_ms"]))
except Exception as e: print("$tag ERR", e, open("gpurun_out/flat_$tag.err").read()[-800:])
PY
done; done
