#!/bin/bash
# quick GPU pass: parity tests + the bench lines
mkdir -p gpurun_out
if [ "$1" != "notest" ]; then timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -6; fi
timeout 300 python bench.py --steps 200 --warmup 10 --cpu-seconds 2 > gpurun_out/bench_config3.json 2> gpurun_out/bench_config3.err; echo rc=$?
timeout 300 python bench.py --steps 200 --warmup 10 --cpu-seconds 1 --l2 flush > gpurun_out/bench_config3_flush.json 2> gpurun_out/bench_config3_flush.err; echo rc=$?
timeout 300 python bench.py --steps 30 --warmup 5 --workload config5 --cpu-seconds 2 > gpurun_out/bench_config5.json 2> gpurun_out/bench_config5.err; echo rc=$?
python - <<'PY'
import json
for f in ("bench_config3","bench_config3_flush","bench_config5"):
    try:
        d=json.loads(open("gpurun_out/%s.json"%f).read().strip().split("\n")[-1])
        print(f, "value %.1f e2e %.1f ms %.4f kernel_ms %.4f asm %.4f frac %.4f cpu %.1f launches %d" % (d["value"], d["e2e"]["value"], d["ms_per_step"], d["roofline"]["kernel_ms"], d["roofline"]["assemble_kernel_ms"], d["roofline"]["frac"], d.get("cpu_baseline",{}).get("value",0), d["gpu_launches"]))
    except Exception as e: print(f, "ERR", e, open("gpurun_out/%s.json"%f).read()[-500:], open("gpurun_out/%s.err"%f).read()[-1500:])
PY
