#!/bin/bash
# WARM ncu capture (--cache-control none) of the config-5 FD kernels: DRAM traffic and stalls with L2 left as the step leaves it.
export MOCOB200_DEVELOPER_SWITCHES=1
O=gpurun_out; mkdir -p $O
python bench.py --steps 3 --warmup 3 --workload config5 --cpu-seconds 0.2 > /dev/null 2>&1 || exit 1
MOCOB200_GRAPH=0 timeout 600 ncu --set full --clock-control none --cache-control none \
    -k regex:"k_points_fd_flat_pair|k_points_fd_vel_pair|k_wlocal_central|k_wpoints_nominal|k_assemble_jac" -c 5 -o /tmp/c5w \
    python bench.py --steps 1 --warmup 3 --workload config5 --cpu-seconds 0.2 > $O/ncu_c5w.log 2>&1; echo ncu rc=$?
ncu -i /tmp/c5w.ncu-rep --page raw --csv > $O/ncu_r02g_config5_warm_raw.csv 2>/dev/null
python scripts/ncu_summary.py $O/ncu_r02g_config5_warm $O/ncu_r02g_config5_warm_raw.csv | cut -c1-330
