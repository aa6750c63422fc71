#!/bin/bash
# ncu --set full captures of the point-function kernels of both bench workloads, exported on the box as CSV (the .ncu-rep
# files are tens of MB; gpurun_out/ is limited to 64 MiB). Run under gpurun from the repo root, after a plain bench run.
export MOCOB200_DEVELOPER_SWITCHES=1  # the library ignores its A-B switches without it
mkdir -p gpurun_out
python bench.py --steps 5 --warmup 3 > /dev/null 2>&1 || exit 1
MOCOB200_GRAPH=0 ncu --set full --clock-control none --import-source on -k regex:"k_wpoints_fd1" -c 8 -o /tmp/c3 \
    python bench.py --steps 1 --warmup 3 > gpurun_out/ncu_c3.log 2>&1
ncu -i /tmp/c3.ncu-rep --page raw --csv > gpurun_out/ncu_r02_config3_raw.csv 2>/dev/null
ncu -i /tmp/c3.ncu-rep --page source --csv --kernel-name regex:"k_wpoints_fd1" --launch-skip 4 --launch-count 1 \
    > gpurun_out/ncu_r02_fd1_0_source.csv 2>/dev/null
MOCOB200_GRAPH=0 ncu --set full --clock-control none --import-source on \
    -k regex:"k_points_fd_flat_pair|k_points_fd_vel_pair|k_wlocal_central|k_wpoints_nominal|k_assemble_jac" -c 5 -o /tmp/c5 \
    python bench.py --steps 1 --warmup 3 --workload config5 > gpurun_out/ncu_c5.log 2>&1
ncu -i /tmp/c5.ncu-rep --page raw --csv > gpurun_out/ncu_r02_config5_raw.csv 2>/dev/null
MOCOB200_GRAPH=0 ncu --set full --clock-control none --import-source on -k regex:"k_wpoints|k_wlocal|k_assemble_jac" -c 5 -o /tmp/c4 \
    python bench.py --steps 1 --warmup 3 --workload config4 > gpurun_out/ncu_c4.log 2>&1
ncu -i /tmp/c4.ncu-rep --page raw --csv > gpurun_out/ncu_r02_config4_raw.csv 2>/dev/null
ls -la gpurun_out/*.csv
