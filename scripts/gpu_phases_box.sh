#!/bin/bash
# Phase stopwatch on the box: the library built with MOCOB200_PROFILE_PHASES=1 (opensim-moco_b200/build_prof/, built beside the
# product library, never loaded by the product) swapped into a scratch copy of the repo. Build it first, in the build container:
#   cp -r /root/repo /tmp/prof && (cd /tmp/prof && MOCOB200_PROFILE_PHASES=1 python opensim-moco_b200/build.py --force) &&
#   mkdir -p opensim-moco_b200/build_prof && cp /tmp/prof/opensim-moco_b200/libmocob200.so opensim-moco_b200/build_prof/
rm -rf /tmp/prof && mkdir -p /tmp/prof && cp -r opensim-moco_b200 include oracle tests scripts __graft_entry__.py bench.py /tmp/prof/
cp opensim-moco_b200/build_prof/libmocob200.so /tmp/prof/opensim-moco_b200/libmocob200.so
mkdir -p gpurun_out
cd /tmp/prof && python scripts/gpu_phases.py > /root/repo/gpurun_out/phases_r02e.txt 2>&1
cat /root/repo/gpurun_out/phases_r02e.txt | cut -c1-400
