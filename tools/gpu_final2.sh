#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
NCU="ncu --set full --import-source on --clock-control none"
timeout 600 $NCU -k regex:path_kernel -s 1 -c 1 -o gpurun_out/c15_path_rtiow -f python tools/gpu_profile_step.py rtiow > gpurun_out/c15_ncu_rtiow.log 2>&1
timeout 600 $NCU -k regex:path_kernel -s 1 -c 1 -o gpurun_out/c15_path_stress -f python tools/gpu_profile_step.py stress 4 > gpurun_out/c15_ncu_stress.log 2>&1
bash tools/gpu_final.sh
