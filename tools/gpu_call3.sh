#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_seams_gpu.py tests/test_render_gpu.py tests/test_full_size_gpu.py tests/test_adaptive_gpu.py -m gpu -x -q > gpurun_out/c3_pytest.log 2>&1; echo "pytest rc $?" >> gpurun_out/c3_pytest.log
AB_MODE_B=1 timeout 900 python tools/gpu_ab.py --stress-spp 8 librtb200.so > gpurun_out/c3_ab.jsonl 2>&1
NCU="ncu --set full --import-source on --clock-control none"
timeout 600 $NCU -k regex:path_kernel -s 1 -c 1 -o gpurun_out/c3_path_bunny_tile -f python tools/gpu_profile_step.py bunny > gpurun_out/c3_ncu_bunny.log 2>&1
timeout 600 $NCU -k regex:path_kernel -s 1 -c 1 -o gpurun_out/c3_path_stress_tile -f python tools/gpu_profile_step.py stress 4 > gpurun_out/c3_ncu_stress.log 2>&1
tail -3 gpurun_out/c3_pytest.log; cat gpurun_out/c3_ab.jsonl
