#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 600 python tools/gpu_create_trace.py > gpurun_out/c5_create_trace.txt 2>&1
timeout 600 python tools/gpu_coherence.py bunny > gpurun_out/c5_coherence_bunny.txt 2>&1
timeout 900 python tools/gpu_ab.py --configs bunny,cornell,rtiow,stress --stress-spp 8 librtb200.so librtb200_b128m7.so librtb200_b128m8.so librtb200_b192m5.so librtb200_b192m4.so librtb200_exit8.so librtb200_refill8.so librtb200_refill24.so librtb200_postpone.so > gpurun_out/c5_ab_variants.jsonl 2>&1
timeout 600 python -m pytest tests/test_device_build_gpu.py tests/test_hits_gpu.py -m gpu -x -q > gpurun_out/c5_pytest.log 2>&1; echo "pytest rc $?" >> gpurun_out/c5_pytest.log
cat gpurun_out/c5_create_trace.txt; cat gpurun_out/c5_coherence_bunny.txt; tail -3 gpurun_out/c5_pytest.log
