#!/bin/bash
# round 2, GPU call 1: full GPU test suite, pool-vs-lane A/B on every config, tuning variants, L2 microbenchmark, ncu of the new kernel
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/c1_smi.txt
python -c "
import __graft_entry__ as g
rt = g.load_package()
print(rt.measure_peaks(l2_mib=(64, 32, 96, 16, 8)))
" > gpurun_out/c1_peaks.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q -s > gpurun_out/c1_pytest.log 2>&1; echo "pytest rc $?" >> gpurun_out/c1_pytest.log
timeout 600 python tools/gpu_ab.py librtb200.so > gpurun_out/c1_ab_default.jsonl 2>&1
timeout 900 python tools/gpu_ab.py --configs bunny,cornell,rtiow,stress librtb200_swap4.so librtb200_swap10.so librtb200_swap16.so librtb200_slots48.so librtb200_slots96.so librtb200_mb3.so librtb200_vote1.so librtb200_exit2.so > gpurun_out/c1_ab_variants.jsonl 2>&1
timeout 600 ncu --set full --import-source on --clock-control none -k regex:pool_kernel -s 1 -c 1 -o gpurun_out/c1_pool_bunny -f python tools/gpu_profile_step.py bunny > gpurun_out/c1_ncu_bunny.log 2>&1
tail -3 gpurun_out/c1_pytest.log; cat gpurun_out/c1_ab_default.jsonl | head -30
