#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
AB_COST_ORDER=1 timeout 900 python tools/gpu_ab.py --stress-spp 8 librtb200.so > gpurun_out/c6_cost_order_ab.jsonl 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/c6_pytest.log 2>&1; echo "pytest rc $?" >> gpurun_out/c6_pytest.log
cat gpurun_out/c6_cost_order_ab.jsonl; tail -5 gpurun_out/c6_pytest.log
