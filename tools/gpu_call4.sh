#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 900 python tools/gpu_ab.py --stress-spp 8 librtb200.so librtb200_smajor.so > gpurun_out/c4_ab.jsonl 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/c4_pytest.log 2>&1; echo "pytest rc $?" >> gpurun_out/c4_pytest.log
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/c4_bench_bunny.json 2> gpurun_out/c4_bench_bunny.err
tail -3 gpurun_out/c4_pytest.log; cat gpurun_out/c4_ab.jsonl; tail -c 600 gpurun_out/c4_bench_bunny.err
