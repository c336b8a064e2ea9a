#!/bin/bash
# round 2, GPU call 2: seam tests, ncu of the default kernel on all five configs (+ smem-top pair, batch, adaptive), sanitizer logs, bench
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_seams_gpu.py tests/test_device_build_gpu.py tests/test_render_gpu.py -m gpu -x -q -s > gpurun_out/c2_pytest.log 2>&1; echo "pytest rc $?" >> gpurun_out/c2_pytest.log
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/c2_bench_bunny.json 2> gpurun_out/c2_bench_bunny.err
NCU="ncu --set full --import-source on --clock-control none"
for cfg in bunny rtiow cornell earth; do
  timeout 600 $NCU -k regex:path_kernel -s 1 -c 1 -o gpurun_out/c2_path_$cfg -f python tools/gpu_profile_step.py $cfg > gpurun_out/c2_ncu_$cfg.log 2>&1
done
timeout 600 $NCU -k regex:path_kernel -s 1 -c 1 -o gpurun_out/c2_path_stress -f python tools/gpu_profile_step.py stress 4 > gpurun_out/c2_ncu_stress.log 2>&1
timeout 600 $NCU -k regex:path_kernel -s 1 -c 1 -o gpurun_out/c2_path_bunny_smemtop -f python tools/gpu_profile_step.py bunny 0 A smemtop > gpurun_out/c2_ncu_smemtop.log 2>&1
timeout 600 $NCU -k regex:intersect_batch_kernel -s 1 -c 1 -o gpurun_out/c2_batch_bunny -f python tools/gpu_profile_step.py bunny 0 A batch > gpurun_out/c2_ncu_batch.log 2>&1
timeout 600 $NCU -k regex:adaptive_replay_kernel -s 3 -c 1 -o gpurun_out/c2_replay_bunny -f python tools/gpu_profile_step.py bunny 0 B adaptive > gpurun_out/c2_ncu_replay.log 2>&1
timeout 900 compute-sanitizer --tool memcheck --print-limit 20 python tools/gpu_sanitize.py > gpurun_out/c2_memcheck.log 2>&1
timeout 900 compute-sanitizer --tool racecheck --print-limit 20 python tools/gpu_sanitize.py > gpurun_out/c2_racecheck.log 2>&1
tail -3 gpurun_out/c2_pytest.log; tail -2 gpurun_out/c2_memcheck.log; tail -2 gpurun_out/c2_racecheck.log; ls -la gpurun_out | tail -25
