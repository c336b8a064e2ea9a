#!/bin/bash
# round 2 final single-GPU evidence: full GPU test suite, default bench line, ncu launch list of the bench command, ncu --set full of the
# dominant kernel on all five configs (-> profiles/r02_path_kernel_<config>.txt, profiles/r02_kernel_counts.json)
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/c9_pytest.log 2>&1; echo "pytest rc $?" >> gpurun_out/c9_pytest.log
timeout 600 python bench.py > gpurun_out/c9_bench_bunny.json 2> gpurun_out/c9_bench_bunny.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/c9_bench_reference.json 2> gpurun_out/c9_bench_reference.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/c9_bench_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/c9_ncu_bench.log 2>&1
NCU="ncu --set full --import-source on --clock-control none"
for cfg in bunny rtiow cornell earth; do
  timeout 600 $NCU -k regex:path_kernel -s 1 -c 1 -o gpurun_out/c9_path_$cfg -f python tools/gpu_profile_step.py $cfg > gpurun_out/c9_ncu_$cfg.log 2>&1
done
timeout 600 $NCU -k regex:path_kernel -s 1 -c 1 -o gpurun_out/c9_path_stress -f python tools/gpu_profile_step.py stress 4 > gpurun_out/c9_ncu_stress.log 2>&1
tail -3 gpurun_out/c9_pytest.log; tail -c 300 gpurun_out/c9_bench_bunny.json; echo; tail -c 400 gpurun_out/c9_bench_reference.json
