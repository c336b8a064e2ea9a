#!/bin/bash
# final single-GPU check of a round: the GPU test suite, smoke(), the default bench line, C5 end to end at 64 spp
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/final_pytest.log 2>&1; echo "pytest rc $?" >> gpurun_out/final_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/final_smoke.log 2>&1; echo "smoke rc $?" >> gpurun_out/final_smoke.log
timeout 600 python bench.py > gpurun_out/final_bench_bunny.json 2> gpurun_out/final_bench_bunny.err
timeout 600 python bench.py --config stress --spp 64 --steps 3 --warmup 3 --lean > gpurun_out/final_bench_stress_64spp.json 2> gpurun_out/final_bench_stress_64spp.err
tail -3 gpurun_out/final_pytest.log; tail -2 gpurun_out/final_smoke.log
python - <<PY
import json
for f in ("final_bench_bunny", "final_bench_stress_64spp"):
    d = json.loads(open(f"gpurun_out/{f}.json").read().strip().split("\n")[-1])
    print(f, round(d["value"], 1), round(d["ms_per_step"], 3), "e2e", round(d["e2e"]["value"], 1), d["e2e"]["ms_parts_rank0"], d["roofline"].get("bound"), d["roofline"].get("frac"))
PY
