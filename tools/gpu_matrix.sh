#!/bin/bash
# The measurement matrix north_star asks for: C1-C5 x N GPUs (this call: N = $1), Msamples/s, Mrays/s, roofline, CPU baseline at N = 1.
#   gpurun --gpus N -- bash tools/gpu_matrix.sh N   ->   gpurun_out/r02_matrix_nN.jsonl (one bench line per config; CFGS="rtiow stress" = a subset, SKIP_TESTS=1 = without the multi-GPU tests)
cd "$GRAFT_REPO_ROOT" || exit 1
N=${1:-1}
mkdir -p gpurun_out
OUT=gpurun_out/r02_matrix_n$N.jsonl
: > $OUT
for cfg in ${CFGS:-rtiow cornell bunny earth stress}; do
  if [ "$cfg" = stress ]; then K=2; W=3; else K=10; W=3; fi
  if [ "$N" = 1 ]; then
    timeout 900 python bench.py --gpus 1 --steps $K --warmup $W --config $cfg --lean >> $OUT 2> gpurun_out/r02_matrix_n${N}_$cfg.err
  else
    timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps $K --warmup $W --config $cfg --lean >> $OUT 2> gpurun_out/r02_matrix_n${N}_$cfg.err
  fi
  echo "$cfg rc $?" >> gpurun_out/r02_matrix_n${N}.log
done
if [ "$N" -ge 2 ] && [ -z "$SKIP_TESTS" ]; then
  timeout 600 python -m pytest tests/test_multi_device_gpu.py tests/test_nccl_gpu.py -m gpu -q > gpurun_out/r02_multi_gpu_tests_n$N.log 2>&1
  tail -3 gpurun_out/r02_multi_gpu_tests_n$N.log
fi
cat gpurun_out/r02_matrix_n${N}.log; python - <<PY
import json
for l in open("$OUT"):
    if l.startswith("{"):
        d = json.loads(l)
        print(d["config"]["workload"][:12], d["n_gpus"], round(d["value"], 1), "Msamples/s", round(d["ms_per_step"], 3), "ms  e2e", round(d["e2e"]["value"], 1), d.get("weak") and round(d["weak"]["msamples_per_s"], 1))
PY
