#!/bin/bash
# bench.py at N = 1, 2, 4, 8 GPUs of one box, launched the way the driver launches it; JSON lines into gpurun_out/.
#   gpurun --gpus 8 --timeout 900 -- 'bash tools/gpu_scaling.sh r1z'
tag=${1:-scale}
mkdir -p gpurun_out
timeout 300 python bench.py --gpus 1 > gpurun_out/${tag}_scale_1.json 2> gpurun_out/${tag}_scale_1.err
for n in 2 4 8; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) \
    bench.py --gpus $n --no-cpu-baseline > gpurun_out/${tag}_scale_$n.json 2> gpurun_out/${tag}_scale_$n.err
done
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29600 \
  bench.py --impl reference --gpus 8 --steps 2 --warmup 1 > gpurun_out/${tag}_ref_8.json 2> gpurun_out/${tag}_ref_8.err
python - "$tag" <<'PY'
import json, sys
tag = sys.argv[1]
for n in (1, 2, 4, 8):
    try:
        d = json.loads(open(f"gpurun_out/{tag}_scale_{n}.json").read().strip().splitlines()[-1])
        e = d["e2e"]
        print(f"N={n}  {d['value']:9.1f} Msamples/s  {d['mrays_per_s']:9.1f} Mrays/s  {d['ms_per_step']:7.3f} ms/step   e2e {e['value']:9.1f} ({e['ms_per_step']:.2f} ms)  sm {d['clocks']['sm_mhz']} MHz {d['clocks']['reasons']}")
    except Exception as ex:
        print("N=%d failed: %s" % (n, ex))
try:
    d = json.loads(open(f"gpurun_out/{tag}_ref_8.json").read().strip().splitlines()[-1])
    print("reference arm under torchrun x8:", d.get("value"), d.get("unit"), d.get("cpu_baseline"))
except Exception as ex:
    print("reference arm failed:", ex)
PY
