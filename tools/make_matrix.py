"""profiles/r02_matrix.json from the per-N bench lines (tools/gpu_matrix.sh): C1-C5 x N in {1, 2, 4, 8} - Msamples/s, Mrays/s,
ms per step, end to end, the weak figure, strong-scaling speed-up over N = 1, and at N = 1 the binding roof and the CPU baseline
of the same run."""
import json
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
out = {"_what": "bench.py --lean lines, one per (config, N); N > 1: spp SPLIT (strong scaling, one image of the config's spp), torchrun, NCCL reduce",
       "configs": {}}
for n in (1, 2, 4, 8):
    f = ROOT / "profiles" / f"r02_matrix_n{n}.jsonl"
    if not f.exists():
        continue
    for ln in f.read_text().split("\n"):
        if not ln.startswith("{"):
            continue
        d = json.loads(ln)
        key = d["config"]["workload"].split(" ")[0]
        c = out["configs"].setdefault(key, {"workload": d["config"]["workload"], "by_n": {}})
        row = {"msamples_per_s": round(d["value"], 1), "mrays_per_s": round(d["mrays_per_s"], 1), "ms_per_step": round(d["ms_per_step"], 3),
               "e2e_msamples_per_s": round(d["e2e"]["value"], 1), "e2e_ms_per_step": round(d["e2e"]["ms_per_step"], 3),
               "reduce_and_resolve_ms": round(d.get("reduce_and_resolve_ms_rank0", 0.0), 3), "steps": d["steps"],
               "clocks_sm_mhz": (d.get("clocks") or {}).get("sm_mhz"), "throttle_reasons": (d.get("clocks") or {}).get("reasons")}
        if d.get("weak"):
            row["weak_msamples_per_s"] = round(d["weak"]["msamples_per_s"], 1)
        if n == 1:
            r = d.get("roofline") or {}
            row["binding_roof"] = {"bound": r.get("bound"), "frac": r.get("frac") and round(r["frac"], 3),
                                   "roofs": {k: round(v["frac"], 3) for k, v in (r.get("roofs") or {}).items()}, "dram_traffic_bytes_per_launch": r.get("traffic")}
            cb = d.get("cpu_baseline") or {}
            row["cpu_baseline"] = {k: cb.get(k) for k in ("value", "unit", "cores", "kind", "sample", "build")}
            if cb.get("value"):
                row["gpu_over_cpu"] = round(d["value"] / cb["value"], 1)
                row["gpu_over_cpu_e2e"] = round(d["e2e"]["value"] / cb["value"], 1)
        c["by_n"][str(n)] = row
for c in out["configs"].values():
    base = c["by_n"].get("1")
    for n, row in c["by_n"].items():
        if base:
            row["speedup_vs_1"] = round(row["msamples_per_s"] / base["msamples_per_s"], 2)
            row["e2e_speedup_vs_1"] = round(row["e2e_msamples_per_s"] / base["e2e_msamples_per_s"], 2)
(ROOT / "profiles" / "r02_matrix.json").write_text(json.dumps(out, indent=1))
for k, c in out["configs"].items():
    print(k, {n: (r["msamples_per_s"], r.get("speedup_vs_1")) for n, r in c["by_n"].items()}, c["by_n"].get("1", {}).get("binding_roof"), c["by_n"].get("1", {}).get("gpu_over_cpu"))
