"""Update profiles/r02_kernel_counts.json (read by bench.py's issue roof) from an ncu summary and the log of the program ncu wrapped:
  python tools/make_kernel_counts.py path:cornell gpurun_out/c16_path_cornell.txt gpurun_out/c16_ncu_cornell.log [...]
summary = tools/ncu_summary.py <rep> --json (its "JSON {...}" line); log = tools/gpu_profile_step.py's output ("<cfg> rep 1 {'samples': .., 'rays': ..}")."""
import ast, json, re, sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
dst = ROOT / "profiles" / "r02_kernel_counts.json"
counts = json.loads(dst.read_text()) if dst.exists() else {}
args = sys.argv[1:]
for key, summary, log in zip(args[0::3], args[1::3], args[2::3]):
    js = next(json.loads(ln[5:]) for ln in open(summary) if ln.startswith("JSON "))
    rep = next(ast.literal_eval(re.search(r"\{.*\}", ln).group(0)) for ln in open(log) if " rep 1 " in ln)
    js["samples"], js["rays"] = int(rep["samples"]), int(rep["rays"])
    js["warp_inst_per_ray"] = js["warp_inst"] / js["rays"]
    js["thread_inst_per_ray"] = js["warp_inst"] * js["lanes_per_inst"] / js["rays"]
    js["dram_bytes_per_ray"] = js["dram_bytes"] / js["rays"]
    counts[key] = js
    print(key, {k: (round(v, 2) if isinstance(v, float) else v) for k, v in js.items() if k in ("kernel", "ms_under_ncu", "lanes_per_inst", "issue_active_pct", "warp_inst_per_ray", "thread_inst_per_ray")})
dst.write_text(json.dumps(counts, indent=1))
