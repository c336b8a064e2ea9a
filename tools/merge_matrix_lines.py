"""Replace the lines of re-measured configs in profiles/r02_matrix_n<N>.jsonl:  python tools/merge_matrix_lines.py gpurun_out/r02_matrix_n1.jsonl 1"""
import json, sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
new = [json.loads(l) for l in open(sys.argv[1]) if l.startswith("{")]
dst = ROOT / "profiles" / f"r02_matrix_n{int(sys.argv[2])}.jsonl"
old = [json.loads(l) for l in open(dst) if l.startswith("{")] if dst.exists() else []
by = {d["config"]["workload"].split(" ")[0]: d for d in old}
for d in new:
    by[d["config"]["workload"].split(" ")[0]] = d
order = ["C1", "C2", "C3", "C4", "C5"]
dst.write_text("".join(json.dumps(by[k]) + "\n" for k in order if k in by))
print("merged", [d["config"]["workload"].split(" ")[0] for d in new], "into", dst.name)
