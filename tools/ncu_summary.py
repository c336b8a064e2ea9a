"""Summarise an .ncu-rep: key raw metrics per launch + instruction share per source region.
  python tools/ncu_summary.py gpurun_out/x.ncu-rep [--lines N]"""
import csv, io, subprocess, sys, collections

rep = sys.argv[1]
top = int(sys.argv[sys.argv.index("--lines") + 1]) if "--lines" in sys.argv else 25
KEYS = ["gpu__time_duration.sum", "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "smsp__cycles_active.avg",
        "sm__cycles_elapsed.max", "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "l1tex__t_sector_hit_rate.pct", "l1tex__t_sector_pipe_lsu_mem_global_op_ld_hit_rate.pct", "lts__t_sector_hit_rate.pct",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__throughput.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio", "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio", "smsp__average_warp_latency_per_inst_issued.ratio",
        "smsp__warps_eligible.avg.per_cycle_active", "smsp__sass_inst_executed_op_local_ld.sum", "smsp__sass_inst_executed_op_local_st.sum",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum"]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, data = rows[0], rows[1], rows[2:]
print("kernel:", [d[hdr.index("Kernel Name")][:60] for d in data])
if "--json" in sys.argv:  # machine-readable counters of the FIRST launch in the report (bench.py reads profiles/r02_kernel_counts.json)
    import json
    def val(k):
        if k not in hdr:
            return None
        i = hdr.index(k)
        v = float(data[0][i].replace(",", ""))
        u = units[i].lower()
        if k.startswith("dram__bytes"):
            v *= {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}.get(u, 1)
        if k == "gpu__time_duration.sum":
            v *= {"ns": 1e-6, "us": 1e-3, "usecond": 1e-3, "ms": 1, "msecond": 1, "s": 1e3, "second": 1e3, "nsecond": 1e-6}.get(u, 1)
        return v
    out = {"kernel": data[0][hdr.index("Kernel Name")][:80], "ms_under_ncu": val("gpu__time_duration.sum"),
           "warp_inst": val("smsp__inst_executed.sum"), "lanes_per_inst": val("smsp__thread_inst_executed_per_inst_executed.ratio"),
           "issue_active_pct": val("smsp__issue_active.avg.pct_of_peak_sustained_active"),
           "dram_bytes": (val("dram__bytes_read.sum") or 0) + (val("dram__bytes_write.sum") or 0),
           "l1_global_ld_hit_pct": val("l1tex__t_sector_pipe_lsu_mem_global_op_ld_hit_rate.pct"), "l2_hit_pct": val("lts__t_sector_hit_rate.pct"),
           "local_ld": val("smsp__sass_inst_executed_op_local_ld.sum"), "local_st": val("smsp__sass_inst_executed_op_local_st.sum"),
           "registers": val("launch__registers_per_thread"), "source": rep.split("/")[-1]}
    print("JSON " + json.dumps(out))
for k in KEYS:
    if k in hdr:
        i = hdr.index(k)
        print(f"{k:90s} {units[i]:14s} " + " | ".join(d[i] for d in data))
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--launch-count", "1"], capture_output=True, text=True).stdout
cur, out = None, []
for r in csv.reader(io.StringIO(src)):
    if r and r[0] == "File Path":
        cur = r[1].split("/")[-1]
        continue
    if len(r) < 10 or r[0] in ("Line No", "Function Name", ""):
        continue
    try:
        out.append((cur, int(r[0]), r[1].strip()[:80], int(r[4]), int(r[7]), int(r[8])))
    except ValueError:
        pass
ti, ts = sum(o[4] for o in out), sum(o[3] for o in out)
print(f"\nsource lines: {ti/1e6:.1f} M warp instructions, {ts} samples (first launch)")
for o in sorted(out, key=lambda o: -o[4])[:top]:
    print(f"{o[0][:16]:16s}{o[1]:4d} inst {100*o[4]/ti:5.1f}% samp {100*o[3]/max(ts,1):5.1f}% lanes {o[5]/max(o[4],1):5.1f} | {o[2]}")

# ---- instruction share per phase of the traversal / path kernels (source line ranges found by marker text)
if "--phases" in sys.argv:
    from pathlib import Path
    root = Path(__file__).resolve().parent.parent / "3360-ray-tracer_b200" / "csrc"
    K = (root / "rtb_kernels.cuh").read_text().split("\n")
    D = (root / "rtb_device.cuh").read_text().split("\n")
    def find(lines, pat):
        for i, l in enumerate(lines):
            if pat in l:
                return i + 1
        return 10 ** 9
    kr = {"node_step": (find(K, "auto node_step = [&]()"), find(K, "auto leaf_step = [&]")),
          "leaf_step": (find(K, "auto leaf_step = [&]"), find(K, "// ---- service phase")),
          "service phase (store/refill control)": (find(K, "// ---- service phase"), find(K, "if (n_trace == 0) break;")),
          "node loop vote": (find(K, "// ---- node phase: descend"), find(K, "// ---- leaf phase, one primitive type")),
          "leaf loop vote": (find(K, "// ---- leaf phase, one primitive type"), find(K, "// ---- leaf phase, one primitive type") + 12),
          "generate_sample": (find(K, "__device__ __forceinline__ void generate_sample"), find(K, "// Slot q_offset + k")),
          "shade_path": (find(K, "__device__ __forceinline__ bool shade_path"), find(K, "// ---- path kernel")),
          "stack / pop / finish": (find(K, "struct TravStack"), find(K, "auto node_step = [&]()"))}
    dr = {"philox": (find(D, "// Philox4x32-10"), find(D, "// Per-lane ray state")),
          "slab + prepare_ray": (find(D, "// Per-lane ray state"), find(D, "// Exact replays.")),
          "exact replays": (find(D, "// Exact replays."), find(D, "// Filters.  Result classes")),
          "triangle_filter": (find(D, "__device__ __forceinline__ int triangle_filter"), find(D, "// Sphere: the cancellation-prone")),
          "sphere_filter": (find(D, "__device__ __forceinline__ int sphere_filter"), find(D, "// Rect: FP32 throughout")),
          "rect_filter": (find(D, "__device__ __forceinline__ int rect_filter"), find(D, "__device__ __forceinline__ uint32_t rank_of")),
          "consider / tie rule": (find(D, "__device__ __forceinline__ uint32_t rank_of"), 10 ** 9)}
    def cat(f, l):
        table = kr if f.startswith("rtb_kernels") else dr if f.startswith("rtb_device") else None
        if f.startswith("rtb_shade"):
            return "shade_path"
        if table:
            for k, (a, b) in table.items():
                if a <= l < b:
                    return k
            return "other " + f
        return "intrinsic headers"
    S = collections.defaultdict(lambda: [0, 0, 0])
    for f, l, _, samp, inst, thr in out:
        c = cat(f, l)
        S[c][0] += samp; S[c][1] += inst; S[c][2] += thr
    print("\nphase                                  warp-inst  samples  active lanes")
    for k, v in sorted(S.items(), key=lambda kv: -kv[1][1]):
        print(f"{k:38s} {100*v[1]/ti:7.1f}% {100*v[0]/max(ts,1):7.1f}% {v[2]/max(v[1],1):8.1f}")
