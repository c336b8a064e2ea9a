#!/usr/bin/env python
"""bench.py — Msamples/s (and Mrays/s) of the path-tracing hot path on N B200s, next to the
reference's CPU path, with the extend kernel's roofline.  Contract: see DESIGN.md "Measurement".

  python bench.py [--gpus N] [--steps K] [--warmup W] [--config bunny|rtiow|cornell|earth|stress] [--impl reference]
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P bench.py --gpus N ...

A "step" is one complete render of the named configuration (all pixels x spp x bounces).  At N>1 the
configuration's spp is SPLIT: rank k renders the slice multigpu.spp_slice(spp, k, N) of the global sample
indices of every pixel (scene replicated, no data-path collective), the float32 radiance accumulators are
summed onto rank 0 with one NCCL reduce and resolved there - one image of the configuration's spp, as the
reference renders it (src/integrator/sampler.h:26-34) - so the total work is fixed and scaling is "strong".
The weak figure (every rank renders the full spp: N x the samples) is measured too and reported under "weak".
"""
from __future__ import annotations

import argparse
import json
import os

HOST_CORES = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)


def pin_openmp():
    """The CPU legs (oracle/_ref: the reference's OpenMP pixel loop) wandered between 6.5 and 9.3 Msamples/s for the same render in
    round 1 (threads migrating between cores, a cold first pass).  OMP_PROC_BIND / OMP_PLACES must be in the environment BEFORE
    libgomp initialises, and they also bind the calling thread to the first place - fatal for a GPU rank (two torchrun ranks
    pinned to core 0 made a 0.15 ms NCCL reduce take 4 ms) - so only the process that runs a CPU leg calls this: `--impl reference`
    itself, and the b200 arm's cpu_baseline by running `bench.py --impl reference` as a child process."""
    os.environ.setdefault("OMP_PROC_BIND", "close")
    os.environ.setdefault("OMP_PLACES", "cores")
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))
import __graft_entry__ as graft  # noqa: E402

# BASELINE.json configs: scene builder, camera preset, image width, spp, depth.
CONFIGS = {
    "rtiow": dict(scene=("rtiow", (11, 1)), cameras="cameras_readme.json", preset="default", width=1000, spp=44, depth=10,
                  label="C1 RTIOW final scene, 484 spheres, README 'default' preset 1000x562, 44 spp, maxDepth 10"),
    "cornell": dict(scene=("cornell", ()), cameras="cameras.json", preset="cornell", width=800, spp=256, depth=20,
                    label="C2 Cornell box (6 rects + 3 spheres), 'cornell' preset 800x533, 256 spp, maxDepth 20"),
    "bunny": dict(scene=("bunny", (50,)), cameras="cameras_readme.json", preset="wide", width=800, spp=100, depth=20,
                  label="C3 Stanford bunny 69451 triangles (SAH BVH) + ground sphere, README 'wide' preset 800x450, 100 spp, maxDepth 20"),
    "earth": dict(scene=("earth_bubbles", ()), cameras="cameras_readme.json", preset="closeup", width=1200, spp=200, depth=15,
                  label="C4 earth-textured sphere + hollow-glass bubble + ground, README 'closeup' preset 1200x675, 200 spp, defocus 0.1"),
    "stress": dict(scene=("stress", (1415, 1414, 1000, 500, 2)), camera=dict(aspect_ratio=16 / 9.0, vfov=60.0, lookfrom=(0, 40, 120), lookat=(0, 0, 0)),
                   width=3840, spp=1024, depth=16,
                   label="C5 synthetic stress: 4.0M-triangle height field + 1M spheres, 3840x2160, 1024 spp, maxDepth 16"),
}
# Algorithmic bytes / flops per traversal event (DESIGN.md "Roofline model")
B_NODE, B_SPHERE, B_TRI, B_RECT, B_REF, B_RAY_IO = 32, 32, 80, 48, 4, 32 + 8
F32_NODE = 18
F64_SPHERE, F64_TRI, F64_RECT = 22, 44, 8


def common_config(cfg, width, height, spp, depth, estimator="A", overridden=False):
    """The `config` object of the bench line: identical in both arms (--impl b200 / reference) for the same workload, so that the driver can
    see that the two lines measure one thing; what differs between the arms goes under `arm`."""
    return {"workload": cfg["label"] + ("" if not overridden else f" [OVERRIDDEN: width {width}, spp {spp}]"), "image": [int(width), int(height)],
            "spp_total": int(spp), "max_depth": int(depth), "estimator": estimator}


def camera_for(rt, cfg, width=None, spp=None):
    if "preset" in cfg:
        c = rt.camera_preset(rt.ASSET_DIR / cfg["cameras"], cfg["preset"])
    else:
        c = rt.camera_config(**cfg["camera"])
    c.image_width = width or cfg["width"]
    c.samples_per_pixel = spp or cfg["spp"]
    c.max_depth = cfg["depth"]
    return c


class ClockSampler:
    """nvidia-smi clocks/throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""
    Q = "index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown," \
        "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, device_index):
        self.dev, self.rows, self.proc = device_index, [], None
        self.window = None  # (t0, t1) of the timed region; samples outside it are dropped

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "50",
                                          "-i", str(self.dev)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), [x.strip() for x in line.split(",")]))

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        rows = [r for t, r in self.rows if self.window is None or self.window[0] <= t <= self.window[1] + 0.15]
        if not rows and self.rows:  # region shorter than the sampling period: the sample nearest to it
            mid = 0.5 * (self.window[0] + self.window[1])
            rows = [min(self.rows, key=lambda tr: abs(tr[0] - mid))[1]]
        for r in rows:
            try:
                sm.append(float(r[1])); mx.append(float(r[2]))
            except (ValueError, IndexError):
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons),
                "samples": len(sm)}


def cpu_reference_leg(rt, cfg_name, budget_s, steps=1, warmup=0):
    """The reference's own CPU path (oracle/_ref: MegaKernel/DefaultSampler/GetPixel compiled from the
    reference sources; else the C restatement), all host threads, on a bounded sample of the workload."""
    from oracle import checker, ref
    cfg = CONFIGS[cfg_name]
    fast = ref.available(fast=True) and _cpu_has("avx2") and _cpu_has("fma")
    hs = rt.HostScene.build(*cfg["scene"], host_only=True)  # librth.so alone: the CUDA library is never mapped by this leg
    ora = checker.make_oracle(hs, fast=fast)
    cam_cfg = camera_for(rt, cfg)
    cam = rt.camera_init(cam_cfg)
    # every host thread this process may use: torchrun exports OMP_NUM_THREADS=1 to its workers, which would silently
    # turn the "all-core" reference into a single-thread run, so the count is passed to omp_set_num_threads explicitly
    cores = HOST_CORES
    # calibrate with whole-image passes at 1 and 2 spp (the first also warms the thread pool), then size
    # the sample to the budget: samples/s is spp-invariant
    ora.render_a(cam_cfg.as_oracle_cfg(), 1, cfg["depth"], seed=1, want_m2=False, threads=cores)
    _, _, secs, _ = ora.render_a(cam_cfg.as_oracle_cfg(), 2, cfg["depth"], seed=1, want_m2=False, threads=cores)
    rate = 2 * cam.width * cam.height / max(secs, 1e-6)
    spp = int(max(1, min(cfg["spp"], budget_s * rate / (cam.width * cam.height))))
    out = []
    for k in range(warmup + steps):
        this_spp = spp if k >= warmup else 1
        _, _, secs, rays = ora.render_a(cam_cfg.as_oracle_cfg(), this_spp, cfg["depth"], seed=2 + k, want_m2=False, threads=cores)
        if k >= warmup:
            out.append((cam.width * cam.height * this_spp, rays, secs))
    samples = sum(o[0] for o in out); rays = sum(o[1] for o in out); secs = sum(o[2] for o in out)
    host = {"nproc": os.cpu_count(), "affinity": cores, "cpu_model": _cpu_model(), "omp_proc_bind": os.environ.get("OMP_PROC_BIND"),
            "omp_places": os.environ.get("OMP_PLACES"), "warm_passes": 2 + warmup}
    return {"value": samples / secs / 1e6, "unit": "Msamples/s", "mrays_per_s": rays / secs / 1e6, "cores": cores, "host": host, "image": [cam.width, cam.height],
            "kind": ora.kind, "build": "-O3 -mavx2 -mfma -fopenmp" if fast else "-O2 -fopenmp (strict IEEE)",
            "sample": f"{cfg_name} {cam.width}x{cam.height} at {spp} of {cfg['spp']} spp, maxDepth {cfg['depth']}, estimator A, "
                      f"{len(out)} pass(es), {secs:.1f} s", "seconds": secs, "ms_per_step": 1e3 * secs / max(1, len(out))}


def cpu_baseline_child(cfg_name, budget_s):
    """cpu_baseline of the b200 arm: the reference arm of this script as a child process (pinned OpenMP threads, librth.so only,
    no CUDA library mapped), one bounded pass; returns its cpu_baseline object."""
    env = {k: v for k, v in os.environ.items() if k not in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "OMP_NUM_THREADS")}
    r = subprocess.run([sys.executable, str(ROOT / "bench.py"), "--impl", "reference", "--config", cfg_name, "--steps", "1", "--warmup", "0",
                        "--cpu-step-budget", str(budget_s)], capture_output=True, text=True, env=env, timeout=900)
    for ln in reversed(r.stdout.strip().split("\n")):
        if ln.startswith("{"):
            d = json.loads(ln)
            return dict(d["cpu_baseline"], mrays_per_s=d.get("mrays_per_s"))
    return {"value": None, "unit": "Msamples/s", "cores": HOST_CORES, "kind": "reference", "sample": "cpu leg failed: " + r.stderr[-300:]}


def _cpu_model():
    try:
        for line in Path("/proc/cpuinfo").read_text().split("\n"):
            if line.startswith("model name"):
                return line.split(":", 1)[1].strip()
    except OSError:
        pass
    return "unknown"


def _cpu_has(flag):
    try:
        return flag in Path("/proc/cpuinfo").read_text().split("flags", 1)[1].split("\n", 1)[0].split()
    except (OSError, IndexError):
        return False


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default="bunny", choices=list(CONFIGS))
    ap.add_argument("--spp", type=int, default=0, help="override the configuration's spp (then the line says so)")
    ap.add_argument("--width", type=int, default=0)
    ap.add_argument("--pool", type=int, default=0)
    ap.add_argument("--mode", default="A", choices=["A", "B"])
    ap.add_argument("--pipeline", default="path", choices=["path", "wavefront"],
                    help="path = the default persistent path kernel; wavefront = generate/extend/shade kernels over queues")
    ap.add_argument("--tree", default="auto", choices=["auto", "host", "device"],
                    help="host = the reference's SAH tree built on the host and compiled for the device; device = linear BVH built on the GPU "
                         "(RTB_SCENE_DEVICE_BUILD); auto = device for the stress scene (faster to build AND to traverse there), host otherwise")
    ap.add_argument("--cpu-budget", type=float, default=15.0, help="seconds of CPU work for the cpu_baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-step-budget", type=float, default=0.0, help="--impl reference: seconds of CPU work per step (0 = sized from --steps)")
    ap.add_argument("--no-roofline", action="store_true")
    ap.add_argument("--lean", action="store_true", help="skip the per-kernel (serial wavefront) and other-pipeline measurements and shorten e2e: "
                                                        "for the C1-C5 x N matrix (tools/gpu_matrix.sh), where only the headline numbers are wanted")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    rt = graft.load_package()
    cfg = CONFIGS[args.config]
    W = max(args.warmup, 0)
    K = max(args.steps, 1)

    if args.impl == "reference":
        if rank != 0:
            return 0
        pin_openmp()
        per_step = args.cpu_step_budget if args.cpu_step_budget > 0 else min(12.0, 150.0 / (K + W))
        leg = cpu_reference_leg(rt, args.config, per_step, steps=K, warmup=W)
        line = {"metric": "msamples_per_s", "value": leg["value"], "unit": "Msamples/s", "n_gpus": args.gpus, "steps": K, "warmup": W,
                "ms_per_step": leg["ms_per_step"], "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
                "data": "synthetic", "impl": "reference", "mrays_per_s": leg["mrays_per_s"],
                "config": common_config(cfg, leg["image"][0], leg["image"][1], cfg["spp"], cfg["depth"]),
                "arm": {"what": "the reference's own CPU path: MegaKernel + DefaultSampler + Camera::GetPixel (estimator A), all host threads", "sample": leg["sample"]},
                "cpu_baseline": {k: leg[k] for k in ("value", "unit", "cores", "kind", "sample", "build", "host")},
                "e2e": {"value": leg["value"], "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "gpu_launches": 0}
        print(json.dumps(line), flush=True)
        return 0

    import torch
    import torch.distributed as dist

    if not torch.cuda.is_available():
        raise SystemExit("bench.py --impl b200 needs a CUDA device: the hot path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    lib = rt.load()

    device_tree = args.tree == "device" or (args.tree == "auto" and args.config == "stress")
    t_build = time.perf_counter()
    hs = rt.HostScene.build(*cfg["scene"], device_tree=device_tree)
    t_build = time.perf_counter() - t_build
    ds = rt.DeviceScene(hs, device=local_rank)
    cam_cfg = camera_for(rt, cfg, args.width or None, args.spp or None)
    cam = rt.camera_init(cam_cfg)
    spp, depth = cam_cfg.samples_per_pixel, cfg["depth"]
    mode = rt.MODE_A if args.mode == "A" else rt.MODE_B
    n_pix = cam.width * cam.height
    accum = torch.zeros(n_pix * 3, dtype=torch.float32, device="cuda")
    rgb8 = torch.zeros(n_pix * 3, dtype=torch.uint8, device="cuda")
    flush = torch.empty(384 << 20, dtype=torch.uint8, device="cuda")  # > 126 MB L2
    tstream = torch.cuda.Stream()  # reduce + resolve run here, bracketed by the two events
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    base_flags = rt.RENDER_WAVEFRONT if args.pipeline == "wavefront" else 0

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step(flags=None, weak=False):
        """one render of the configuration (this rank's slice of its spp) + reduce + resolve; returns (device ms, stats, reduce ms)"""
        flags = base_flags if flags is None else flags
        b_, e_ = rt.multigpu.weak_slice(spp, rank) if weak else rt.multigpu.spp_slice(spp, rank, world)
        st = ds.render_device(cam, accum.data_ptr(), b_, e_, depth, mode=mode, pool_size=args.pool, flags=flags)  # synchronous
        with torch.cuda.stream(tstream):
            ev[0].record(tstream)
            if world > 1:
                dist.reduce(accum, dst=0, op=dist.ReduceOp.SUM)
            if rank == 0:
                rt.binding._check(lib, lib.rtb_resolve_device(accum.data_ptr(), n_pix, spp * world if weak else spp, rgb8.data_ptr(), tstream.cuda_stream))
            ev[1].record(tstream)
        torch.cuda.synchronize()
        tail = ev[0].elapsed_time(ev[1])
        return st["ms_total"] + tail, st, tail

    clocks = ClockSampler(local_rank)
    if rank == 0:
        clocks.start()  # nvidia-smi takes a few hundred ms to deliver its first sample: start it before the warm-up
    for _ in range(W):
        flush.fill_(1)
        barrier()
        step()
    dev_ms, wall_ms, stats, reduce_ms = [], [], [], []
    t_region = time.perf_counter()
    for _ in range(K):
        flush.fill_(1)  # evict the scene and the path pool from L2 between timed steps (untimed)
        barrier()
        t0 = time.perf_counter()
        ms, st, red = step()
        wall_ms.append(1e3 * (time.perf_counter() - t0))
        dev_ms.append(ms)
        stats.append(st)
        reduce_ms.append(red)
    barrier()
    clocks.window = (t_region, time.perf_counter())
    clk = clocks.stop() if rank == 0 else None
    # per-kernel times: the same step with ONE wavefront in flight and an event pair around every launch
    # (RTB_RENDER_SERIAL); in the timed steps above several wavefronts overlap, so kernels cannot be timed singly
    serial_stats = []
    for _ in range(0 if args.lean and args.pipeline == "path" else min(K, 3)):
        flush.fill_(1)
        barrier()
        serial_stats.append(step(rt.RENDER_SERIAL)[1])
    # the other pipeline, for the record (not the headline)
    other_flags = 0 if args.pipeline == "wavefront" else rt.RENDER_WAVEFRONT
    other_ms = []
    for k in range(0 if args.lean else 1 + min(K, 5)):
        flush.fill_(1)
        barrier()
        ms_o = step(other_flags)[0]
        if k:
            other_ms.append(ms_o)
    # weak scaling for the record: every rank renders the configuration's full spp (N x the samples, one reduce)
    weak_ms, weak_samples = [], 0
    if world > 1:
        for k in range(1 + min(K, 5)):
            flush.fill_(1)
            barrier()
            ms_w, st_w, _ = step(weak=True)
            if k:
                weak_ms.append(ms_w)
        wk = torch.tensor([sum(weak_ms)], dtype=torch.float64, device="cuda")
        dist.all_reduce(wk, op=dist.ReduceOp.MAX)
        weak_ms = wk.item() / max(1, min(K, 5))
        weak_samples = sum_over_ranks(dist, world, st_w["samples"])
    # max over ranks of the K-step time
    tot = torch.tensor([sum(dev_ms), sum(wall_ms)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(tot, op=dist.ReduceOp.MAX)
    total_dev_ms, total_wall_ms = tot.tolist()
    samples_per_step = sum_over_ranks(dist, world, stats[-1]["samples"])
    rays_per_step = sum_over_ranks(dist, world, stats[-1]["rays"])

    # ---- e2e: the public C-ABI call with HOST buffers: scene upload (H2D) + render + image readback (D2H), every step
    desc = hs.desc
    h2d = scene_bytes(desc)
    d2h = n_pix * 12
    e2e_ms = []
    import ctypes as C
    host_img = torch.empty(n_pix * 3, dtype=torch.float32).pin_memory()
    KE = min(K, 3 if args.lean else 10)
    e2e_parts = []
    for k in range(min(W, 1) + KE):
        barrier()
        t0 = time.perf_counter()
        ds2 = rt.DeviceScene(hs, device=local_rank)  # rtb_scene_create: copies every scene array host -> device
        t_created = time.perf_counter()
        if world == 1:
            prm = rt.binding.RtbRenderParams(0, spp, depth, mode, 0x3360, args.pool, base_flags)
            s2 = rt.binding.RtbStats()
            rt.binding._check(lib, lib.rtb_render(ds2._h, C.byref(cam), C.byref(prm), host_img.data_ptr(), C.byref(s2)))
        else:
            b_, e_ = rt.multigpu.spp_slice(spp, rank, world)
            ds2.render_device(cam, accum.data_ptr(), b_, e_, depth, mode=mode, pool_size=args.pool, flags=base_flags)
            dist.reduce(accum, dst=0, op=dist.ReduceOp.SUM)
            if rank == 0:
                host_img.copy_(accum, non_blocking=False)
        t_rendered = time.perf_counter()
        ds2.close()
        torch.cuda.synchronize()
        if k >= min(W, 1):
            t_end = time.perf_counter()
            e2e_ms.append(1e3 * (t_end - t0))
            e2e_parts.append((1e3 * (t_created - t0), 1e3 * (t_rendered - t_created), 1e3 * (t_end - t_rendered)))
    e2e_tot = torch.tensor([sum(e2e_ms)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(e2e_tot, op=dist.ReduceOp.MAX)

    if rank == 0:
        ms_per_step = total_dev_ms / K
        value = samples_per_step / (ms_per_step * 1e-3) / 1e6
        line = {
            "metric": "msamples_per_s", "value": value, "unit": "Msamples/s", "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32 shading / f64 exact hit replay", "data": "synthetic", "impl": "b200",
            "mrays_per_s": rays_per_step / (ms_per_step * 1e-3) / 1e6,
            "wall_ms_per_step": total_wall_ms / K,
            "config": common_config(cfg, cam.width, cam.height, spp, depth, args.mode, overridden=bool(args.spp or args.width)),
            "arm": {"why_this_config": "north_star quotes its headline target (>=100x CPU, traversal roofline) on the bunny mesh scene",
                    "spp_per_gpu": [list(rt.multigpu.spp_slice(spp, r, world)) for r in range(world)],
                    "parallelism": f"spp SPLIT x{world} (one image of {spp} spp), scene replicated, 1 NCCL reduce(sum,f32) of {n_pix * 12} B onto rank 0, resolve there",
                    "l2": "384 MiB buffer written between timed steps (L2 flushed); scene itself is L2-resident by design",
                    "pipeline": "persistent path kernel (generate + traverse + shade per lane, one launch per step)" if args.pipeline == "path"
                    else f"wavefront: generate / extend / shade+compact over queues, pool {args.pool or (1 << 21)} paths x {os.environ.get('RTB_WAVES', '4')} wavefronts in flight",
                    "tree": "linear BVH built on the GPU inside rtb_scene_create (RTB_SCENE_DEVICE_BUILD)" if device_tree
                    else "host SAH tree (the reference's algorithm) compiled to sibling pairs",
                    "scene_build_s_untimed": round(t_build, 3)},
            "e2e": {"value": samples_per_step / (e2e_tot.item() / KE * 1e-3) / 1e6, "unit": "Msamples/s", "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": d2h, "ms_per_step": e2e_tot.item() / KE, "steps": KE,
                    "ms_parts_rank0": dict(zip(("scene_create", "render_and_readback", "scene_destroy"), (float(x) for x in np.mean(e2e_parts, axis=0)))),
                    "what": "rtb_scene_create (all scene arrays H2D) + rtb_render with a host rgb buffer (D2H) + rtb_scene_destroy, wall clock"},
            "gpu_launches": int(sum(s["launches"] for s in stats) + K),
            "other_pipeline": None if not other_ms else {"name": "path" if args.pipeline == "wavefront" else "wavefront", "ms_per_step": float(np.mean(other_ms)),
                                                          "msamples_per_s": samples_per_step / float(np.mean(other_ms)) / 1e3},
            "reduce_and_resolve_ms_rank0": float(np.mean(reduce_ms)),
            "weak": None if world == 1 else {"what": f"every rank renders the full {spp} spp ({world}x the samples), one reduce", "ms_per_step": weak_ms,
                                              "msamples_per_s": weak_samples / (weak_ms * 1e-3) / 1e6, "spp_total": spp * world},
            "clocks": clk,
            "kernel_ms_per_step": {"path_kernel": float(np.mean([s["ms_total"] for s in stats])) if args.pipeline == "path" else None,
                                   "wavefront_serial": None if not serial_stats else {k: float(np.mean([s[k] for s in serial_stats])) for k in ("ms_generate", "ms_extend", "ms_shade", "ms_drain", "ms_total")},
                                   "how": "path_kernel: CUDA events around the one launch of each timed step; wavefront_serial: the wavefront pipeline re-run with one "
                                          "wavefront in flight (RTB_RENDER_SERIAL) and events around every launch"},
        }
        if not args.no_roofline and world == 1:
            line["roofline"] = roofline(rt, hs, cam, cfg, stats if args.pipeline == "path" else serial_stats, mode, args.pool, args.pipeline, device_tree,
                                        sm_mhz=(clk or {}).get("sm_mhz"))
        elif not args.no_roofline:
            line["roofline"] = {"note": "measured at N=1 only (one process, one GPU)"}
        if not args.no_cpu_baseline and world == 1:
            line["cpu_baseline"] = cpu_baseline_child(args.config, args.cpu_budget)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def sum_over_ranks(dist, world, v):
    if world == 1:
        return int(v)
    import torch
    t = torch.tensor([int(v)], dtype=torch.int64, device="cuda")
    dist.all_reduce(t)
    return int(t.item())


def scene_bytes(d):
    return int(d.n_nodes * 32 + d.n_prim_refs * 4 + d.n_spheres * (32 + 16 + 12) + d.n_triangles * (80 + 48 + 8) + d.n_rects * (48 + 8)
               + d.n_materials * 32 + d.n_textures * 32)


def roofline(rt, hs, cam, cfg, stats, mode, pool, pipeline, device_tree=False, sm_mhz=None):
    """The dominant kernel - path_kernel (default pipeline: one launch per step) or extend_kernel (wavefront pipeline) -
    against three roofs; `bound` / `achieved` / `peak` / `frac` at the top level are those of the roof that BINDS (the
    largest model time), the HBM form the bench contract describes stays under roofs.hbm:
      hbm    algorithmic bytes (SURVEY 8d per-ray figure x rays of the launch + accumulator RMW) / launch time against
             MEASURED_PEAKS.json hbm_gbs.  For C1-C4 the scene is L2-resident and this roof does not bind (`traffic`,
             the DRAM bytes ncu measured for the launch, is ~1000x below the algorithmic bytes);
      l2     the same bytes against the L2 read bandwidth measured in this run with an ld.global.cg sweep of 64 MiB
             (cannot hit the L1s; round 1 used __ldg over 32 MiB = 221 KB per SM, which partly measured the L1);
      issue  executed thread-instructions of the launch against 4 warp-instructions x 32 lanes per clock per SM: the
             fraction of the SMs' lane-issue slots that did work = (issue-slot use) x (active lanes / 32).  Instruction
             counts per ray come from the committed ncu capture of this kernel on this config
             (profiles/r02_kernel_counts.json, from profiles/r02_path_kernel_<config>.txt - NOT measured in-run).
    Traversal counts per ray are measured in this run with the instrumented build (librtb200_counters.so)."""
    import torch
    peaks_file = ROOT / "MEASURED_PEAKS.json"
    hbm_peak, src, sm_max = 6650.0, "fallback (B200_PROFILING.md)", 1965.0
    if peaks_file.exists():
        mpj = json.loads(peaks_file.read_text())
        hbm_peak, src, sm_max = float(mpj["hbm_gbs"]), "MEASURED_PEAKS.json", float(mpj.get("sm_max_mhz", sm_max))
    mp = rt.measure_peaks()
    n_sm = torch.cuda.get_device_properties(torch.cuda.current_device()).multi_processor_count
    f_sm = 1e6 * float(sm_mhz or sm_max)
    # per-ray traversal counts from the instrumented build on the same workload at reduced spp (counts per ray are spp-invariant)
    hs_c = rt.HostScene.build(*cfg["scene"], counters=True, device_tree=device_tree)
    ds_c = rt.DeviceScene(hs_c)
    _, sc = ds_c.render(cam, spp=max(1, min(4, cfg["spp"])), max_depth=cfg["depth"], mode=mode, pool_size=pool,
                        flags=rt.RENDER_WAVEFRONT if pipeline == "wavefront" else 0)
    ds_c.close()
    per_ray = {k: sc[k] / sc["rays"] for k in ("node_visits", "sphere_tests", "triangle_tests", "rect_tests")}
    refs = per_ray["sphere_tests"] + per_ray["triangle_tests"] + per_ray["rect_tests"]
    io_ray = B_RAY_IO if pipeline == "wavefront" else 0  # the path kernel keeps rays and hits in registers
    bytes_ray = (B_NODE * per_ray["node_visits"] + B_SPHERE * per_ray["sphere_tests"] + B_TRI * per_ray["triangle_tests"]
                 + B_RECT * per_ray["rect_tests"] + B_REF * refs + io_ray)
    f32_ray = F32_NODE * per_ray["node_visits"]
    f64_ray = F64_SPHERE * per_ray["sphere_tests"] + F64_TRI * per_ray["triangle_tests"] + F64_RECT * per_ray["rect_tests"]
    rays = float(np.mean([s["rays"] for s in stats]))
    samples = float(np.mean([s["samples"] for s in stats]))
    if pipeline == "path":
        kernel, launches = "path_kernel (generate + closest-hit BVH traversal + shade, persistent)", float(np.mean([s["launches"] for s in stats]))
        k_ms = float(np.mean([s["ms_total"] for s in stats]))
        extra = 12.0 * 2 * samples  # >= one float3 read-modify-write of the accumulator per sample
    else:
        kernel, launches = "extend_kernel (closest-hit BVH traversal)", float(np.mean([s["iterations"] for s in stats]))
        k_ms = float(np.mean([s["ms_extend"] for s in stats]))
        extra = 0.0
    total_bytes = bytes_ray * rays + extra
    secs = k_ms * 1e-3
    gbs = total_bytes / secs / 1e9
    roofs = {"hbm": {"achieved": gbs, "peak": hbm_peak, "unit": "GB/s", "frac": gbs / hbm_peak, "model_ms": 1e3 * total_bytes / (hbm_peak * 1e9),
                     "peak_source": src},
             "l2": {"achieved": gbs, "peak": mp["l2_gbs"], "unit": "GB/s", "frac": gbs / mp["l2_gbs"], "model_ms": 1e3 * total_bytes / (mp["l2_gbs"] * 1e9),
                    "peak_source": "ld.global.cg read sweep of 64 MiB, this run (rtb_measure_l2_gbs)"},
             "fp": {"achieved": (f32_ray + f64_ray) * rays / secs / 1e12, "unit": "TFLOP/s", "fp32_peak": mp["fp32_tflops"], "fp64_peak": mp["fp64_tflops"],
                    "model_ms": 1e3 * (f32_ray * rays / (mp["fp32_tflops"] * 1e12) + f64_ray * rays / (mp["fp64_tflops"] * 1e12))}}
    roofs["fp"]["frac"] = roofs["fp"]["model_ms"] / k_ms
    traffic, counts_note = None, None
    cf = ROOT / "profiles" / "r02_kernel_counts.json"  # ncu --set full capture of this kernel per config (committed; not measured in-run)
    if cf.exists():
        scene_key = {"earth_bubbles": "earth"}.get(cfg["scene"][0], cfg["scene"][0])
        kc = json.loads(cf.read_text()).get(f"{pipeline}:{scene_key}")
        if kc:
            traffic = kc["dram_bytes_per_ray"] * rays / max(launches, 1)
            lane_peak = 128.0 * n_sm * f_sm  # 4 schedulers x 32 lanes per clock per SM
            tinst = kc["thread_inst_per_ray"] * rays
            roofs["issue"] = {"achieved": tinst / secs / 1e12, "peak": lane_peak / 1e12, "unit": "T thread-instructions/s", "frac": tinst / secs / lane_peak,
                              "model_ms": 1e3 * tinst / lane_peak, "warp_issue_frac": kc["warp_inst_per_ray"] * rays / secs / (4.0 * n_sm * f_sm),
                              "lanes_per_instruction": kc["lanes_per_inst"], "thread_inst_per_ray": kc["thread_inst_per_ray"],
                              "warp_inst_per_ray": kc["warp_inst_per_ray"], "sm_count": n_sm, "sm_mhz": f_sm / 1e6,
                              "counts_source": "profiles/r02_kernel_counts.json (ncu capture " + str(kc.get("source")) + "; from profiles/, not measured in-run)"}
            roofs["hbm"]["dram_traffic_bytes_per_launch"] = traffic
            roofs["hbm"]["traffic_model_ms"] = 1e3 * traffic * max(launches, 1) / (hbm_peak * 1e9)  # what HBM actually has to deliver
            counts_note = "traffic = dram__bytes_read.sum + dram__bytes_write.sum per ray from the committed ncu capture x rays of this launch"
    # which roof binds: the largest model time, where HBM is charged with the DRAM traffic the launch really causes (ncu) when that
    # is known - an L2-resident scene does not move its algorithmic bytes through HBM
    binding = max(roofs, key=lambda k: roofs[k].get("traffic_model_ms", roofs[k]["model_ms"]))
    b = roofs[binding]
    return {"kernel": kernel, "bound": binding, "achieved": b["achieved"], "peak": b.get("peak", b.get("fp32_peak")), "unit": b["unit"], "frac": b["frac"],
            "traffic": traffic, "traffic_note": counts_note,
            "note": "bound = the roof with the largest model time; every roof is listed under roofs (hbm = the bench contract's form)",
            "roofs": roofs, "measured_peaks": mp,
            "per_ray": {**per_ray, "bytes": bytes_ray, "fp32_flops": f32_ray, "fp64_flops": f64_ray},
            "rays_per_step": rays, "launches_per_step": launches, "kernel_ms_per_step": k_ms,
            "avg_launch_ms": k_ms / max(launches, 1), "algorithmic_bytes_per_launch": total_bytes / max(launches, 1)}


if __name__ == "__main__":
    sys.exit(main())
