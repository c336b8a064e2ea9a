"""Closest-hit fuzz against the compiled reference with seeds the test-suite does not use: per scene, random rays from
jittered eyes, two generations of bounce rays, grazing rays along box faces.  Prints mismatch counts (all must be 0).
  python tools/gpu_fuzz_hits.py [n_rays] [n_seeds]"""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))
import __graft_entry__ as g
import helpers
from oracle import checker
from test_host_scene import SCENES
from golden.make_golden import EYES

rt = g.load_package()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000
seeds = int(sys.argv[2]) if len(sys.argv) > 2 else 3
bad_total = 0
for name, make in SCENES.items():
    hs = make(rt)
    ds = rt.DeviceScene(hs)
    ora = checker.make_oracle(hs)
    eye, spread, jitter = EYES[name]
    rays_total = bad = 0
    for seed in range(1000, 1000 + seeds):
        batches = [helpers.ray_batch(rt, n, seed, eye=eye, spread=spread, jitter=jitter)]
        t, gid = ora.intersect(batches[0])
        b1 = helpers.secondary_rays(rt, batches[0], t, gid >= 0, seed + 50)
        if len(b1):
            batches.append(b1)
            t1, g1 = ora.intersect(b1)
            b2 = helpers.secondary_rays(rt, b1, t1, g1 >= 0, seed + 60)
            if len(b2):
                batches.append(b2)
        # axis-parallel rays through random points: they graze box faces and run inside slabs
        rng = np.random.default_rng(seed)
        r = np.zeros(n // 4, rt.RAY_DTYPE)
        o = rng.uniform(-4, 4, (len(r), 3)) + np.asarray(eye) * 0.0
        d = np.zeros((len(r), 3)); d[np.arange(len(r)), rng.integers(0, 3, len(r))] = rng.choice([-1.0, 1.0], len(r))
        r["o"], r["d"], r["tmin"], r["tmax"] = o.astype(np.float32), d.astype(np.float32), 0.001, np.inf
        batches.append(r)
        for rays in batches:
            t_ref, gid_ref = ora.intersect(rays)
            h = ds.intersect(rays)
            ok = (h["prim_id"] == gid_ref) & ((gid_ref < 0) | (np.abs(h["t"] - t_ref) <= 1e-5 * np.abs(t_ref)))
            bad += int((~ok).sum())
            rays_total += len(rays)
    print(f"{name:12s} {rays_total:9d} rays  mismatches {bad}", flush=True)
    bad_total += bad
print("TOTAL mismatches", bad_total)
sys.exit(1 if bad_total else 0)
