"""Degenerate rays through rtb_intersect_batch against the compiled reference: zero / NaN / infinite components, empty and
inverted intervals, denormal directions.  Prints, per class, how many (id, t) pairs differ.  Run under `timeout`."""
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

rt = g.load_package()
rng = np.random.default_rng(77)
n = 4096
for scene in ("mixed", "bunny"):
    hs = SCENES[scene](rt)
    ds = rt.DeviceScene(hs)
    ora = checker.make_oracle(hs)
    o = rng.uniform(-3, 3, (n, 3))
    d = rng.normal(size=(n, 3))
    nan_o = o.copy(); nan_o[np.arange(n), rng.integers(0, 3, n)] = np.nan
    nan_d = d.copy(); nan_d[np.arange(n), rng.integers(0, 3, n)] = np.nan
    inf_o = o.copy(); inf_o[np.arange(n), rng.integers(0, 3, n)] = np.inf
    inf_d = d.copy(); inf_d[np.arange(n), rng.integers(0, 3, n)] = -np.inf
    one_den = d.copy(); one_den[np.arange(n), rng.integers(0, 3, n)] = rng.choice([1e-41, -1e-41, 1e-30, -3e-25, 0.0, -0.0], n)
    two_den = np.full((n, 3), 1e-41) * rng.choice([-1.0, 1.0], (n, 3)); two_den[np.arange(n), rng.integers(0, 3, n)] = rng.normal(size=n)
    cases = {
        "one tiny component": (o, one_den, 0.001, np.inf),
        "two tiny components": (o, two_den, 0.001, np.inf),
        "zero direction": (o, np.zeros((n, 3)), 0.001, np.inf),
        "nan origin": (nan_o, d, 0.001, np.inf),
        "nan direction": (nan_d, d * 0 + nan_d, 0.001, np.inf),
        "inf origin": (inf_o, d, 0.001, np.inf),
        "inf direction": (o, inf_d, 0.001, np.inf),
        "inverted interval": (o, d, 5.0, 1.0),
        "empty interval": (o, d, 2.0, 2.0),
        "negative tmin": (o, d, -10.0, np.inf),
        "nan tmax": (o, d, 0.001, np.nan),
        "denormal direction": (o, d * 1e-41, 0.001, np.inf),
        "all nan": (o * np.nan, d * np.nan, np.nan, np.nan),
    }
    for name, (oo, dd, tmin, tmax) in cases.items():
        r = np.zeros(n, rt.RAY_DTYPE)
        r["o"], r["d"], r["tmin"], r["tmax"] = oo.astype(np.float32), dd.astype(np.float32), tmin, tmax
        h = ds.intersect(r)
        t_ref, gid_ref = ora.intersect(r)
        same_id = h["prim_id"] == gid_ref
        same_t = (h["t"] == t_ref) | (np.isnan(h["t"]) & np.isnan(t_ref)) | ~(gid_ref >= 0)
        print(f"{scene:6s} {name:20s} ref hits {int((gid_ref >= 0).sum()):5d}  gpu hits {int((h['prim_id'] >= 0).sum()):5d}  "
              f"id differs {int((~same_id).sum()):5d}  t differs {int((same_id & ~same_t).sum()):5d}", flush=True)
