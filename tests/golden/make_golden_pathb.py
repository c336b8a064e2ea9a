"""Golden fixtures for estimator B from the UNMODIFIED reference: WavefrontRenderer::Render
(src/renderer/wavefront.cc:40-242) over CPURayIntegrator, exactly what the reference's main() runs, with its
adaptive stop (kRelThresh 0.05, kMinSamples 16).  Run here, in the build container:

    python tests/golden/make_golden_pathb.py

Output (committed): pathb_<scene>.npz with, over RUNS independently seeded runs,
    rgb8_mean / rgb8_std   per-pixel mean and spread of the reference's 8-bit P3 output
    active_mean / active_std   pixels still sampling in pass s (the reference logs them as "queue=")
"""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))
import __graft_entry__ as g  # noqa: E402
from make_golden import CAMERAS, SCENES, digest  # noqa: E402
from oracle import ref  # noqa: E402

rt = g.load_package()
OUT = Path(__file__).resolve().parent
RUNS = 12
MAX_SAMPLES = {"cornell": 96, "rtiow": 64, "mixed": 64, "earth_bubbles": 64}


def main():
    for name, max_samples in MAX_SAMPLES.items():
        hs = SCENES[name]()
        prims, mats, texs, img = hs.export()
        rs = ref.RefScene(prims, mats, texs, img)
        cfg = rt.camera_config(**CAMERAS[name])
        imgs, act = [], []
        for run in range(RUNS):
            rgb8, active, secs = rs.render_b(cfg.as_oracle_cfg(), cfg.max_depth, max_samples, seed=100 + run)
            imgs.append(rgb8.astype(np.float64))
            act.append(active.astype(np.float64))
        imgs, act = np.stack(imgs), np.stack(act)
        np.savez_compressed(OUT / f"pathb_{name}.npz", rgb8_mean=imgs.mean(0).astype(np.float32), rgb8_std=imgs.std(0, ddof=1).astype(np.float32),
                            active_mean=act.mean(0), active_std=act.std(0, ddof=1), runs=RUNS, max_samples=max_samples,
                            max_depth=cfg.max_depth, scene=digest(prims))
        print(name, imgs.shape, "mean spp", act.mean(0).sum() / (imgs.shape[1] * imgs.shape[2]), "active", act.mean(0)[[0, 15, 16, 20, 32, max_samples - 1]], f"{secs:.1f}s/run")


if __name__ == "__main__":
    main()
