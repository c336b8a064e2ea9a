"""Golden vectors for the adaptive stop rule from the reference's OWN integrator::RecordSample / IsConverged
(src/integrator/pixel_state.h:22-72), compiled unmodified (oracle/_ref).  Run here, in the build container:

    python tests/golden/make_golden_pixel_state.py

Output (committed): pixel_state.npz - the generator parameters of 4096 sample sequences (regenerated from the seed by
`sequences()`, which the tests import) and, for four (threshold, min_spp) settings, the sample count at which the
reference stopped each of them plus its PixelState::sum and ::m2 there.
"""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
OUT = Path(__file__).resolve().parent
SETTINGS = [(0.05, 16), (0.02, 8), (0.1, 32), (0.002, 16)]  # the last one: the "final render" thresholds sampler.h suggests
N_SEQ, LEN, SEED = 4096, 128, 20261104


def sequences(n_seq=N_SEQ, length=LEN, seed=SEED):
    """Sample-radiance sequences shaped like a renderer's: a per-pixel mean colour (some channels near zero), noise from
    almost none to heavy-tailed, occasional exact zeros (absorbed paths) and fireflies.  float32 values, so that the
    device (float) and the reference (double) fold EXACTLY the same numbers."""
    rng = np.random.default_rng(seed)
    mean = rng.uniform(0.0, 1.5, (n_seq, 1, 3)) * (rng.uniform(size=(n_seq, 1, 3)) > 0.15)
    rel = 10.0 ** rng.uniform(-3.0, 0.3, (n_seq, 1, 1))
    x = mean * (1.0 + rel * rng.standard_normal((n_seq, length, 3)))
    x = np.abs(x)
    x *= rng.uniform(size=(n_seq, length, 1)) > rng.uniform(0.0, 0.3, (n_seq, 1, 1))      # absorbed paths
    fire = rng.uniform(size=(n_seq, length, 1)) < 0.003
    x = np.where(fire, x * 40.0 + 5.0, x)
    return x.astype(np.float32)


def main():
    from oracle import ref
    x = sequences()
    out = {"n_seq": N_SEQ, "len": LEN, "seed": SEED, "settings": np.array(SETTINGS), "checksum": float(x.astype(np.float64).sum())}
    for i, (thr, mn) in enumerate(SETTINGS):
        stop, s, m2 = ref.pixel_state_run(x.astype(np.float64), thr, mn)
        out[f"stop_{i}"], out[f"sum_{i}"], out[f"m2_{i}"] = stop, s, m2
        print(f"thr {thr} min {mn}: stopped before {LEN}: {(stop < LEN).mean():.3f}, at min: {(stop == mn).mean():.3f}, mean n {stop.mean():.1f}")
    np.savez_compressed(OUT / "pixel_state.npz", **out)


if __name__ == "__main__":
    main()
