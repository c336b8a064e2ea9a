"""Multi-GPU host logic on CPU: spp sharding tiles the sample range, and the N-rank reduce of the
per-rank accumulators equals the single-rank sum (gloo, world_size 2 and 3)."""
import os
import socket

import numpy as np
import pytest


def test_spp_slices_tile_the_range(rt):
    for spp in (0, 1, 7, 100, 1024):
        for world in (1, 2, 3, 8):
            got = [rt.multigpu.spp_slice(spp, r, world) for r in range(world)]
            assert got[0][0] == 0 and got[-1][1] == spp
            assert all(a[1] == b[0] for a, b in zip(got, got[1:]))
            sizes = [e - b for b, e in got]
            assert max(sizes) - min(sizes) <= 1
    assert rt.multigpu.weak_slice(100, 3) == (300, 400)
    with pytest.raises(ValueError):
        rt.multigpu.spp_slice(10, 2, 2)


def _fake_radiance(n_pix, b, e):
    """Deterministic stand-in for a rank's render: sum over its global sample indices of f(pixel, sample)."""
    s = np.arange(b, e, dtype=np.float64)[None, :]
    p = np.arange(n_pix * 3, dtype=np.float64)[:, None]
    return np.sin(0.37 * p + 1.3 * s).sum(axis=1).astype(np.float32) if e > b else np.zeros(n_pix * 3, np.float32)


def _worker(rank, world, port, spp, n_pix, out_path):
    import sys
    from pathlib import Path
    root = Path(__file__).resolve().parents[1]
    sys.path.insert(0, str(root))
    import torch
    import torch.distributed as dist
    import __graft_entry__ as g
    rt = g.load_package()
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    b, e = rt.multigpu.spp_slice(spp, rank, world)
    accum = torch.from_numpy(_fake_radiance(n_pix, b, e))
    rt.multigpu.reduce_accumulator(accum, dist)
    if rank == 0:
        np.save(out_path, accum.numpy())
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_reduce_of_sharded_accumulators_equals_single_rank(rt, tmp_path, world):
    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    spp, n_pix = 37, 101
    out = tmp_path / "sum.npy"
    mp.spawn(_worker, args=(world, port, spp, n_pix, str(out)), nprocs=world, join=True)
    got = np.load(out)
    want = _fake_radiance(n_pix, 0, spp)
    assert np.allclose(got, want, rtol=1e-5, atol=1e-4)


def _worker_adaptive(rank, world, port, n_pix, out_path):
    import sys
    from pathlib import Path
    root = Path(__file__).resolve().parents[1]
    sys.path.insert(0, str(root))
    import torch
    import torch.distributed as dist
    import __graft_entry__ as g
    rt = g.load_package()
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    # stand-in for a rank's adaptive render: pixel p stopped after (p + rank) % 5 + 1 samples of value (rank + 1)
    cnt = torch.tensor([(p + rank) % 5 + 1 for p in range(n_pix)], dtype=torch.int32)
    accum = (cnt.to(torch.float32) * (rank + 1)).repeat_interleave(3)
    rt.multigpu.reduce_adaptive(accum, cnt, dist)
    if rank == 0:
        np.savez(out_path, sum=accum.numpy(), count=cnt.numpy())
    dist.barrier()
    dist.destroy_process_group()


def test_adaptive_reduce_sums_radiance_and_counts(rt, tmp_path):
    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    world, n_pix = 2, 64
    out = tmp_path / "adaptive.npz"
    mp.spawn(_worker_adaptive, args=(world, port, n_pix, str(out)), nprocs=world, join=True)
    got = np.load(out)
    cnts = np.array([[(p + r) % 5 + 1 for p in range(n_pix)] for r in range(world)])
    assert np.array_equal(got["count"], cnts.sum(0))
    want = (cnts * np.arange(1, world + 1)[:, None]).sum(0).repeat(3).astype(np.float32)
    assert np.allclose(got["sum"], want)
    # the image is sum / count: a count-weighted mean of the per-rank means
    mean = got["sum"].reshape(-1, 3)[:, 0] / got["count"]
    assert np.all((mean >= 1.0) & (mean <= world))


def test_bench_reference_arm_line_has_the_contract_keys_and_the_common_config():
    """bench.py --impl reference on the host cores: one JSON line with the contract's keys; `config` is the object the b200 arm prints for
    the same workload (bench.common_config), so that the driver can see both arms measure one thing."""
    import json
    import subprocess
    import sys
    from pathlib import Path
    root = Path(__file__).resolve().parents[1]
    from oracle import ref
    if not ref.available(fast=True) and not ref.available():
        pytest.skip("oracle/_ref not built")
    r = subprocess.run([sys.executable, str(root / "bench.py"), "--impl", "reference", "--config", "rtiow", "--steps", "1", "--warmup", "0",
                        "--cpu-step-budget", "1"], capture_output=True, text=True, timeout=600, cwd=root)
    assert r.returncode == 0, r.stderr[-500:]
    line = json.loads([ln for ln in r.stdout.split("\n") if ln.startswith("{")][-1])
    for k in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline", "dtype", "data", "config",
              "impl", "cpu_baseline", "e2e"):
        assert k in line, k
    assert line["impl"] == "reference" and line["unit"] == "Msamples/s" and line["scaling"] == "strong" and line["value"] > 0
    assert line["e2e"] == {"value": line["value"], "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert line["cpu_baseline"]["kind"] in ("reference", "port") and line["cpu_baseline"]["cores"] >= 1
    import bench
    assert line["config"] == bench.common_config(bench.CONFIGS["rtiow"], 1000, 562, 44, 10)
