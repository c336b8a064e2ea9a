"""How much of the traversal cost is incoherence?  Same ray set, three orders: tiled (coherent), random, all-identical."""
import sys, time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np, torch
import bench, __graft_entry__ as g
m = g.load_package()
name = sys.argv[1] if len(sys.argv) > 1 else "bunny"
cfg = bench.CONFIGS[name]
hs = m.HostScene.build(*cfg["scene"]); ds = m.DeviceScene(hs)
cam = m.camera_init(bench.camera_for(m, cfg))
W, H = cam.width, cam.height
rng = np.random.default_rng(3)
spp = 11
ys, xs = np.mgrid[0:H, 0:W]
# 8x4 tiles like the generate kernel
order = (ys // 4 * ((W + 7) // 8) + xs // 8) * 32 + (ys % 4) * 8 + xs % 8
idx = np.argsort(order.ravel(), kind="stable")
px = np.tile(xs.ravel()[idx], spp).astype(np.float32) + rng.uniform(-.5, .5, W * H * spp).astype(np.float32)
py = np.tile(ys.ravel()[idx], spp).astype(np.float32) + rng.uniform(-.5, .5, W * H * spp).astype(np.float32)
c = np.float32(cam.center[:]); p00 = np.float32(cam.pixel00[:]); du = np.float32(cam.delta_u[:]); dv = np.float32(cam.delta_v[:])
n = px.size
rays = np.zeros(n, m.RAY_DTYPE)
rays["o"] = c; rays["d"] = p00 + px[:, None] * du + py[:, None] * dv - c
rays["tmin"] = 0.001; rays["tmax"] = np.inf
# secondary rays: diffuse bounce off the primary hits
hits = ds.intersect(rays)
hit = hits["prim_id"] >= 0
P = rays["o"] + hits["t"][:, None].astype(np.float32) * rays["d"]
sec = rays[hit].copy(); sec["o"] = P[hit]
v = rng.normal(size=(hit.sum(), 3)).astype(np.float32); v /= np.linalg.norm(v, axis=1, keepdims=True)
v[:, 1] = np.abs(v[:, 1])  # roughly the upper hemisphere (ground + bunny top); good enough for a coherence probe
sec["d"] = v
def run(label, r):
    d_rays = torch.from_numpy(r.view(np.float32).reshape(-1, 8)).cuda()
    d_hits = torch.zeros(len(r) * 4, dtype=torch.int32, device="cuda")
    stream = torch.cuda.Stream(); st = stream.cuda_stream
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ts = []
    for rep in range(5):
        e0.record(stream); ds.intersect_device(d_rays.data_ptr(), len(r), d_hits.data_ptr(), st); e1.record(stream); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    t = min(ts[1:])
    print("%-28s %9d rays %.3f ms  %.3f ns/ray  %.2f Grays/s" % (label, len(r), t, 1e6 * t / len(r), len(r) / t / 1e6), flush=True)
run("primary tiled", rays)
run("primary random order", rays[rng.permutation(n)])
run("primary all identical", np.repeat(rays[n // 2 + 1234:n // 2 + 1235], n))
run("secondary (parent order)", sec)
run("secondary random order", sec[rng.permutation(len(sec))])
k = np.lexsort((sec["d"][:, 2] > 0, sec["d"][:, 0] > 0))
run("secondary sorted by octant", sec[k])
# The bound on what ANY regrouping of rays between lanes / warps could buy: the same ray population, but every warp's 32 lanes
# trace one and the same ray (chunks of 32 consecutive batch entries are what a warp claims), i.e. zero SIMT divergence and
# 32 identical addresses per load.  time(x32) / time(as is) is the floor a perfectly coherent scheduler could reach.
run("primary, each ray x32", np.repeat(rays[::32], 32)[:n])
run("secondary, each ray x32", np.repeat(sec[::32], 32)[:len(sec)])
mix = np.concatenate([rays[: len(rays) // 2], sec[: len(sec) // 2]])
mix = mix[rng.permutation(len(mix))]
run("primary+secondary shuffled", mix)
run("primary+secondary, x32", np.repeat(mix[::32], 32)[:len(mix)])
print("--- launch floor: small batches")
for k in (32, 1024, 32768):
    run("primary first %d" % k, rays[n // 2:n // 2 + k])
    run("secondary first %d" % k, sec[len(sec) // 2:len(sec) // 2 + k])
