"""First-contact GPU script: peaks, hit parity on three scenes, one render."""
import sys, time, json
sys.path.insert(0, '.')
import numpy as np
import __graft_entry__ as g
m = g.load_package()
from oracle import checker
print('peaks', m.measure_peaks(), flush=True)

def rays_for(name, n, rng):
    r = np.zeros(n, m.RAY_DTYPE)
    if name == 'cornell':
        r['o'] = np.float32([5, 5, -3.6]) + rng.uniform(-1, 1, (n, 3)).astype(np.float32)
        r['d'] = np.float32([0, 0, 1]) + rng.uniform(-0.6, 0.6, (n, 3)).astype(np.float32)
    else:
        r['o'] = rng.normal(size=(n, 3)).astype(np.float32) * 3 + np.float32([0, 2, 8])
        r['d'] = -r['o'] + rng.normal(size=(n, 3)).astype(np.float32) * 2
    r['tmin'] = 0.001; r['tmax'] = np.inf
    return r

import torch
for name, params in [('cornell', ()), ('rtiow', (11, 1)), ('bunny', (50,))]:
    hs = m.HostScene.build(name, params)
    ds = m.DeviceScene(hs)
    rng = np.random.default_rng(1)
    rays = rays_for(name, 1 << 20, rng)
    hits = ds.intersect(rays)
    ora = checker.make_oracle(hs)
    t0 = time.time(); t_ref, gid = ora.intersect(rays); cpu_s = time.time() - t0
    bad = np.nonzero(hits['prim_id'] != gid)[0]
    hit = gid >= 0
    rel = np.abs(hits['t'][hit] - t_ref[hit]) / np.abs(t_ref[hit])
    print(name, 'n', len(rays), 'hit frac %.3f' % hit.mean(), 'id mismatches', len(bad), 'max rel t err', rel.max() if rel.size else 0,
          't bit-equal', np.array_equal(hits['t'][hit], t_ref[hit]), 'cpu Mrays/s %.2f' % (len(rays) / cpu_s / 1e6), flush=True)
    for b in bad[:5]:
        print('   ray', b, rays[b], 'gpu', hits[b], 'ref', t_ref[b], gid[b])
    # device-resident timing
    d_rays = torch.from_numpy(rays.view(np.float32).reshape(-1, 8)).cuda()
    d_hits = torch.zeros(len(rays) * 4, dtype=torch.int32, device='cuda')
    for rep in range(3):
        torch.cuda.synchronize(); t0 = time.time()
        ds.intersect_device(d_rays.data_ptr(), len(rays), d_hits.data_ptr()); 
        m.load().rtb_shutdown if False else None
        import ctypes
        torch.cuda.synchronize()
        # library stream is non-blocking: sync it through a host-buffer call of zero cost
        ds.intersect(rays[:1])
        dt = time.time() - t0
    print('   gpu intersect_batch %.2f Mrays/s (wall, incl. 1 tiny sync call)' % (len(rays) / dt / 1e6), flush=True)

hs = m.HostScene.build('rtiow', (11, 1)); ds = m.DeviceScene(hs)
cfg = m.camera_preset('assets/cameras_readme.json', 'default'); cfg.image_width = 400
cam = m.camera_init(cfg)
for rep in range(2):
    img, st = ds.render(cam, spp=16, max_depth=10)
print('render rtiow', cam.width, cam.height, st, 'Msamples/s %.1f' % (st['samples'] / st['ms_total'] / 1e3), 'mean', (img / 16).mean(axis=(0, 1)))
rgb8 = m.resolve(img, 16); m.write_ppm('gpurun_out/rtiow_first.ppm', rgb8)
hs = m.HostScene.build('bunny', (50,)); ds = m.DeviceScene(hs)
cfg = m.camera_preset('assets/cameras_readme.json', 'wide'); cam = m.camera_init(cfg)
for rep in range(2):
    img, st = ds.render(cam, spp=16, max_depth=20)
print('render bunny', cam.width, cam.height, st, 'Msamples/s %.1f' % (st['samples'] / st['ms_total'] / 1e3), 'Mrays/s %.1f' % (st['rays'] / st['ms_total'] / 1e3))
m.write_ppm('gpurun_out/bunny_first.ppm', m.resolve(img, 16))
