"""Debug helper: per-iteration timing and traversal counters of one bunny render (counters build)."""
import os, sys
sys.path.insert(0, '.')
os.environ['RTB_DEBUG_COUNTERS'] = '1'
os.environ['RTB_TRACE_ITERS'] = '1'
import __graft_entry__ as g
m = g.load_package()
name = sys.argv[1] if len(sys.argv) > 1 else 'bunny'
counters = (sys.argv[2] != '0') if len(sys.argv) > 2 else True
import bench
cfg = bench.CONFIGS[name]
hs = m.HostScene.build(*cfg['scene'], counters=counters)
ds = m.DeviceScene(hs)
cam = m.camera_init(bench.camera_for(m, cfg))
for rep in range(2):
    img, st = ds.render(cam, spp=cfg['spp'], max_depth=cfg['depth'], flags=int(os.environ.get('RTB_FLAGS', '4')))
    print(st, flush=True)
