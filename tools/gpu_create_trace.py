"""rtb_scene_create of the stress scene (device-built tree), three times, with the library's own lap timer (RTB_TRACE_CREATE)
and the wall clock: where the time of the C5 end-to-end path goes before the first ray."""
import os, sys, time
from pathlib import Path
os.environ["RTB_TRACE_CREATE"] = "1"
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import bench, __graft_entry__ as g
m = g.load_package()
cfg = bench.CONFIGS["stress"]
t = time.perf_counter(); hs = m.HostScene.build(*cfg["scene"], device_tree=True); print("host scene (no tree) %.2f s" % (time.perf_counter() - t), flush=True)
for rep in range(3):
    t = time.perf_counter(); ds = m.DeviceScene(hs); dt = time.perf_counter() - t
    print("rtb_scene_create #%d: %.1f ms" % (rep, 1e3 * dt), flush=True)
    t = time.perf_counter(); ds.close(); print("rtb_scene_destroy: %.1f ms" % (1e3 * (time.perf_counter() - t)), flush=True)
os.environ["RTB_HOST_SHADOWS"] = "1"
for rep in range(2):
    t = time.perf_counter(); ds = m.DeviceScene(hs); dt = time.perf_counter() - t
    print("with host-side float shadows #%d: %.1f ms" % (rep, 1e3 * dt), flush=True)
    ds.close()
