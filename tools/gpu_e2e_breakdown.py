"""Where the end-to-end step (rtb_scene_create + rtb_render to a host buffer + rtb_scene_destroy) spends its wall time."""
import sys, time, ctypes as C
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
import bench, __graft_entry__ as g
m = g.load_package()
lib = m.load()
name = sys.argv[1] if len(sys.argv) > 1 else "bunny"
cfg = bench.CONFIGS[name]
hs = m.HostScene.build(*cfg["scene"])
cam = m.camera_init(bench.camera_for(m, cfg))
import torch
host_img = torch.empty(cam.width * cam.height * 3, dtype=torch.float32).pin_memory()
pageable = np.zeros(cam.width * cam.height * 3, np.float32)
for rep in range(6):
    t0 = time.perf_counter()
    ds = m.DeviceScene(hs)
    t1 = time.perf_counter()
    prm = m.binding.RtbRenderParams(0, cfg["spp"], cfg["depth"], m.MODE_A, 0x3360, 0, 0)
    st = m.binding.RtbStats()
    buf = host_img.data_ptr() if rep % 2 == 0 else pageable.ctypes.data
    m.binding._check(lib, lib.rtb_render(ds._h, C.byref(cam), C.byref(prm), buf, C.byref(st)))
    t2 = time.perf_counter()
    ds.close()
    t3 = time.perf_counter()
    print(f"{name} rep {rep} ({'pinned' if rep % 2 == 0 else 'pageable'} image): create {1e3*(t1-t0):.2f} ms  render {1e3*(t2-t1):.2f} ms (kernel {st.ms_total:.2f})  destroy {1e3*(t3-t2):.2f} ms  total {1e3*(t3-t0):.2f}")
