#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
export RTB_LIB=librtb200_tailmerge.so
timeout 60 python -c "
import __graft_entry__ as g, numpy as np
m = g.load_package()
ds = m.DeviceScene(m.HostScene.build('cornell'))
cam = m.camera_init(m.camera_config(aspect_ratio=1.5, image_width=60, vfov=40.0, lookfrom=(5, 5, -3.6), lookat=(5, 5, 5), focus_dist=9.0))
for spp in (1, 4, 32):
    img, st = ds.render(cam, spp=spp, max_depth=10)
    print('tiny', spp, st['rays'], st['ms_total'], float(img.sum()), flush=True)
" > gpurun_out/c11_tiny.txt 2>&1
echo "tiny rc $?" >> gpurun_out/c11_tiny.txt
cat gpurun_out/c11_tiny.txt
if grep -q "tiny rc 0" gpurun_out/c11_tiny.txt; then
  timeout 240 python tools/gpu_tail_ab.py bunny rtiow cornell earth stress > gpurun_out/c11_tail_merge.txt 2>&1; echo "ab rc $?" >> gpurun_out/c11_tail_merge.txt
  cat gpurun_out/c11_tail_merge.txt
fi
