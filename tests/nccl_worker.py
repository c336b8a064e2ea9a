"""torchrun worker of tests/test_nccl_gpu.py: the torch.distributed (NCCL) sharding of multigpu.py on real GPUs.
Rank 0 writes what the ranks produced together to the path in argv[1]."""
import os
import sys
from pathlib import Path

import numpy as np
import torch
import torch.distributed as dist

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
import __graft_entry__ as g  # noqa: E402

rt = g.load_package()
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
hs = rt.HostScene.build("rtiow", (11, 1))
ds = rt.DeviceScene(hs, device=local)
cam = rt.camera_init(rt.camera_config(aspect_ratio=1.777, image_width=192, vfov=45.0, lookfrom=(0, 2, 5), lookat=(0, 0, 0)))
n = cam.width * cam.height
spp, depth = 40, 10
accum = torch.zeros(n * 3, dtype=torch.float32, device="cuda")
st = rt.multigpu.render_sharded(ds, cam, spp, depth, accum, dist)
rays = torch.tensor([st["rays"]], dtype=torch.int64, device="cuda")
dist.all_reduce(rays)
# adaptive: every rank applies the rule to its own slice; sums, sums of squares and counts are reduced
a_sum = torch.zeros(n * 3, dtype=torch.float32, device="cuda")
a_sq = torch.zeros(n * 3, dtype=torch.float32, device="cuda")
a_cnt = torch.zeros(n, dtype=torch.int32, device="cuda")
rt.multigpu.render_sharded_adaptive(ds, cam, 64 * world, depth, a_sum, a_sq, a_cnt, dist, rel_threshold=0.05, min_samples=16, mode=rt.MODE_A)
if rank == 0:
    np.savez(sys.argv[1], fixed=accum.cpu().numpy(), rays=int(rays.item()), a_sum=a_sum.cpu().numpy(), a_cnt=a_cnt.cpu().numpy(),
             shape=np.array([cam.height, cam.width]), world=world, spp=spp)
dist.barrier()
dist.destroy_process_group()
