"""Pick the strongest oracle available: the compiled reference (oracle/_ref) when it was built,
else the C restatement (oracle/port.c).  TEST INFRASTRUCTURE ONLY."""
from __future__ import annotations

from . import ref


class _RefOracle:
    kind = "reference"

    def __init__(self, host_scene, fast=False):
        prims, mats, texs, img = host_scene.export()
        self.scene = ref.RefScene(prims, mats, texs, img, fast=fast)

    def intersect(self, rays, threads=0):
        return self.scene.intersect(rays, threads)

    def render_a(self, cfg16, spp, max_depth, **kw):
        return self.scene.render_a(cfg16, spp, max_depth, **kw)


def make_oracle(host_scene, prefer="reference", fast=False):
    if prefer == "reference" and ref.available(fast):
        return _RefOracle(host_scene, fast)
    from . import port  # noqa: WPS433

    return port.PortOracle(host_scene)
