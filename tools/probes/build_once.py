#!/usr/bin/env python
"""One SetGeometry of the C2 mesh with the default (device, binned SAH) builder and two refits: the workload of the ncu capture
profiles/r02_ncu_build.txt (tools/profile_r02.sh)."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np  # noqa: E402
from lighthouse2_b200 import core, scenes  # noqa: E402

verts = scenes.sphere_field(int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000)
v4 = np.ones((verts.shape[0] * 3, 4), np.float32)
v4[:, :3] = verts.reshape(-1, 3)
rc = core.RenderCoreB200(0)
t0 = time.perf_counter()
rc.SetGeometry(0, v4)
print("build: %.1f ms, %d wide nodes" % ((time.perf_counter() - t0) * 1e3, rc.GetCoreStats()["wideNodeCount"]))
for _ in range(2):
    rc.SetGeometry(0, v4)  # same triangle count: refit
print("refit: %.3f ms" % rc.GetCoreStats()["refitTime"])
rc.Shutdown()
