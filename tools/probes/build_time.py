#!/usr/bin/env python
"""SetGeometry wall time of the default builder, best of 5 (B200_LIB picks a build variant).  Usage: build_time.py [tris]"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np  # noqa: E402
from lighthouse2_b200 import core, scenes  # noqa: E402

for gen in (scenes.sphere_field, scenes.triangle_soup):
    verts = gen(int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000)
    v4 = np.ones((verts.shape[0] * 3, 4), np.float32)
    v4[:, :3] = verts.reshape(-1, 3)
    best = 1e9
    for _ in range(5):
        rc = core.RenderCoreB200(0)
        t0 = time.perf_counter()
        rc.SetGeometry(0, v4)
        best = min(best, time.perf_counter() - t0)
        n = rc.GetCoreStats()["wideNodeCount"]
        rc.Shutdown()
    # a scene that rebuilds: the same core, the triangle count changes from call to call (so SetGeometry cannot refit)
    rc = core.RenderCoreB200(0)
    rc.SetGeometry(0, v4)
    again = 1e9
    for k in range(6):
        t0 = time.perf_counter()
        rc.SetGeometry(0, v4[:-3] if k % 2 == 0 else v4)
        again = min(again, time.perf_counter() - t0)
    rc.Shutdown()
    print("%s %d triangles: fresh core %.2f ms, rebuild in the same core %.2f ms, %d wide nodes" % (gen.__name__, verts.shape[0], best * 1e3, again * 1e3, n))
