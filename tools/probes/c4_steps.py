#!/usr/bin/env python
"""Steps per ray on the C4 scene (1000 instances of a 100 K-triangle sphere): wide nodes visited and triangles tested per primary ray,
through the general two-level kernel (Setting countSteps)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np  # noqa: E402
from lighthouse2_b200 import core, scenes  # noqa: E402

n_inst, n_tris = 1000, 100_000
rc = core.RenderCoreB200(0)
rc.SetGeometry(0, scenes.deforming_sphere(n_tris, time=0.0))
M = scenes.lattice_transforms(n_inst, spacing=8.0, time=0.0)
for i in range(n_inst):
    rc.SetInstance(i, 0, M[i])
rc.SetInstance(n_inst, -1)
rc.UpdateToplevel()
view = scenes.view_pyramid(width=1920, height=1080, **scenes.C4_CAMERA)
O, D = scenes.primary_rays(view, 1920, 1080)
O4, D4 = core.pack_rays(O, D)
for rep in range(3):
    h = rc.TraceClosest(O4, D4)
print("plain wave: %.3f ms for %d rays = %.0f Mrays/s (host buffers; kernel time from the stats below)" % (0, len(O), 0))
print(rc.GetCoreStats()["lastTraceMs"] if "lastTraceMs" in rc.GetCoreStats() else "")
rc.Setting("countSteps", 1)
h = core.unpack_hits(rc.TraceClosest(O4, D4))
nn, nt = rc.StepCounters(len(O))
hit = h["prim"] >= 0
print("rays %d, hit %.1f %%; nodes per ray %.2f (hit %.2f, miss %.2f), triangles per ray %.2f (hit %.2f); nodes percentiles 50/90/99: %s" % (
    len(O), 100 * hit.mean(), nn.mean(), nn[hit].mean(), nn[~hit].mean(), nt.mean(), nt[hit].mean(), np.percentile(nn, [50, 90, 99]).tolist()))
st = rc.GetCoreStats()
print("BLAS nodes %d, TLAS nodes %d" % (st["wideNodeCount"], st["tlasNodeCount"]))
rc.Shutdown()
