#!/usr/bin/env python
"""Filter mode (Setting filterCore) on the C2 scene: device time of the reconstruction filter + TAA and of the whole frame, 1080p and 4K, TAA on / off.
Usage: python tools/bench_filter.py"""
import sys, json, numpy as np
sys.path.insert(0, __import__('os').path.dirname(__import__('os').path.dirname(__import__('os').path.abspath(__file__))))
from lighthouse2_b200 import core, scenes
for (w, h, spp) in ((1920, 1080, 1), (3840, 2160, 1)):
    verts = scenes.sphere_field(1_000_000)
    v4 = np.ones((verts.shape[0] * 3, 4), np.float32); v4[:, :3] = verts.reshape(-1, 3)
    rc = core.RenderCoreB200(0)
    rc.SetGeometry(0, v4); rc.SetInstance(0, 0); rc.SetInstance(1, -1); rc.UpdateToplevel()
    rc.SetTarget(w, h, spp)
    rc.SetPointLight((0.0, 50.0, 0.0), (3000.0, 3000.0, 3000.0))
    rc.Setting("filterCore", 1)
    view = scenes.view_pyramid(width=w, height=h, **scenes.C2_CAMERA)
    for taa in (1, 0):
        rc.Setting("TAA", taa)
        ts, rs = [], []
        for k in range(6):
            rc.Render(view, converge=0)
            st = rc.GetCoreStats(); ts.append(st["filterTime"]); rs.append(st["renderTime"])
        print(json.dumps({"config": "C2 scene %dx%d %d spp, filterCore" % (w, h, spp), "TAA": taa, "filter_ms": round(float(np.median(ts[1:])), 4), "render_ms": round(float(np.median(rs[1:])), 4),
                          "pixels": w * h}), flush=True)
    rc.Shutdown()
