#!/usr/bin/env python
"""Extend-wave time vs ray count on the diffuse-bounce rays of the C2 pass: T(n) = fixed cost + n / marginal rate.  GPU box only."""
import os, sys, json
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
from lighthouse2_b200 import core, scenes, workload

rc = core.RenderCoreB200(0)
verts = scenes.sphere_field(1_000_000)
rc.SetGeometry(0, verts); rc.SetInstance(0, 0); rc.SetInstance(1, -1); rc.UpdateToplevel()
view = scenes.view_pyramid(width=1920, height=1080, **scenes.C2_CAMERA)
O, D = scenes.primary_rays(view, 1920, 1080)
def tr(o, d):
    h = core.unpack_hits(rc.TraceClosest(*core.pack_rays(o, d, tmin=workload.EPS))); return h["t"], h["prim"]
waves = workload.make_pass(tr, verts, O, D, scenes.C2_LIGHT, 2, seed=1)
dev = torch.device("cuda", 0)
rc.SetStream(torch.cuda.current_stream().cuda_stream)
for w in (waves[0], waves[2]):
    n_all = w["O4"].shape[0]
    perm = np.random.default_rng(0).permutation(n_all) if "--shuffle" in sys.argv else np.arange(n_all)
    dO, dD = torch.from_numpy(w["O4"][perm]).to(dev), torch.from_numpy(w["D4"][perm]).to(dev)
    out = torch.empty((n_all, 4), dtype=torch.float32, device=dev)
    res = []
    for n in (32, 1024, 10_000, 50_000, 113_664, 227_328, 500_000, 1_000_000, n_all):
        n = min(n, n_all)
        for _ in range(3): rc.TraceClosestDevice(dO.data_ptr(), dD.data_ptr(), n, out.data_ptr())
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(10): rc.TraceClosestDevice(dO.data_ptr(), dD.data_ptr(), n, out.data_ptr())
        b.record(); torch.cuda.synchronize()
        ms = a.elapsed_time(b) / 10
        res.append((n, round(ms * 1e3, 1), round(n / ms / 1e3)))
    print(w["cls"], "(rays, us, Mrays/s):", res, flush=True)
