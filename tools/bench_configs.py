#!/usr/bin/env python
"""The other BASELINE.json configs through Render() (full wavefront frame), reported the way the reference UI does:
rays / trace time per wave class from CoreStats (apps/imguiapp/main.cpp:230-233).  One JSON line per config.
  C1  1K-triangle mesh, 640x360
  C3  10M-triangle soup, 1080p, primary + 4 diffuse bounces (+ NEE shadow rays)
  C4  1000 instances x 100K-triangle BLAS, 1080p x 2 spp, per frame: GPU refit of the BLAS + TLAS rebuild + primary/shadow
Usage: python tools/bench_configs.py [c1] [c3] [c3gpu] [c4] [refit] [--frames N]   (c3gpu: the same with Setting gpuBuild = 1)"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from lighthouse2_b200 import core, scenes  # noqa: E402


def stats_line(name, rc, frames, extra):
    # per-class rates: waves one after the other (overlapConnect = 0, one chain), so that each event pair times one wave alone;
    # frame time: the default schedule (connect wave of bounce k beside the extend wave of bounce k+1)
    acc = {}
    rc.Setting("overlapConnect", 0)
    rc.Setting("chains", 1)  # one chain, so that no wave of another chain runs inside a wave's event span
    for _ in range(2):
        rc.Render(extra["view"], 1)
    for _ in range(frames):
        rc.Render(extra["view"], 1)
        st = rc.GetCoreStats()
        for k, v in st.items():
            if isinstance(v, (int, float)):
                acc[k] = acc.get(k, 0) + v
    serial_render_ms = acc["renderTime"] / frames
    rc.Setting("overlapConnect", 1)
    rc.Setting("chains", 0)
    for _ in range(2):
        rc.Render(extra["view"], 1)
    t0 = time.perf_counter()
    overlapped = 0.0
    for _ in range(frames):
        rc.Render(extra["view"], 1)
        overlapped += rc.GetCoreStats()["renderTime"]
    wall = (time.perf_counter() - t0) / frames * 1e3
    acc["renderTime"] = overlapped
    extra = dict(extra, render_ms_waves_serial=serial_render_ms)
    m = lambda n, t: (acc[n] / (acc[t] * 1e-3) / 1e6) if acc.get(t, 0) > 0 else None
    line = {"config": name, "frames": frames, "ms_per_frame_wall": wall, "render_ms": acc["renderTime"] / frames,
            "rays_per_frame": acc["totalRays"] / frames,
            "mrays_s": {"primary": m("primaryRayCount", "traceTime0"), "bounce1": m("bounce1RayCount", "traceTime1"),
                        "deep": m("deepRayCount", "traceTimeX"), "shadow": m("totalShadowRays", "shadowTraceTime"),
                        "frame_total": acc["totalRays"] / (acc["renderTime"] * 1e-3) / 1e6},
            "shade_ms": acc["shadeTime"] / frames}
    line.update({k: v for k, v in extra.items() if k != "view"})
    print(json.dumps(line), flush=True)


def refit_traffic(n_tris, nodes):
    """Algorithmic bytes of one refit (refit.cu): per triangle 16 B record tag + 48 B vertices + 16 B shading record read, 48 B record +
    16 B shading record written; per node 80 B read + 80 B written + 32 B float box written (and read once by the parent)."""
    return 144 * n_tris + 224 * nodes


def refit_c2(frames):
    """GPU refit of the 1M-triangle C2 mesh (vertices already on the device): the HBM-bound size of the refit kernel."""
    import torch
    rc = core.RenderCoreB200(0)
    verts = scenes.sphere_field(1_000_000)
    rc.SetGeometry(0, verts); rc.SetInstance(0, 0); rc.SetInstance(1, -1); rc.UpdateToplevel()
    v4 = np.ones((verts.shape[0] * 3, 4), np.float32); v4[:, :3] = verts.reshape(-1, 3) * 1.001
    dv = torch.from_numpy(v4).cuda()
    torch.cuda.synchronize()
    ms = []
    for _ in range(frames + 2):
        rc.RefitDevice(0, dv.data_ptr(), v4.shape[0])
        ms.append(rc.GetCoreStats()["refitTime"])
    ms = float(np.mean(ms[2:]))
    nodes = rc.GetCoreStats()["wideNodeCount"]
    b = refit_traffic(verts.shape[0], nodes)
    print(json.dumps({"config": "refit of the C2 mesh (1M triangles), device-resident vertices", "refit_ms_device": ms, "refit_bytes": b,
                      "refit_gbs": b / (ms * 1e-3) / 1e9, "hbm_frac_of_6543": b / (ms * 1e-3) / 1e9 / 6543.1, "nodes": nodes}), flush=True)
    rc.Shutdown()


def c1(frames):
    rc = core.RenderCoreB200(0)
    rc.SetGeometry(0, scenes.grid_mesh()); rc.SetInstance(0, 0); rc.SetInstance(1, -1); rc.UpdateToplevel()
    rc.SetTarget(640, 360, 1); rc.SetPointLight((0.5, 3.0, -1.0), (40, 36, 30)); rc.SetSky((0.3, 0.4, 0.6)); rc.Setting("maxPathLength", 1)
    view = scenes.view_pyramid(width=640, height=360, **scenes.C1_CAMERA)
    stats_line("C1 1K tris 640x360 primary+shadow", rc, frames, {"view": view})
    rc.Shutdown()


EXTRA_SETTINGS = [a.split("=") for a in sys.argv[1:] if "=" in a and not a.startswith("--")]  # name=value applied to the C3 core


def c3(frames, n_tris=10_000_000, gpu_build=0):
    rc = core.RenderCoreB200(0)
    rc.Setting("gpuBuild", gpu_build)
    for k, v in EXTRA_SETTINGS:
        rc.Setting(k, float(v))
    t0 = time.perf_counter()
    verts = scenes.triangle_soup(n_tris)
    t1 = time.perf_counter()
    rc.SetGeometry(0, verts); rc.SetInstance(0, 0); rc.SetInstance(1, -1); rc.UpdateToplevel()
    t2 = time.perf_counter()
    st = rc.GetCoreStats()
    rc.SetTarget(1920, 1080, 1); rc.SetPointLight((0.0, 120.0, -60.0), (2e4, 2e4, 1.8e4)); rc.SetSky((0.4, 0.5, 0.7)); rc.Setting("maxPathLength", 5); rc.Setting("russianRoulette", 0)
    view = scenes.view_pyramid(width=1920, height=1080, **scenes.C3_CAMERA)
    stats_line(f"C3 {n_tris} tri soup 1080p 4 bounces" + (" (BVH built on the GPU)" if gpu_build else ""), rc, frames,
               {"view": view, "scene_gen_s": t1 - t0, "bvh_build_upload_s": t2 - t1, "wide_nodes": st["wideNodeCount"],
                "scene_mb": (st["wideNodeCount"] * 80 + st["triangleCount"] * 48) / 1e6})
    rc.Shutdown()


def c4(frames, n_inst=1000, n_tris=100_000):
    rc = core.RenderCoreB200(0)
    for k, v in EXTRA_SETTINGS:
        rc.Setting(k, float(v))
    rc.SetGeometry(0, scenes.deforming_sphere(n_tris, time=0.0))
    M = scenes.lattice_transforms(n_inst, spacing=8.0, time=0.0)
    for i in range(n_inst):
        rc.SetInstance(i, 0, M[i])
    rc.SetInstance(n_inst, -1); rc.UpdateToplevel()
    rc.SetTarget(1920, 1080, 2); rc.SetPointLight((0.0, 80.0, -40.0), (1e4, 1e4, 9e3)); rc.SetSky((0.4, 0.5, 0.7)); rc.Setting("maxPathLength", 1)
    view = scenes.view_pyramid(width=1920, height=1080, **scenes.C4_CAMERA)
    refit_ms = tlas_ms = host_ms = 0.0
    acc = {}
    nodes = rc.GetCoreStats()["wideNodeCount"]
    for f in range(frames + 2):
        tm = 0.1 * (f + 1)
        v = scenes.deforming_sphere(n_tris, time=tm)           # host animation (outside the timed parts)
        M = scenes.lattice_transforms(n_inst, spacing=8.0, time=tm)
        t0 = time.perf_counter()
        rc.SetGeometry(0, v)                                   # same triangle count -> GPU refit (incl. the H2D of 4.8 MB)
        for i in range(n_inst):
            rc.SetInstance(i, 0, M[i])
        rc.SetInstance(n_inst, -1); rc.UpdateToplevel()
        t1 = time.perf_counter()
        rc.Render(view, 1)
        st = rc.GetCoreStats()
        if f >= 2:
            refit_ms += st["refitTime"]; tlas_ms += st["tlasTime"]; host_ms += (t1 - t0) * 1e3
            for k, v_ in st.items():
                if isinstance(v_, (int, float)):
                    acc[k] = acc.get(k, 0) + v_
    m = lambda n, t: (acc[n] / (acc[t] * 1e-3) / 1e6) if acc.get(t, 0) > 0 else None
    refit_bytes = refit_traffic(n_tris, nodes)
    print(json.dumps({"config": f"C4 {n_inst} inst x {n_tris} tri BLAS 1080p 2spp, refit + TLAS rebuild per frame", "frames": frames,
                      "refit_ms_device": refit_ms / frames, "refit_gbs": refit_bytes / (refit_ms / frames * 1e-3) / 1e9, "refit_bytes": refit_bytes,
                      "tlas_rebuild_ms_host": tlas_ms / frames, "scene_update_ms_wall": host_ms / frames, "render_ms": acc["renderTime"] / frames,
                      "rays_per_frame": acc["totalRays"] / frames,
                      "mrays_s": {"primary": m("primaryRayCount", "traceTime0"), "shadow": m("totalShadowRays", "shadowTraceTime"),
                                  "frame_total": acc["totalRays"] / (acc["renderTime"] * 1e-3) / 1e6},
                      "blas_nodes": nodes, "tlas_nodes": st["tlasNodeCount"]}), flush=True)
    rc.Shutdown()


if __name__ == "__main__":
    args = [a for a in sys.argv[1:] if not a.startswith("--") and not a.isdigit() and "=" not in a] or ["c1", "c3", "c4"]
    frames = int(sys.argv[sys.argv.index("--frames") + 1]) if "--frames" in sys.argv else 5
    for a in args:
        {"c1": c1, "c3": c3, "c3gpu": lambda f: c3(f, gpu_build=1), "c4": c4, "refit": refit_c2}[a](frames)
