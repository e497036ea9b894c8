#!/usr/bin/env python
"""BVH build time, host SAH builder vs the GPU builder (Setting gpuBuild), as SetGeometry wall time (upload included) on the C2
sphere field and a C3-class soup.  Usage: python tools/bench_build.py [tris ...]"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
from lighthouse2_b200 import core, scenes  # noqa: E402

# --cache: also time SetGeometry when the host-built tree comes from the on-disk cache (B200_BVH_CACHE, bvh_cache.cpp)
use_cache = "--cache" in sys.argv
if use_cache:
    import tempfile
    os.environ["B200_BVH_CACHE"] = tempfile.mkdtemp(prefix="b200_bvh_cache_", dir=os.path.join(ROOT, "gpurun_out"))
sizes = [int(a) for a in sys.argv[1:] if a.isdigit()] or [1_000_000, 10_000_000]
for n in sizes:
    for name, gen in (("sphere field", scenes.sphere_field), ("soup", scenes.triangle_soup)):
        verts = gen(n)
        v4 = np.ones((verts.shape[0] * 3, 4), np.float32); v4[:, :3] = verts.reshape(-1, 3)
        row = {"scene": name, "triangles": int(verts.shape[0])}
        for gpu in (0, 1, 2, 3, 4):
            best, first, hits = 1e9, None, 0
            for rep in range(2 if gpu == 0 and n > 2_000_000 else 3):
                rc = core.RenderCoreB200(0)
                rc.Setting("gpuBuild", gpu)
                t0 = time.perf_counter()
                rc.SetGeometry(0, v4)
                dt = time.perf_counter() - t0
                first = dt if first is None else first
                best = min(best, dt)
                st = rc.GetCoreStats()
                nodes, hits = st["wideNodeCount"], hits + st["bvhCacheHits"]
                rc.Shutdown()
            if gpu == 0 and use_cache:
                row["host_build_first_s"], row["host_from_cache_s"], row["cache_hits"] = round(first, 4), round(best, 4), hits
            name = ("host", "gpu", "ploc", "sahb", "sah")[gpu]
            row[name + "_build_s"], row[name + "_nodes"] = round(best, 4), nodes
        print(json.dumps(row), flush=True)
