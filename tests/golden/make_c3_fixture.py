"""Golden hit buffer for BASELINE.json configs[2] AT SIZE: the 10 000 000-triangle soup, built and traced by the REFERENCE ITSELF.

Run in the build container (needs /root/reference to compile oracle/_ref; the reference's BVH2::Rebuild of 10 M triangles takes a few
minutes and ~3 GB):
    python tests/golden/make_c3_fixture.py
Writes tests/golden/c3_10m.npz: 40 000 primary rays (sampled pixels of the 1080p C3 view) and the 4 diffuse-bounce waves that follow them
(rays stored explicitly: their directions come from numpy sin / cos, whose last bits may differ between machines), with (t, prim) of every
ray from oracle/_ref = the unmodified lib/RenderCore_Bart/bvh.cpp.  tests/test_fullsize_gpu.py::test_c3_10m_against_the_reference_fixture
regenerates the soup from its seed on the GPU box and compares the CUDA waves with it: that the reported triangle is the CLOSEST one of
ten million is exactly what the size-independent properties of test_c3_class_soup_properties cannot show.
"""
import hashlib
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import oracle  # noqa: E402
from oracle import Oracle  # noqa: E402
from lighthouse2_b200 import scenes, workload  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
N_TRIS, N_PIXELS, BOUNCES = 10_000_000, 40_000, 4


def main():
    oracle.build()
    kind = "ref" if oracle.have_ref() else None
    assert kind, "oracle/_ref/liboracle_ref.so missing: /root/reference is required to make goldens"
    verts = scenes.triangle_soup(N_TRIS)
    t0 = time.time()
    o = Oracle(kind)
    o.SetGeometry(0, verts); o.SetInstance(0, 0); o.UpdateToplevel()
    print("reference BVH build of %d triangles: %.0f s" % (N_TRIS, time.time() - t0), flush=True)
    view = scenes.view_pyramid(width=1920, height=1080, **scenes.C3_CAMERA)
    O, D = scenes.primary_rays(view, 1920, 1080)
    pix = np.sort(np.random.default_rng(2024).choice(len(O), N_PIXELS, replace=False))
    tr = lambda o_, d_: o.trace(o_, d_, nthreads=os.cpu_count() or 1)[:2]
    waves = workload.make_pass(tr, verts, O[pix], D[pix], (0.0, 120.0, -60.0), BOUNCES, seed=77)
    out = {}
    k = 0
    for w in waves:
        if w["kind"] != "closest":
            continue
        t, prim, _ = o.trace(np.ascontiguousarray(w["O4"][:, :3]), np.ascontiguousarray(w["D4"][:, :3]), nthreads=os.cpu_count() or 1)
        out["O%d" % k], out["D%d" % k] = np.ascontiguousarray(w["O4"][:, :3]), np.ascontiguousarray(w["D4"][:, :3])
        out["t%d" % k], out["prim%d" % k] = t, prim
        print("wave %d (%s %d): %d rays, %d hits" % (k, w["cls"], w["bounce"], len(t), int((prim >= 0).sum())), flush=True)
        k += 1
    sha = hashlib.sha256(b"".join(out[n].tobytes() for n in sorted(out))).hexdigest()
    path = os.path.join(HERE, "c3_10m.npz")
    np.savez_compressed(path, waves=np.int32(k), pixels=pix.astype(np.int32), sha256=np.frombuffer(bytes.fromhex(sha), np.uint8), **out)
    print("-> %s (%.0f KiB), sha256 %s" % (os.path.relpath(path, ROOT), os.path.getsize(path) / 1024, sha[:16]))
    o.close()


if __name__ == "__main__":
    main()
