"""Parity at BASELINE.json's full sizes (C2: 1M-triangle sphere field, 1920x1080; C3-class: 10M-triangle soup) through
sampled comparison with the CPU oracle and through size-independent properties: any-hit <=> closest t < tmax, permutation
invariance, idempotence, chunked host path == device path, refit with unchanged vertices == identity."""
import numpy as np
import pytest

import cases
import parity
from lighthouse2_b200 import core, scenes, workload

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def c2():
    verts = scenes.sphere_field(1_000_000)
    rc = core.RenderCoreB200(0)
    rc.SetGeometry(0, verts); rc.SetInstance(0, 0); rc.SetInstance(1, -1); rc.UpdateToplevel()
    view = scenes.view_pyramid(width=1920, height=1080, **scenes.C2_CAMERA)
    O, D = scenes.primary_rays(view, 1920, 1080)

    def tr(o, d):
        h = core.unpack_hits(rc.TraceClosest(*core.pack_rays(o, d, tmin=workload.EPS)))
        return h["t"], h["prim"]
    waves = workload.make_pass(tr, verts, O, D, scenes.C2_LIGHT, bounces=2, seed=3)
    yield dict(rc=rc, verts=verts, waves=waves)
    rc.Shutdown()


def test_c2_sampled_against_oracle(c2):
    """20 000 random rays of every wave of the full-size pass vs the reference traversal: ids and t bit-exact."""
    case = dict(meshes=[c2["verts"]], instances=[(0, None)])
    port = cases.load_oracle(case, "port")
    rng = np.random.default_rng(11)
    total = exceptions = 0
    for w in c2["waves"]:
        idx = np.sort(rng.choice(w["O4"].shape[0], 20000, replace=False))
        O4, D4 = np.ascontiguousarray(w["O4"][idx]), np.ascontiguousarray(w["D4"][idx])
        t, prim, _ = port.trace(O4[:, :3], D4[:, :3], nthreads=8)
        if w["kind"] == "closest":
            h = core.unpack_hits(c2["rc"].TraceClosest(O4, D4))
            bad = (h["prim"] != prim) | (h["t"].view(np.uint32) != t.view(np.uint32))
        else:
            bad = c2["rc"].TraceAny(O4, D4) != (t < D4[:, 3])
        total += len(idx); exceptions += int(bad.sum())
    print("C2 sampled parity:", exceptions, "exceptions in", total, "rays")
    assert exceptions <= max(1, int(1e-5 * total))


def test_c2_any_hit_consistent_with_closest(c2):
    rc = c2["rc"]
    w = c2["waves"][2]  # first diffuse wave, all rays
    h = core.unpack_hits(rc.TraceClosest(w["O4"], w["D4"]))
    rng = np.random.default_rng(5)
    D4 = w["D4"].copy()
    D4[:, 3] = rng.uniform(0.5, 60.0, len(D4)).astype(np.float32)
    occ = rc.TraceAny(w["O4"], D4)
    assert np.array_equal(occ, h["t"] < D4[:, 3])
    # shadow wave of the pass itself: occluded <=> a closest hit exists below tmax
    s = c2["waves"][1]
    hs = core.unpack_hits(rc.TraceClosest(s["O4"], np.concatenate([s["D4"][:, :3], np.full((len(s["D4"]), 1), 1e34, np.float32)], axis=1)))
    assert np.array_equal(rc.TraceAny(s["O4"], s["D4"]), hs["t"] < s["D4"][:, 3])


def test_c2_permutation_invariance_and_idempotence(c2):
    """A ray's hit must not depend on which rays share its warp (triangle tests are done by other lanes) nor on the run."""
    rc = c2["rc"]
    w = c2["waves"][2]
    a = rc.TraceClosest(w["O4"], w["D4"])
    b = rc.TraceClosest(w["O4"], w["D4"])
    assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
    perm = np.random.default_rng(9).permutation(len(a))
    c = rc.TraceClosest(np.ascontiguousarray(w["O4"][perm]), np.ascontiguousarray(w["D4"][perm]))
    assert np.array_equal(c.view(np.uint32), a.view(np.uint32)[perm])


def test_c2_exact_ties_do_not_depend_on_the_warp(c2):
    """Rays aimed exactly at mesh vertices and edge midpoints pass through shared edges: several triangles report bit-identical t.
    The kernel's tie rule (smallest leaf-order index, traverse.cu::intersectTriangle) must make the answer independent of which rays
    share a warp -- the property the bit-identical multi-GPU frame check of bench.py rests on at hundreds of millions of rays -- and the
    answer must be one of the tied triangles (checked with the reference arithmetic through the oracle's exhaustive search)."""
    rc, verts = c2["rc"], c2["verts"]
    rng = np.random.default_rng(17)
    tri = verts[rng.choice(len(verts), 400_000, replace=False)]
    targets = np.concatenate([tri[:200_000, 0], (0.5 * (tri[200_000:, 0] + tri[200_000:, 1])).astype(np.float32)])
    O = np.tile(np.asarray(scenes.C2_CAMERA["pos"], np.float32), (len(targets), 1)) + rng.normal(0, 0.05, (len(targets), 3)).astype(np.float32)
    D = scenes._normalize_ref((targets - O).astype(np.float32))
    O4, D4 = core.pack_rays(O, D, tmin=1e-4)
    a = rc.TraceClosest(O4, D4)
    for seed in (1, 2, 3):
        perm = np.random.default_rng(seed).permutation(len(a))
        b = rc.TraceClosest(np.ascontiguousarray(O4[perm]), np.ascontiguousarray(D4[perm]))
        assert np.array_equal(b.view(np.uint32), a.view(np.uint32)[perm]), "a hit depends on the rays it shares a warp with"
    # against the oracle on a sample: identical, or a documented tie (different triangle, bit-identical reference t)
    port = cases.load_oracle(dict(meshes=[verts], instances=[(0, None)]), "port")
    idx = np.sort(rng.choice(len(a), 40_000, replace=False))
    t, prim, inst = port.trace(O[idx], D[idx], nthreads=8)
    rep = parity.compare("c2-ties", a[idx], port, O[idx], D[idx], t, prim, inst)
    print("rays through vertices / edge midpoints:", rep)
    assert rep["wrong"] == 0, rep  # every difference is a documented class: an exact tie or a box the reference's slab test culled (tests/parity.py)


def test_c2_host_chunks_equal_device_wave(c2):
    import torch
    rc = c2["rc"]
    w = c2["waves"][0]  # 2 073 600 primary rays: the host path splits them into pipelined chunks
    host = rc.TraceClosest(w["O4"], w["D4"])
    dev = torch.device("cuda", 0)
    dO, dD = torch.from_numpy(w["O4"]).to(dev), torch.from_numpy(w["D4"]).to(dev)
    out = torch.empty((len(host), 4), dtype=torch.float32, device=dev)
    torch.cuda.synchronize()
    rc.TraceClosestDevice(dO.data_ptr(), dD.data_ptr(), len(host), out.data_ptr())
    torch.cuda.synchronize()
    assert np.array_equal(out.cpu().numpy().view(np.uint32), host.view(np.uint32))
    bits = rc.TraceAny(w["O4"], np.concatenate([w["D4"][:, :3], np.full((len(host), 1), 70.0, np.float32)], axis=1))
    assert np.array_equal(bits, core.unpack_hits(host)["t"] < np.float32(70.0))


def test_c2_pinned_host_wave_equals_device_wave(c2):
    """Pinned host planes take the zero-copy path (the kernel fetches rays over PCIe, several chunk launches on two
    streams); results must equal the pageable-buffer path, the device-pointer wave, and the path with zeroCopy off."""
    import torch
    rc = c2["rc"]
    lib, h = rc.lib, rc.h
    rc.Setting("zeroCopyBelow", 1 << 30); rc.Setting("pinnedChunks", 4)  # force the zero-copy path for the big waves too
    for w in c2["waves"][:3]:  # primary (4 chunks), shadow (any-hit), diffuse
        n = w["O4"].shape[0]
        hO, hD = torch.from_numpy(w["O4"]).pin_memory(), torch.from_numpy(w["D4"]).pin_memory()
        if w["kind"] == "closest":
            ref = rc.TraceClosest(w["O4"], w["D4"])  # pageable numpy buffers: staged chunks
            outs = []
            for zc in (1, 0, 1):
                rc.Setting("zeroCopy", zc)
                out = torch.full((n, 4), -1.0).pin_memory()
                assert lib.b200_trace_closest(h, hO.data_ptr(), hD.data_ptr(), n, out.data_ptr()) == 0
                outs.append(out.numpy().view(np.uint32).copy())
            for o in outs:
                assert np.array_equal(o, ref.view(np.uint32))
        else:
            ref = rc.TraceAny(w["O4"], w["D4"])
            for zc in (1, 0):
                rc.Setting("zeroCopy", zc)
                bits = torch.full(((n + 31) // 32,), -1, dtype=torch.int32).pin_memory()
                assert lib.b200_trace_any(h, hO.data_ptr(), hD.data_ptr(), n, bits.data_ptr()) == 0
                got = np.unpackbits(bits.numpy().view(np.uint8), bitorder="little")[:n].astype(bool)
                assert np.array_equal(got, ref)
    rc.Setting("zeroCopy", 1)
    # ragged sizes around the chunk / bitmask-word boundaries
    w = c2["waves"][0]
    for n in (1, 31, 33, 1025, 262143, 262145, 524289 + 17):
        hO, hD = torch.from_numpy(w["O4"][:n].copy()).pin_memory(), torch.from_numpy(w["D4"][:n].copy()).pin_memory()
        out = torch.full((n, 4), -1.0).pin_memory()
        assert lib.b200_trace_closest(h, hO.data_ptr(), hD.data_ptr(), n, out.data_ptr()) == 0
        assert np.array_equal(out.numpy().view(np.uint32), rc.TraceClosest(w["O4"][:n], w["D4"][:n]).view(np.uint32))
    rc.Setting("zeroCopyBelow", 400000); rc.Setting("pinnedChunks", 1)


def test_c2_refit_with_same_vertices_is_identity(c2):
    rc = c2["rc"]
    w = c2["waves"][2]
    n = 300_000
    before = rc.TraceClosest(w["O4"][:n], w["D4"][:n])
    rc.SetGeometry(0, c2["verts"])  # same triangle count -> GPU refit; boxes are re-quantised from identical positions
    rc.UpdateToplevel()
    st = rc.GetCoreStats()
    assert st["refitTime"] > 0
    print("refit of 1M triangles: %.3f ms device" % st["refitTime"])
    after = rc.TraceClosest(w["O4"][:n], w["D4"][:n])
    assert np.array_equal(before.view(np.uint32), after.view(np.uint32))


def test_c3_10m_against_the_reference_fixture():
    """BASELINE.json configs[2] at size: the 10 000 000-triangle soup against hit buffers made by the unmodified reference
    (tests/golden/make_c3_fixture.py: BVH2::Rebuild + Traverse of lib/RenderCore_Bart/bvh.cpp over the same soup): 40 000 primary rays
    and the four diffuse-bounce waves behind them.  Bit-exact t and triangle ids except the documented classes (tests/parity.py)."""
    import os
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "c3_10m.npz")
    if not os.path.exists(path):
        pytest.skip("tests/golden/c3_10m.npz missing (python tests/golden/make_c3_fixture.py)")
    g = np.load(path)
    verts = scenes.triangle_soup(10_000_000)
    rc = core.RenderCoreB200(0)
    rc.SetGeometry(0, verts); rc.SetInstance(0, 0); rc.SetInstance(1, -1); rc.UpdateToplevel()
    total = exceptions = 0
    for k in range(int(g["waves"])):
        O, D, t, prim = g["O%d" % k], g["D%d" % k], g["t%d" % k], g["prim%d" % k]
        h = core.unpack_hits(rc.TraceClosest(*core.pack_rays(O, D, tmin=workload.EPS)))
        bad = np.flatnonzero((h["prim"] != prim) | (h["t"].view(np.uint32) != t.view(np.uint32)))
        # a disagreement is acceptable only as a documented class: our hit is CLOSER and is what the reference arithmetic gives for that
        # triangle (a box the reference's slab test culled), or an exact tie; never a farther hit
        for i in bad:
            tri = verts[h["prim"][i]] if h["prim"][i] >= 0 else None
            assert tri is not None and h["t"][i] <= t[i], (k, i, h["t"][i], t[i], h["prim"][i], prim[i])
        total += len(t); exceptions += len(bad)
        print("C3 10M wave %d: %d rays, %d hits, %d exceptions" % (k, len(t), int((prim >= 0).sum()), len(bad)))
    assert exceptions <= max(1, int(1e-5 * total)), (exceptions, total)
    rc.Shutdown()


def test_c3_class_soup_properties():
    """10M-triangle soup (scene not L2-resident): permutation invariance, any-hit consistency, and a sampled check against
    the reference arithmetic by exhaustive search over the triangles near each reported hit is too slow here -- the oracle
    comparison at this class of scene runs at 200K triangles in test_parity_gpu.py."""
    verts = scenes.triangle_soup(10_000_000)
    rc = core.RenderCoreB200(0)
    rc.SetGeometry(0, verts); rc.SetInstance(0, 0); rc.SetInstance(1, -1); rc.UpdateToplevel()
    O, D = scenes.random_rays(1_000_000, 55.0, seed=13)
    O4, D4 = core.pack_rays(O, D, tmin=1e-4)
    a = rc.TraceClosest(O4, D4)
    h = core.unpack_hits(a)
    assert 0.05 < (h["prim"] >= 0).mean() < 1.0
    perm = np.random.default_rng(1).permutation(len(a))
    b = rc.TraceClosest(np.ascontiguousarray(O4[perm]), np.ascontiguousarray(D4[perm]))
    assert np.array_equal(b.view(np.uint32), a.view(np.uint32)[perm])
    D4s = D4.copy(); D4s[:, 3] = 20.0
    assert np.array_equal(rc.TraceAny(O4, D4s), h["t"] < np.float32(20.0))
    # the reported t must be what the reference triangle test computes for the reported triangle (CPU restatement, numpy float32)
    idx = np.flatnonzero(h["prim"] >= 0)[:50_000]
    tri = verts[h["prim"][idx]]
    e1, e2 = tri[:, 1] - tri[:, 0], tri[:, 2] - tri[:, 0]
    d, o = scenes._normalize_ref(D[idx]), O[idx]  # the TLAS leaf re-normalises the direction (bvh.cpp:607-610)
    hx = d[:, 1] * e2[:, 2] - d[:, 2] * e2[:, 1]; hy = d[:, 2] * e2[:, 0] - d[:, 0] * e2[:, 2]; hz = d[:, 0] * e2[:, 1] - d[:, 1] * e2[:, 0]
    aa = (e1[:, 0] * hx + e1[:, 1] * hy) + e1[:, 2] * hz
    f = np.float32(1.0) / aa
    s = o - tri[:, 0]
    qx = s[:, 1] * e1[:, 2] - s[:, 2] * e1[:, 1]; qy = s[:, 2] * e1[:, 0] - s[:, 0] * e1[:, 2]; qz = s[:, 0] * e1[:, 1] - s[:, 1] * e1[:, 0]
    tt = f * ((e2[:, 0] * qx + e2[:, 1] * qy) + e2[:, 2] * qz)
    assert np.array_equal(tt.astype(np.float32).view(np.uint32), h["t"][idx].view(np.uint32))
    rc.Shutdown()
