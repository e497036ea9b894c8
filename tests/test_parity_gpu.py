"""GPU parity tests: every call goes through the C ABI of libRenderCore_B200.so (host buffers in, host
buffers out) and is compared with the golden hit buffers / the CPU oracle on the same seeded inputs."""
import os

import numpy as np
import pytest

import cases
import parity
from lighthouse2_b200 import core, scenes

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture(scope="module")
def oracles():
    import oracle
    oracle.build()
    return oracle


def _run_case(case, oracles, golden=True, isect=0):
    O, D = case["rays"]
    rc = core.RenderCoreB200(0)
    rc.Setting("isectMode", isect)
    cases.load_core(rc, case)
    O4, D4 = core.pack_rays(O, D)
    hits = rc.TraceClosest(O4, D4)
    port = cases.load_oracle(case, "port")
    if golden:
        g = np.load(os.path.join(GOLDEN, case["name"] + ".npz"))
        ref_t, ref_prim, ref_inst = g["t"], g["prim"], g["inst"]
    else:
        ref_t, ref_prim, ref_inst = port.trace(O, D, nthreads=8)
    rep = parity.compare(case["name"], hits, port, O, D, ref_t, ref_prim, ref_inst)
    rc.Shutdown()
    return rep, hits


def test_c1_golden_hit_buffer(oracles):
    """BASELINE configs[0]: 1K-triangle mesh, 640x360 primary rays vs the golden buffer made by the reference."""
    rep, hits = _run_case(cases.case_c1(), oracles)
    parity.assert_parity(rep)
    assert rep["rays"] == 230400


@pytest.mark.parametrize("make", [cases.case_soup, cases.case_instanced, cases.case_axis_aligned, cases.case_single_triangle])
def test_small_cases_golden(oracles, make):
    rep, _ = _run_case(make(), oracles)
    parity.assert_parity(rep)


@pytest.mark.parametrize("scale", [None, (0.5, 1.8)])
def test_rotated_and_scaled_instances_vs_exhaustive_search(oracles, scale):
    """Rotated (and non-uniformly scaled) instances through the general two-level kernels.  The reference's own TLAS is wrong for
    them -- its leaf bounds transform only two corners of the mesh box (bvh.cpp:579-587) and it misses geometry -- so the checker is
    the exhaustive search with the reference's instance entry and triangle arithmetic (oracle brute mode: bvh.cpp:607-610 +
    common.h:19-50 over every triangle of every instance; t is compared across instances in object-space units, as the reference
    does).  Bar: t bit-exact and (instance, triangle) identical on every ray except exact ties.  The `inst27rot` golden buffer (made
    by the unmodified reference) must agree wherever the reference found a hit at all: what it found is never closer than ours."""
    case = cases.case_instanced(rotate=True, scale=scale)
    O, D = case["rays"]
    rc = cases.load_core(core.RenderCoreB200(0), case)
    h = core.unpack_hits(rc.TraceClosest(*core.pack_rays(O, D)))
    port = cases.load_oracle(case, "port")
    tb, pb, ib = port.trace(O, D, nthreads=8, mode="brute")
    same = (h["t"].view(np.uint32) == tb.view(np.uint32)) & (h["prim"] == pb) & ((h["inst"] == ib) | (pb < 0))
    bad = np.flatnonzero(~same)
    # what is left must be exact ties: same t bits, another (instance, triangle) whose reference-arithmetic t is that t
    t_choice = port.t_of(O[bad], D[bad], h["inst"][bad], h["prim"][bad])
    assert np.array_equal(h["t"][bad].view(np.uint32), tb[bad].view(np.uint32)) and np.array_equal(t_choice.view(np.uint32), tb[bad].view(np.uint32)), \
        (len(bad), bad[:5], h["t"][bad][:5], tb[bad][:5])
    assert len(bad) <= max(1, int(1e-4 * len(O))), len(bad)
    assert (h["prim"] >= 0).mean() > 0.2
    if scale is None:
        g = np.load(os.path.join(GOLDEN, "inst27rot.npz"))
        ref_hit = g["prim"] >= 0
        assert np.all(h["t"][ref_hit] <= g["t"][ref_hit]), "the reference reports a hit closer than the exhaustive search"
        agree = (h["t"].view(np.uint32) == g["t"].view(np.uint32)) & (h["prim"] == g["prim"])
        print("rotated instances: the reference's own TLAS agrees on %.2f %% of the rays (it misses geometry under rotation)" % (100 * agree.mean()))
    # any-hit through the same kernels
    tmax = np.random.default_rng(3).uniform(0.5, 12.0, len(O)).astype(np.float32)
    assert np.array_equal(rc.TraceAny(*core.pack_rays(O, D, tmax=tmax)), tb < tmax)
    print("rotated%s instances: %d rays, %d exact ties" % (" + scaled" if scale else "", len(O), len(bad)))
    rc.Shutdown()


def test_c4_class_refit_and_toplevel_rebuild_vs_oracle(oracles):
    """BASELINE.json configs[3] at size: 1000 translation-only instances of a 100 000-triangle deforming sphere; three frames, each
    with new vertex positions (GPU refit through SetGeometry) and new instance transforms (top-level rebuild), 200 000 primary
    and 200 000 shadow rays per frame against the reference traversal (the oracle rebuilds its BVHs every frame, which is what the
    reference does: rendercore.cpp:78,199)."""
    n_inst, n_tris, n_rays = 1000, 100_000, 200_000
    rc = core.RenderCoreB200(0)
    rng = np.random.default_rng(41)
    view = scenes.view_pyramid(width=1920, height=1080, **scenes.C4_CAMERA)
    # rays through an irrational sub-pixel position: through the pixel centres 3 of 200 000 rays of frame 0 run exactly along an edge shared
    # by two triangles of the UV sphere (bit-identical t on both: the tie class of tests/parity.py, 1.5e-5 of the rays against a bar of
    # 1e-5); SURVEY.md section 7 names the remedy used here
    Oall, Dall = scenes.primary_rays(view, 1920, 1080, offset=(0.6180339887, 0.7548776662))
    port = None
    for frame in range(3):
        tm = 0.1 * frame
        case = cases.case_c4_class(n_inst, n_tris, time=tm)
        rc.SetGeometry(0, case["meshes"][0])  # frame 0 builds, frames 1 and 2 refit on the device
        for i, (mesh, mat) in enumerate(case["instances"]):
            rc.SetInstance(i, mesh, mat)
        rc.SetInstance(n_inst, -1)
        rc.UpdateToplevel()
        if frame:
            assert rc.GetCoreStats()["refitTime"] > 0
        if port is not None:
            port.close()
        port = cases.load_oracle(case, "port")
        idx = np.sort(rng.choice(len(Oall), n_rays, replace=False))
        O, D = np.ascontiguousarray(Oall[idx]), np.ascontiguousarray(Dall[idx])
        hits = rc.TraceClosest(*core.pack_rays(O, D))
        t, prim, inst = port.trace(O, D, nthreads=8)
        rep = parity.compare("c4 frame %d" % frame, hits, port, O, D, t, prim, inst)
        # (1000 instances of a tessellated sphere: rays through shared edges.  At most 2 of 200 000 rays with the host-built tree, 5 with the
        # device-built one that is the default now -- the count is a property of who visits which triangle first, not of an error)
        parity.assert_parity(rep, max_tie_rate=1e-4)
        assert (prim >= 0).mean() > 0.3
        # one shadow ray per hit towards a point light (tmax = distance - 2 eps), any-hit <=> reference closest t < tmax
        hit = prim >= 0
        P = O[hit] + D[hit] * t[hit, None] * np.float32(0.999)
        L = np.asarray((0.0, 80.0, -40.0), np.float32) - P
        dist = np.linalg.norm(L, axis=1).astype(np.float32)
        Ld = scenes._normalize_ref(L.astype(np.float32))
        ts, _, _ = port.trace(P, Ld, nthreads=8)
        occ = rc.TraceAny(*core.pack_rays(P, Ld, tmin=1e-4, tmax=dist - np.float32(2e-4)))
        bad = np.flatnonzero(occ != (ts < dist - np.float32(2e-4)))
        assert len(bad) <= max(1, int(1e-5 * len(P))), (len(bad), bad[:5])
    st = rc.GetCoreStats()
    print("C4 class: %d instances, %d BLAS nodes, %d TLAS nodes, refit %.3f ms, top level %.3f ms" % (st["instanceCount"], st["wideNodeCount"], st["tlasNodeCount"], st["refitTime"], st["tlasTime"]))
    port.close()
    rc.Shutdown()


@pytest.mark.parametrize("n_inst", [1, 2, 3, 9, 27, 500, 2048])
def test_device_toplevel_equals_host_toplevel(n_inst):
    """The top level built by one thread block on the device (build.cu::tlasBuildKernel: Morton keys, bitonic sort, radix tree, SAH
    collapse, quantised planes, instance records) gives the hits of the host-built one, bit for bit (any conservative hierarchy does:
    the closest hit does not depend on the tree), for rotated and scaled instances, from 1 instance to the largest count one block takes."""
    sph = scenes.uv_sphere(12, 7, 1.0)
    grid = scenes.grid_mesh(4, 4)
    M = scenes.lattice_transforms(n_inst, spacing=3.0, rotate=True, seed=8, scale=(0.6, 1.5))
    case = dict(name="tlas", meshes=[sph, grid], instances=[(i % 2, M[i]) for i in range(n_inst)])
    O, D = scenes.random_rays(100_000, 2.0 + 1.5 * round(n_inst ** (1 / 3)), seed=12)
    out = []
    for device in (0, 1):
        rc = core.RenderCoreB200(0)
        rc.Setting("deviceTlas", device)
        cases.load_core(rc, case)
        out.append(rc.TraceClosest(*core.pack_rays(O, D)))
        st = rc.GetCoreStats()
        assert st["instanceCount"] == n_inst and st["tlasNodeCount"] >= 1
        print("instances %d, deviceTlas %d: %d top-level nodes, %.3f ms" % (n_inst, device, st["tlasNodeCount"], st["tlasTime"]))
        rc.Shutdown()
    assert (core.unpack_hits(out[0])["prim"] >= 0).mean() > 0.02
    assert np.array_equal(out[0].view(np.uint32), out[1].view(np.uint32))


def test_bigger_soup_vs_oracle(oracles):
    case = cases.case_soup(n_tris=200_000, n_rays=300_000, seed=21)
    rep, _ = _run_case(case, oracles, golden=False)
    parity.assert_parity(rep)


def test_any_hit_matches_closest(oracles):
    """occluded(tmax) <=> oracle closest t < tmax (Appendix A.2), and == the GPU's own closest-hit wave."""
    case = cases.case_soup(n_tris=50_000, n_rays=200_000, seed=31)
    O, D = case["rays"]
    rc = cases.load_core(core.RenderCoreB200(0), case)
    rng = np.random.default_rng(1)
    tmax = rng.uniform(0.5, 12.0, len(O)).astype(np.float32)
    O4, D4 = core.pack_rays(O, D, tmin=0.0, tmax=tmax)
    occ = rc.TraceAny(O4, D4)
    port = cases.load_oracle(case, "port")
    t, prim, _ = port.trace(O, D, nthreads=8)
    expect = t < tmax
    bad = np.flatnonzero(occ != expect)
    # a ray whose closest t equals tmax to the last bit cannot disagree (strict <); anything else is an error
    assert len(bad) <= max(1, int(1e-5 * len(O))), (len(bad), bad[:10])
    O4c, D4c = core.pack_rays(O, D)
    h = core.unpack_hits(rc.TraceClosest(O4c, D4c))
    assert np.array_equal(occ, h["t"] < tmax)
    rc.Shutdown()


def test_queued_waves_equal_the_synchronous_calls():
    """b200_trace_*_async + b200_wait: seven waves queued back to back (more than the three staging slots, sizes on both sides of the
    zero-copy and the chunking thresholds, closest and any-hit mixed) return what the synchronous calls return, bit for bit."""
    case = cases.case_soup(n_tris=40_000, n_rays=1_300_000, seed=77)
    O, D = case["rays"]
    rc = cases.load_core(core.RenderCoreB200(0), case)
    sizes = [1_300_000, 70_001, 450_000, 33, 600_000, 1_000_000, 0, 250_000]
    kinds = ["closest", "any", "closest", "closest", "any", "closest", "closest", "any"]
    waves, expect = [], []
    for i, (n, kind) in enumerate(zip(sizes, kinds)):
        lo = (i * 37_003) % (len(O) - n + 1)
        O4, D4 = core.pack_rays(O[lo:lo + n], D[lo:lo + n], tmax=1e34 if kind == "closest" else 6.0)
        waves.append((kind, O4, D4))
        expect.append(rc.TraceClosest(O4, D4) if kind == "closest" else rc.TraceAny(O4, D4))
    for _ in range(2):  # the second round reuses every slot
        got = rc.TraceQueued(waves)
        for kind, e, g in zip(kinds, expect, got):
            assert e.shape == g.shape
            assert np.array_equal(e.view(np.uint32), g.view(np.uint32)) if kind == "closest" else np.array_equal(e, g)
    rc.Shutdown()


def test_one_instance_kernels_equal_the_two_level_kernels():
    """A scene of one instance runs the SINGLE instances of the kernels (rays enter the instance when fetched, no top-level
    step); Setting singleInstanceShortcut=0 sends the same scene through the general two-level kernels.  Same bits, closest and any hit,
    also for a transformed instance."""
    base = cases.case_soup(n_tris=30_000, n_rays=150_000, seed=91)
    O, D = base["rays"]
    M = np.eye(4, dtype=np.float32)
    M[:3, 3] = (0.25, -0.5, 0.125)
    for mat in (None, M.reshape(16)):
        case = dict(base, instances=[(0, mat)])
        out = []
        for shortcut in (1, 0):
            rc = core.RenderCoreB200(0)
            rc.Setting("singleInstanceShortcut", shortcut)
            cases.load_core(rc, case)
            O4, D4 = core.pack_rays(O, D)
            hits = rc.TraceClosest(O4, D4)
            O4s, D4s = core.pack_rays(O, D, tmax=5.0)
            out.append((hits, rc.TraceAny(O4s, D4s)))
            rc.Shutdown()
        assert np.array_equal(out[0][0].view(np.uint32), out[1][0].view(np.uint32))
        assert np.array_equal(out[0][1], out[1][1])
        assert (core.unpack_hits(out[0][0])["prim"] >= 0).mean() > 0.2


def test_empty_mesh_zero_rays_and_misses():
    rc = core.RenderCoreB200(0)
    case = cases.case_empty()
    cases.load_core(rc, case)
    O4, D4 = core.pack_rays(*case["rays"])
    h = core.unpack_hits(rc.TraceClosest(O4, D4))
    assert (h["prim"] == -1).all() and (h["t"] == np.float32(1e34)).all()
    assert rc.TraceClosest(O4[:0], D4[:0]).shape == (0, 4)
    assert not rc.TraceAny(O4, D4).any()
    rc.Shutdown()


def test_ragged_ray_counts(oracles):
    """Ray counts that are not multiples of the warp / block size, through the 1-bit output too."""
    case = cases.case_soup(n_tris=3000, n_rays=5000, seed=41)
    O, D = case["rays"]
    rc = cases.load_core(core.RenderCoreB200(0), case)
    port = cases.load_oracle(case, "port")
    t, prim, _ = port.trace(O, D, nthreads=4)
    for n in (1, 31, 33, 127, 129, 4097):
        O4, D4 = core.pack_rays(O[:n], D[:n])
        h = core.unpack_hits(rc.TraceClosest(O4, D4))
        assert np.array_equal(h["prim"], prim[:n]) and np.array_equal(h["t"].view(np.uint32), t[:n].view(np.uint32))
        O4, D4 = core.pack_rays(O[:n], D[:n], tmax=3.0)
        assert np.array_equal(rc.TraceAny(O4, D4), t[:n] < 3.0)
    rc.Shutdown()


def test_tmin_tmax_window(oracles):
    """The wave's own (tmin, tmax) interval on top of the reference's fixed (1e-4, 1e4) window."""
    case = cases.case_soup(n_tris=3000, n_rays=20000, seed=43)
    O, D = case["rays"]
    rc = cases.load_core(core.RenderCoreB200(0), case)
    port = cases.load_oracle(case, "port")
    t, prim, _ = port.trace(O, D, nthreads=4)
    O4, D4 = core.pack_rays(O, D, tmin=0.0, tmax=2.0)
    h = core.unpack_hits(rc.TraceClosest(O4, D4))
    inside = t < 2.0
    assert np.array_equal(h["prim"][inside], prim[inside])
    assert (h["prim"][~inside] == -1).all()
    rc.Shutdown()


def test_refit_equals_rebuild(oracles):
    """GPU refit (animated mesh) gives the same hits as a fresh build of the deformed mesh, and as the oracle
    (which rebuilds: the reference has no refit, rendercore.cpp:78)."""
    n_tris = 20000
    v0 = scenes.deforming_sphere(n_tris, time=0.0, radius=3.0)
    v1 = scenes.deforming_sphere(n_tris, time=1.3, radius=3.0)
    O, D = scenes.random_rays(100_000, 5.0, seed=77)
    O4, D4 = core.pack_rays(O, D)
    rc = core.RenderCoreB200(0)
    rc.SetGeometry(0, v0); rc.SetInstance(0, 0); rc.SetInstance(1, -1); rc.UpdateToplevel()
    rc.SetGeometry(0, v1)  # same triangle count -> GPU refit
    rc.UpdateToplevel()
    st = rc.GetCoreStats()
    assert st["refitTime"] > 0
    hits = rc.TraceClosest(O4, D4)
    case = dict(name="refit", meshes=[v1], instances=[(0, None)], rays=(O, D))
    port = cases.load_oracle(case, "port")
    t, prim, inst = port.trace(O, D, nthreads=8)
    rep = parity.compare("refit", hits, port, O, D, t, prim, inst)
    parity.assert_parity(rep)
    rc2 = cases.load_core(core.RenderCoreB200(0), case)
    h2 = rc2.TraceClosest(O4, D4)
    assert np.array_equal(hits.view(np.uint32)[:, 1:], h2.view(np.uint32)[:, 1:])
    rc.Shutdown(); rc2.Shutdown()


def test_step_counters(oracles):
    case = cases.case_soup(n_tris=20000, n_rays=20000, seed=51)
    rc = cases.load_core(core.RenderCoreB200(0), case)
    rc.Setting("countSteps", 1)
    O4, D4 = core.pack_rays(*case["rays"])
    rc.TraceClosest(O4, D4)
    nodes, tris = rc.StepCounters(len(O4))
    assert nodes.min() >= 1 and nodes.mean() < 200 and tris.mean() < 200
    rc.Shutdown()


def test_errors_are_reported_not_swallowed():
    """The C ABI returns non-zero with a message (the plugin shim turns that into the reference's FatalError + exit,
    lib/platform/system.cpp:221-236); nothing falls back silently."""
    rc = core.RenderCoreB200(0)
    tri = scenes.grid_mesh(2, 2)
    v4 = np.ones((len(tri) * 3, 4), np.float32)
    lib, h = rc.lib, rc.h
    assert lib.b200_set_geometry(h, 0, v4.ctypes.data, len(tri) * 3 - 1, len(tri), None, None) != 0  # vertexCount != 3 T
    assert b"vertexCount" in lib.b200_last_error()
    assert lib.b200_set_geometry(h, 3, v4.ctypes.data, len(tri) * 3, len(tri), None, None) != 0      # mesh indices must be dense
    assert lib.b200_set_geometry(h, 0, None, len(tri) * 3, len(tri), None, None) != 0                # null vertices
    rc.SetGeometry(0, tri)
    with pytest.raises(core.B200Error):
        rc.SetInstance(0, 5)                                                                          # unknown mesh
    with pytest.raises(core.B200Error):
        rc.Refit(0, tri[:-1])                                                                         # triangle count changed
    with pytest.raises(core.B200Error):
        rc.Setting("noSuchKey", 1)
    with pytest.raises(core.B200Error):
        rc.Setting("maxPathLength", 99)
    view = scenes.view_pyramid(width=16, height=16, **scenes.C1_CAMERA)
    with pytest.raises(core.B200Error):
        rc.Render(view, 1)                                                                            # before SetTarget
    with pytest.raises(core.B200Error):
        rc.SetTarget(0, 16, 1)
    rc.SetInstance(0, 0); rc.SetInstance(1, -1); rc.UpdateToplevel()
    O, D = scenes.random_rays(8, 2.0, seed=1)
    O4, D4 = core.pack_rays(O, D)
    assert lib.b200_trace_closest(h, O4.ctypes.data, D4.ctypes.data, -1, None) != 0                   # negative count
    hits = rc.TraceClosest(O4, D4)                                                                    # and the core still works afterwards
    assert hits.shape == (8, 4)
    # the queued waves and the peer-memory gather report misuse the same way
    assert lib.b200_wait(h) == 0                                                                      # nothing queued: not an error
    assert lib.b200_trace_closest_async(h, O4.ctypes.data, D4.ctypes.data, -1, None) != 0
    buf = (core.C.c_char * 128)()
    assert lib.b200_peer_handle(h, buf) != 0 and b"SetTarget" in lib.b200_last_error()                # no target yet
    assert lib.b200_gather_finalize(h, 1) != 0
    rc.SetTarget(16, 16, 1)
    assert lib.b200_gather_finalize(h, 1) != 0 and b"OpenPeers" in lib.b200_last_error()              # peers not mapped
    assert lib.b200_open_peers(h, buf, 1, 0) != 0 and b"PeerHandle" in lib.b200_last_error()          # handle not taken yet
    assert lib.b200_peer_handle(h, buf) == 0
    assert lib.b200_open_peers(h, buf, 40, 0) != 0                                                    # more ranks than the peer table holds
    assert lib.b200_open_peers(h, buf, 1, 0) == 0                                                     # a world of one: nothing to map
    rc.Render(view, 1)
    assert lib.b200_gather_finalize(h, 0) != 0                                                        # zero samples
    assert lib.b200_gather_finalize(h, 1) == 0                                                        # ... and works as plain finalize
    assert np.isfinite(rc.ReadFrame()).all()
    assert lib.b200_close_peers(h) == 0
    rc.Shutdown()
