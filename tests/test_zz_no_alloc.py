"""b200_trace_closest_noalloc -- the extend wave over host buffers with the kernel's scratch in a file-scope static device array
(traverse.cu), for callers that cannot call cudaMalloc.  The entry point has no counterpart in the reference (it was added
), so the checker is the existing entry point b200_trace_closest: same inputs, results bit for bit
(the bar of every integer / index / reference-arithmetic comparison in tests/), on two calls in a row that share the array.

This file is named to be collected LAST (a failure here must not hide other tests) and passes both
alone in a fresh process and after the other GPU tests."""

import numpy as np
import pytest

import cases
from lighthouse2_b200 import core, scenes

pytestmark = pytest.mark.gpu

MAX_RAYS = 65536  # B200_NOALLOC_MAX_RAYS of include/rendercore_b200.h


def _noalloc(rc, O4, D4):
    O4 = np.ascontiguousarray(O4, np.float32).reshape(-1, 4)
    D4 = np.ascontiguousarray(D4, np.float32).reshape(-1, 4)
    hits = np.full((O4.shape[0], 4), -7.0, np.float32)
    code = rc.lib.b200_trace_closest_noalloc(rc.h, O4.ctypes.data, D4.ctypes.data, O4.shape[0], hits.ctypes.data)
    return code, hits


def test_header_states_the_limit():
    import os
    hdr = open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "include", "rendercore_b200.h")).read()
    assert "#define B200_NOALLOC_MAX_RAYS %d" % MAX_RAYS in hdr


def test_noalloc_wave_equals_the_existing_entry_point_twice_in_a_row():
    # two-level scene (27 translated instances of two meshes) so that the general kernels and the instance table are exercised
    case = cases.case_instanced(n_inst=27, n_rays=MAX_RAYS)
    rc = cases.load_core(core.RenderCoreB200(0), case)
    O, D = case["rays"]
    O4, D4 = core.pack_rays(O, D)
    want = rc.TraceClosest(O4, D4)
    assert (core.unpack_hits(want)["prim"] >= 0).mean() > 0.1
    # call 1: the largest wave the array takes
    code, got = _noalloc(rc, O4, D4)
    assert code == 0, rc.lib.b200_last_error().decode()
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    # call 2, right behind it: a different, shorter wave -- the array still holds call 1's rays, hits and a spent fetch counter
    O2, D2 = scenes.random_rays(20_001, 6.0, seed=99)
    O42, D42 = core.pack_rays(O2, D2, tmin=1e-4)
    want2 = rc.TraceClosest(O42, D42)
    code, got2 = _noalloc(rc, O42, D42)
    assert code == 0, rc.lib.b200_last_error().decode()
    assert np.array_equal(got2.view(np.uint32), want2.view(np.uint32))
    # and the first wave once more: nothing of call 2 leaks into it
    code, got3 = _noalloc(rc, O4, D4)
    assert code == 0 and np.array_equal(got3.view(np.uint32), want.view(np.uint32))
    rc.Shutdown()


def test_noalloc_single_instance_scene_and_empty_wave():
    # one-instance scenes run the SINGLE kernel instances (C1: 1K-triangle mesh)
    case = cases.case_c1(160, 90)
    rc = cases.load_core(core.RenderCoreB200(0), case)
    O4, D4 = core.pack_rays(*case["rays"])
    want = rc.TraceClosest(O4, D4)
    for _ in range(2):
        code, got = _noalloc(rc, O4, D4)
        assert code == 0 and np.array_equal(got.view(np.uint32), want.view(np.uint32))
    code, got = _noalloc(rc, O4[:0], D4[:0])
    assert code == 0 and got.shape == (0, 4)
    rc.Shutdown()


def test_noalloc_rejects_what_it_cannot_do_without_allocating():
    case = cases.case_soup(n_tris=2000, n_rays=MAX_RAYS + 1)
    rc = core.RenderCoreB200(0)
    for i, m in enumerate(case["meshes"]):
        rc.SetGeometry(i, m)
    rc.SetInstance(0, 0)
    rc.SetInstance(1, -1)
    O4, D4 = core.pack_rays(*case["rays"])
    # the top level has not been built: rebuilding it could grow pools, so the call refuses instead
    code, hits = _noalloc(rc, O4[:100], D4[:100])
    assert code != 0 and b"b200_update_toplevel" in rc.lib.b200_last_error()
    assert (hits == -7.0).all()
    rc.UpdateToplevel()
    # one ray more than the fixed array holds
    code, hits = _noalloc(rc, O4, D4)
    msg = rc.lib.b200_last_error()
    assert code != 0 and b"B200_NOALLOC_MAX_RAYS" in msg and str(MAX_RAYS).encode() in msg
    assert (hits == -7.0).all()
    assert rc.lib.b200_trace_closest_noalloc(rc.h, None, None, 5, None) != 0
    # the core keeps working after the errors
    code, got = _noalloc(rc, O4[:MAX_RAYS], D4[:MAX_RAYS])
    assert code == 0 and np.array_equal(got.view(np.uint32), rc.TraceClosest(O4[:MAX_RAYS], D4[:MAX_RAYS]).view(np.uint32))
    rc.Shutdown()
