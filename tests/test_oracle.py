"""CPU tests of the checkers themselves: the plain-C port (oracle/bart_port.c) against the golden vectors
generated from the unmodified reference (tests/golden/make_golden.py), against the live reference library
where it exists, and hand-computed known-answer tests of the two primitives the path is made of."""
import os

import numpy as np
import pytest

import cases
import oracle
from oracle import Oracle

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
ALL_CASES = [cases.case_c1, cases.case_soup, cases.case_instanced, lambda: cases.case_instanced(rotate=True),
             cases.case_axis_aligned, cases.case_single_triangle]


def _kinds():
    return ["port", "ref"] if oracle.have_ref() else ["port"]


@pytest.mark.parametrize("make", ALL_CASES)
def test_port_matches_golden(built_oracles, make):
    case = make()
    g = np.load(os.path.join(GOLDEN, case["name"] + ".npz"))
    o = cases.load_oracle(case, "port")
    t, prim, inst = o.trace(*case["rays"], nthreads=4)
    assert np.array_equal(t.view(np.uint32), g["t"].view(np.uint32)), "t must be bit-identical to the reference"
    assert np.array_equal(prim, g["prim"])
    assert np.array_equal(inst, g["inst"])


@pytest.mark.skipif(not oracle.have_ref(), reason="oracle/_ref not built (no /root/reference here)")
@pytest.mark.parametrize("make", ALL_CASES)
def test_reference_library_matches_golden(built_oracles, make):
    case = make()
    g = np.load(os.path.join(GOLDEN, case["name"] + ".npz"))
    o = cases.load_oracle(case, "ref")
    t, prim, _ = o.trace(*case["rays"], nthreads=4)
    assert np.array_equal(t.view(np.uint32), g["t"].view(np.uint32))
    assert np.array_equal(prim, g["prim"])


def test_bvh_equals_exhaustive_for_translations(built_oracles):
    """With translation-only instances the reference BVH result equals the exhaustive search over every
    triangle (same arithmetic), except rays that graze a box (the reference slab test is not conservative)."""
    for make in (cases.case_soup, cases.case_instanced):
        case = make()
        o = cases.load_oracle(case, "port")
        t, prim, inst = o.trace(*case["rays"], nthreads=4)
        tb, pb, ib = o.trace(*case["rays"], nthreads=4, mode="brute")
        bad = (t.view(np.uint32) != tb.view(np.uint32)) | (prim != pb)
        assert bad.mean() < 1e-4, bad.sum()


def test_empty_and_ragged(built_oracles):
    """Zero rays, and rays that miss everything.  (A mesh with ZERO triangles is undefined behaviour in the
    reference -- BVH2::Rebuild writes pool[0] of a `new BVHNode[0]`, bvh.cpp:64-69 -- so the oracles reject it;
    the CUDA core defines it as `every ray misses`, tested in test_parity_gpu.py.)"""
    case = cases.case_single_triangle()
    for kind in _kinds():
        o = cases.load_oracle(case, kind)
        t, prim, _ = o.trace(np.zeros((0, 3), np.float32), np.zeros((0, 3), np.float32))
        assert t.shape == (0,)
        O = np.tile(np.float32([0, 0, 10]), (33, 1))
        D = np.tile(np.float32([0, 0, 1]), (33, 1))
        t, prim, _ = o.trace(O, D)
        assert (prim == -1).all() and (t == np.float32(1e34)).all()
    with pytest.raises(AssertionError):
        Oracle("port").SetGeometry(0, np.zeros((0, 3, 3), np.float32))


# ---- known-answer tests: Ray::IntersectTriangle (lib/RenderCore_Bart/common.h:19-50) ---------------------
TRI = [0, 0, 5, 4, 0, 5, 0, 4, 5]  # right triangle in the plane z=5, legs 4


@pytest.mark.parametrize("kind", ["port", "ref"])
def test_kat_triangle(built_oracles, kind):
    if kind == "ref" and not oracle.have_ref():
        pytest.skip("no reference library here")
    o = Oracle(kind)
    Z = [0, 0, 1]
    hit, t = o.intersect_triangle([1, 1, 0], Z, TRI)
    assert hit and t == 5.0                                    # interior: t = 5 exactly
    assert o.intersect_triangle([0, 1, 0], Z, TRI)[0]          # u == 0 edge is inclusive (u < 0.0 rejects)
    assert o.intersect_triangle([1, 0, 0], Z, TRI)[0]          # v == 0 edge inclusive
    assert o.intersect_triangle([2, 2, 0], Z, TRI)[0]          # u + v == 1 hypotenuse inclusive (u+v > 1.0 rejects)
    assert not o.intersect_triangle([2.01, 2.01, 0], Z, TRI)[0]
    assert not o.intersect_triangle([-0.01, 1, 0], Z, TRI)[0]
    assert o.intersect_triangle([1, 1, 10], [0, 0, -1], TRI)[0]  # two-sided
    assert not o.intersect_triangle([1, 1, 5.0], Z, TRI)[0]      # t = 0 <= EPSILON
    assert not o.intersect_triangle([1, 1, 4.99995], Z, TRI)[0]  # t = 5e-5 <= EPSILON (1e-4)
    assert o.intersect_triangle([1, 1, 4.9998], Z, TRI)[0]       # t = 2e-4 > EPSILON
    far = [0, 0, 10005, 4, 0, 10005, 0, 4, 10005]
    assert not o.intersect_triangle([1, 1, 0], Z, far)[0]        # t = 10005 >= 1/EPSILON
    assert o.intersect_triangle([1, 1, 10], Z, far)[0]           # t = 9995
    assert not o.intersect_triangle([1, 1, 0], [1, 0, 0], TRI)[0]  # parallel: a == 0
    # absolute determinant cut: |a| < 1e-4 rejects small triangles outright (a = 2*area*cos)
    small = [0, 0, 5, 0.009, 0, 5, 0, 0.009, 5]                  # a = 8.1e-5
    assert not o.intersect_triangle([0.002, 0.002, 0], Z, small)[0]
    ok = [0, 0, 5, 0.011, 0, 5, 0, 0.011, 5]                     # a = 1.21e-4
    assert o.intersect_triangle([0.002, 0.002, 0], Z, ok)[0]
    # a miss leaves t untouched
    hit, t = o.intersect_triangle([9, 9, 0], Z, TRI, t0=123.0)
    assert not hit and t == 123.0


# ---- known-answer tests: BVH::IntersectAABB (lib/RenderCore_Bart/bvh.cpp:7-42) ---------------------------
@pytest.mark.parametrize("kind", ["port", "ref"])
def test_kat_aabb(built_oracles, kind):
    if kind == "ref" and not oracle.have_ref():
        pytest.skip("no reference library here")
    o = Oracle(kind)
    lo, hi = [-1, -1, -1], [1, 1, 1]
    d = o.normalize([1, 1, 1])
    hit, t = o.intersect_aabb([-3, -3, -3], d, lo, hi)
    assert hit and abs(t - 2 * np.sqrt(3)) < 1e-5                # entry distance
    hit, t = o.intersect_aabb([0, 0, 0], d, lo, hi)
    assert hit and t < 0                                         # origin inside: NEGATIVE entry t is returned
    hit, t = o.intersect_aabb([3, 3, 3], d, lo, hi)
    assert not hit and t < 0                                     # box behind: returns t_max < 0
    # axis-parallel ray: 1/0 = inf; inside the other two slabs -> hit, outside -> miss
    assert o.intersect_aabb([0.5, 0.5, -5], [0, 0, 1], lo, hi) == (True, 4.0)
    assert not o.intersect_aabb([1.5, 0.5, -5], [0, 0, 1], lo, hi)[0]
    # origin exactly on a slab plane of a parallel axis: (1-1)*inf = NaN; std::min/max keep the first
    # argument on NaN comparisons, which makes this particular ray a HIT in the reference
    hit, _ = o.intersect_aabb([1.0, 0.5, -5], [0, 0, 1], lo, hi)
    ref_like = Oracle("port").intersect_aabb([1.0, 0.5, -5], [0, 0, 1], lo, hi)[0]
    assert hit == ref_like
    # flat box (zero thickness), hit straight on
    assert o.intersect_aabb([0, 5, 0], [0, -1, 0], [-1, 0, -1], [1, 0, 1]) == (True, 5.0)


@pytest.mark.skipif(not oracle.have_ref(), reason="needs the reference library")
def test_port_equals_reference_on_random_primitives(built_oracles):
    """Primitive-level agreement incl. NaN/inf lanes: 20k random boxes/triangles, axis-aligned directions mixed in."""
    rng = np.random.default_rng(123)
    ref, port = Oracle("ref"), Oracle("port")
    for i in range(4000):
        O = rng.uniform(-2, 2, 3).astype(np.float32)
        D = rng.normal(size=3).astype(np.float32)
        if i % 3 == 0:
            D[rng.integers(0, 3)] = 0.0
        if i % 7 == 0:
            D[:] = 0
            D[rng.integers(0, 3)] = 1.0
            O = np.round(O)  # origins on integer planes -> exact 0*inf cases
        nrm = np.linalg.norm(D)
        D = port.normalize(D) if nrm > 0 else D
        lo = np.round(rng.uniform(-2, 0, 3)).astype(np.float32)
        hi = lo + np.round(rng.uniform(0, 3, 3)).astype(np.float32)
        a, b = ref.intersect_aabb(O, D, lo, hi), port.intersect_aabb(O, D, lo, hi)
        assert a[0] == b[0] and (a[1] == b[1] or (np.isnan(a[1]) and np.isnan(b[1])))
        tri = rng.uniform(-2, 2, 9).astype(np.float32)
        a, b = ref.intersect_triangle(O, D, tri), port.intersect_triangle(O, D, tri)
        assert a == b


def test_inverse_matches_reference(built_oracles):
    """Instance.inv_transform = mat4::Inverted (common_types.h:568-611)."""
    M = cases.scenes.lattice_transforms(16, rotate=True)
    for kind in _kinds():
        o = Oracle(kind)
        o.SetGeometry(0, cases.scenes.grid_mesh(2, 2))
        for i in range(16):
            o.SetInstance(i, 0, M[i])
        inv = np.stack([o.inv_transform(i) for i in range(16)])
        prod = inv.reshape(-1, 4, 4) @ M.reshape(-1, 4, 4)
        assert np.abs(prod - np.eye(4)).max() < 1e-5
        if kind == "port":
            port_inv = inv
        else:
            assert np.array_equal(inv.view(np.uint32), port_inv.view(np.uint32))
