"""GPU BVH build (Setting gpuBuild = 1: Morton sort + binary radix tree, = 2: Morton sort + agglomerative clustering along the
curve (PLOC); lighthouse2_b200/csrc/build.cu), then the 8-wide collapse on the device, boxes by the refit.  Every test runs with both.  A different tree than the host SAH builder's, the same hits: checked against the golden
buffers / the CPU oracle and against the SAH-built core, including duplicate centroids, instancing, refit after the build,
and a second mesh appended behind a GPU-built one."""
import os

import numpy as np
import pytest

import cases
import parity
from lighthouse2_b200 import core, scenes

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
BUILDER = 1


@pytest.fixture(autouse=True, params=[1, 2, 3, 4, 5], ids=["lbvh", "ploc", "sahb", "sahb-all-levels", "sah"])
def _builder(request):
    global BUILDER
    BUILDER = request.param


def _choose_builder(rc):
    """Setting gpuBuild = 1 (radix tree), 2 (PLOC), 3 (SAH inside the radix tree's cells: the bottoms only, or -- 4 here -- every level),
    4 (binned SAH from the top; 5 here)."""
    rc.Setting("gpuBuild", 4 if BUILDER == 5 else min(BUILDER, 3))
    rc.Setting("gpuBuildLevels", 64 if BUILDER == 4 else 1)


def _gpu_core(case):
    rc = core.RenderCoreB200(0)
    _choose_builder(rc)
    return cases.load_core(rc, case)


def _against_oracle(case, golden):
    O, D = case["rays"]
    rc = _gpu_core(case)
    hits = rc.TraceClosest(*core.pack_rays(O, D))
    port = cases.load_oracle(case, "port")
    if golden:
        g = np.load(os.path.join(GOLDEN, case["name"] + ".npz"))
        ref = g["t"], g["prim"], g["inst"]
    else:
        ref = port.trace(O, D, nthreads=8)
    rep = parity.compare(case["name"], hits, port, O, D, *ref)
    st = rc.GetCoreStats()
    rc.Shutdown()
    return rep, st


def test_c1_golden_with_gpu_built_tree():
    rep, st = _against_oracle(cases.case_c1(), golden=True)
    parity.assert_parity(rep)
    assert rep["rays"] == 230400 and st["wideNodeCount"] > 10


@pytest.mark.parametrize("make", [cases.case_soup, cases.case_instanced])
def test_small_cases_golden_with_gpu_built_tree(make):
    rep, _ = _against_oracle(make(), golden=True)
    parity.assert_parity(rep)


@pytest.mark.parametrize("n_tris", [64, 65, 100, 255, 256, 257, 258, 300, 511, 513, 1000, 4097])
def test_sizes_around_the_builders_thresholds(n_tris):
    """64 triangles is the smallest mesh the device builds, 256 the largest subtree a warp finishes in shared memory: every size
    around them, against the exhaustive search of the oracle (which needs no tree)."""
    case = cases.case_soup(n_tris=n_tris, n_rays=20_000, seed=100 + n_tris)
    O, D = case["rays"]
    rc = _gpu_core(case)
    h = core.unpack_hits(rc.TraceClosest(*core.pack_rays(O, D)))
    port = cases.load_oracle(case, "port")
    tb, pb, _ = port.trace(O, D, nthreads=8, mode="brute")
    rc.Shutdown()
    assert np.array_equal(h["t"].view(np.uint32), tb.view(np.uint32))
    assert (h["prim"] != pb).mean() < 1e-3 and (pb >= 0).sum() > 50  # (ids may differ on exact ties only: t is bit-equal)


@pytest.mark.parametrize("shape", ["identical", "one-line", "flat", "two-clumps", "huge-and-tiny"])
def test_degenerate_distributions(shape):
    """Inputs on which a split search finds nothing to separate, or almost nothing: all triangles the same, all centroids on a line or in
    a plane, two far clumps, one triangle a million times the size of the others.  The build must end and the hits must be the exhaustive
    search's."""
    rng = np.random.default_rng(77)
    n = 3000
    tri = np.asarray([[0, 0, 0], [1, 0, 0], [0, 1, 0]], np.float32)
    if shape == "identical":
        verts = np.repeat(tri[None], n, axis=0)
    elif shape == "one-line":
        verts = tri[None] * np.float32(0.3) + np.stack([np.linspace(-5, 5, n), np.zeros(n), np.zeros(n)], 1).astype(np.float32)[:, None, :]
    elif shape == "flat":
        off = np.stack([rng.uniform(-5, 5, n), rng.uniform(-5, 5, n), np.zeros(n)], 1).astype(np.float32)
        verts = tri[None] * np.float32(0.3) + off[:, None, :]
    elif shape == "two-clumps":
        off = rng.normal(0, 0.05, (n, 3)).astype(np.float32) + np.where(np.arange(n)[:, None] % 2 == 0, -4.0, 4.0).astype(np.float32)
        verts = tri[None] * np.float32(0.1) + off[:, None, :]
    else:
        verts = scenes.triangle_soup(n, seed=3, extent=1.0, edge=1e-3)
        verts[0] = np.asarray([[-4e3, -4e3, 2], [4e3, -4e3, 2], [0, 6e3, 2]], np.float32)
    verts = np.ascontiguousarray(verts, np.float32)
    O, D = scenes.random_rays(30_000, 6.0, seed=9)
    case = dict(name=shape, meshes=[verts], instances=[(0, None)], rays=(O, D))
    rc = _gpu_core(case)
    h = core.unpack_hits(rc.TraceClosest(*core.pack_rays(O, D)))
    port = cases.load_oracle(case, "port")
    tb, pb, _ = port.trace(O, D, nthreads=8, mode="brute")
    rc.Shutdown()
    assert np.array_equal(h["t"].view(np.uint32), tb.view(np.uint32)), shape
    assert (pb >= 0).any()


def test_bigger_soup_vs_oracle_with_gpu_built_tree():
    rep, _ = _against_oracle(cases.case_soup(n_tris=200_000, n_rays=300_000, seed=21), golden=False)
    parity.assert_parity(rep)


def test_gpu_tree_equals_sah_tree_on_the_sphere_field():
    """300K-triangle sphere field: t bit-identical between the two trees on every ray; the id differs only where two triangles
    are hit at exactly the same t (shared edges; the reference keeps whichever its own tree visits first)."""
    verts = scenes.sphere_field(300_000)
    view = scenes.view_pyramid(width=960, height=540, **scenes.C2_CAMERA)
    O, D = scenes.primary_rays(view, 960, 540)
    rng = np.random.default_rng(4)
    Or, Dr = scenes.random_rays(400_000, 60.0, seed=8)
    O, D = np.concatenate([O, Or]), np.concatenate([D, Dr])
    out = []
    for gpu in (0, BUILDER):
        rc = core.RenderCoreB200(0)
        _choose_builder(rc) if gpu else rc.Setting("gpuBuild", 0)
        rc.SetGeometry(0, verts); rc.SetInstance(0, 0); rc.SetInstance(1, -1); rc.UpdateToplevel()
        out.append(core.unpack_hits(rc.TraceClosest(*core.pack_rays(O, D, tmin=1e-4))))
        occ = rc.TraceAny(*core.pack_rays(O, D, tmin=1e-4, tmax=80.0))
        assert np.array_equal(occ, out[-1]["t"] < np.float32(80.0))
        rc.Shutdown()
    a, b = out
    assert np.array_equal(a["t"].view(np.uint32), b["t"].view(np.uint32))
    differ = a["prim"] != b["prim"]
    print("sphere field: hits", int((a["prim"] >= 0).sum()), "of", len(O), "; same-t different-id:", int(differ.sum()))
    assert differ.mean() < 1e-4


def test_duplicate_centroids_and_coincident_triangles():
    """Hundreds of triangles with the same centroid (equal Morton codes: the radix tree falls back to the index) and exact
    copies of one triangle."""
    rng = np.random.default_rng(12)
    base = scenes.triangle_soup(3000, seed=5, extent=4.0, edge=0.6)
    star = []
    for k in range(400):  # triangles of different shapes around one centroid
        d = rng.normal(size=(3, 3)).astype(np.float32) * 0.4
        d -= d.mean(axis=0, keepdims=True)
        star.append(np.array([0.5, 0.25, -0.75], np.float32) + d)
    copies = np.repeat(base[:1], 300, axis=0)
    verts = np.concatenate([base, np.asarray(star, np.float32), copies]).astype(np.float32)
    O, D = scenes.random_rays(100_000, 5.0, seed=13)
    case = dict(name="dups", meshes=[verts], instances=[(0, None)], rays=(O, D))
    rc = _gpu_core(case)
    hits = rc.TraceClosest(*core.pack_rays(O, D))
    port = cases.load_oracle(case, "port")
    rep = parity.compare("dups", hits, port, O, D, *port.trace(O, D, nthreads=8))
    rc.Shutdown()
    assert rep["wrong"] == 0, rep  # coincident copies tie by construction, so the tie class is large here


def test_refit_after_gpu_build_and_second_mesh():
    """A GPU-built mesh refits like a host-built one, and a mesh appended behind it lands at the right pool offsets."""
    sph = scenes.uv_sphere(64, 33, 1.0)
    grid = scenes.grid_mesh(32, 32)
    M = scenes.lattice_transforms(8, spacing=3.0, rotate=False, seed=2)
    inst = [(i % 2, M[i]) for i in range(8)]
    O, D = scenes.random_rays(60_000, 6.0, seed=31)
    moved = (sph * np.float32(1.0) + np.float32(0.15) * np.sin(sph[..., ::-1] * np.float32(3.0))).astype(np.float32)
    rc = core.RenderCoreB200(0)
    _choose_builder(rc)
    cases.load_core(rc, dict(meshes=[sph, grid], instances=inst))
    rc.SetGeometry(0, moved)  # same triangle count: refit of the GPU-built tree
    rc.UpdateToplevel()
    hits = rc.TraceClosest(*core.pack_rays(O, D))
    case = dict(name="gpu_refit", meshes=[moved, grid], instances=inst, rays=(O, D))
    port = cases.load_oracle(case, "port")
    rep = parity.compare("gpu_refit", hits, port, O, D, *port.trace(O, D, nthreads=8))
    rc.Shutdown()
    parity.assert_parity(rep)


def test_gpu_built_tree_is_the_same_bytes_every_time():
    """The device collapse places children and primitives by a running sum over the level (build.cu: collapseCountKernel + scan), not by
    the order in which thread groups reach an atomic: the same mesh gives the same node and triangle bytes on every build."""
    verts = scenes.triangle_soup(300_000, seed=5)
    trees = []
    for _ in range(3):
        rc = core.RenderCoreB200(0)
        _choose_builder(rc)
        cases.load_core(rc, dict(meshes=[verts], instances=[(0, None)]))
        trees.append(rc.ReadTree())
        rc.Shutdown()
    for nodes, tris in trees[1:]:
        assert nodes.shape == trees[0][0].shape and np.array_equal(nodes, trees[0][0])
        assert np.array_equal(tris, trees[0][1])
    assert np.array_equal(np.sort(trees[0][1][:, 3]), np.arange(verts.shape[0], dtype=np.uint32))  # every primitive exactly once


def test_bvh_disk_cache_round_trip(tmp_path):
    """B200_BVH_CACHE (bvh_cache.cpp): the second core takes the host-built tree from disk and traces the
    same hits; a damaged file is ignored and rebuilt.  (The directory is read once per process: subprocesses.)"""
    import subprocess
    import sys
    script = r'''
import sys, json, numpy as np
sys.path.insert(0, %r); sys.path.insert(0, %r)
from lighthouse2_b200 import core, scenes
verts = scenes.sphere_field(60_000)
O, D = scenes.random_rays(50_000, 20.0, seed=5)
rc = core.RenderCoreB200(0)
rc.SetGeometry(0, verts); rc.SetInstance(0, 0); rc.SetInstance(1, -1); rc.UpdateToplevel()
h = rc.TraceClosest(*core.pack_rays(O, D))
st = rc.GetCoreStats()
np.save(sys.argv[1], h)
print(json.dumps(dict(hits=int(st["bvhCacheHits"]), nodes=int(st["wideNodeCount"]), build_ms=float(st["bvhBuildTime"]))))
''' % (os.path.dirname(os.path.dirname(os.path.abspath(__file__))), os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, B200_BVH_CACHE=str(tmp_path))

    def run(tag):
        r = subprocess.run([sys.executable, "-c", script, str(tmp_path / (tag + ".npy"))], env=env, capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, r.stderr
        import json
        return json.loads(r.stdout.strip().splitlines()[-1]), np.load(tmp_path / (tag + ".npy"))
    a, ha = run("first")
    files = [f for f in os.listdir(tmp_path) if f.endswith(".b2wb")]
    assert a["hits"] == 0 and len(files) == 1
    b, hb = run("second")
    assert b["hits"] == 1 and b["nodes"] == a["nodes"]
    assert np.array_equal(ha.view(np.uint32), hb.view(np.uint32))
    print("bvh cache: build", a["build_ms"], "ms, from cache", b["build_ms"], "ms")
    # flip one byte in the node payload: the checksum rejects the file, the tree is rebuilt (and the file rewritten)
    path = tmp_path / files[0]
    raw = bytearray(path.read_bytes())
    raw[len(raw) // 2] ^= 0x40
    path.write_bytes(bytes(raw))
    c, hc = run("third")
    assert c["hits"] == 0 and np.array_equal(ha.view(np.uint32), hc.view(np.uint32))
    d, _ = run("fourth")
    assert d["hits"] == 1
