"""Seeded parity cases shared by the oracle tests (CPU) and the GPU parity tests.

A case = meshes (verts[T,3,3]), instances [(mesh, mat16 or None)], rays (O[n,3], D[n,3]).
The same arrays go to the CPU oracle and through the C ABI to the CUDA kernels.
"""
import numpy as np

from lighthouse2_b200 import scenes


def case_c1(width=640, height=360):
    """BASELINE.json configs[0]: 1K-triangle procedural mesh, 640x360 primary rays."""
    view = scenes.view_pyramid(width=width, height=height, **scenes.C1_CAMERA)
    O, D = scenes.primary_rays(view, width, height)
    return dict(name="c1", meshes=[scenes.grid_mesh()], instances=[(0, None)], rays=(O, D), view=view, size=(width, height))


def case_soup(n_tris=2000, n_rays=20000, seed=11):
    soup = scenes.triangle_soup(n_tris, seed=seed, extent=5.0, edge=0.5)
    O, D = scenes.random_rays(n_rays, 6.0, seed=seed + 1)
    return dict(name=f"soup{n_tris}", meshes=[soup], instances=[(0, None)], rays=(O, D))


def case_instanced(n_inst=27, n_rays=20000, rotate=False, seed=5, scale=None):
    sph = scenes.uv_sphere(20, 11, 1.0)
    grid = scenes.grid_mesh(8, 8)
    M = scenes.lattice_transforms(n_inst, spacing=3.0, rotate=rotate, seed=seed, scale=scale)
    inst = [(i % 2, M[i]) for i in range(n_inst)]
    O, D = scenes.random_rays(n_rays, 6.0, seed=seed + 2)
    return dict(name="inst%d%s%s" % (n_inst, "rot" if rotate else "", "scaled" if scale else ""), meshes=[sph, grid], instances=inst, rays=(O, D))


def case_c4_class(n_inst=1000, n_tris=100_000, time=0.0, rotate=False):
    """BASELINE.json configs[3] at size: 1000 instances of one 100 000-triangle deforming sphere on a 10 x 10 x 10 lattice
    (translation-only = the bit-exact parity class; rotate = the second, rotated run of SURVEY.md 8d C4)."""
    mesh = scenes.deforming_sphere(n_tris, time=time)
    M = scenes.lattice_transforms(n_inst, spacing=8.0, time=time, rotate=rotate)
    return dict(name="c4", meshes=[mesh], instances=[(0, M[i]) for i in range(n_inst)])


def case_axis_aligned(n_rays=20000, seed=3):
    """Edge cases: flat axis-aligned quads (zero-thickness boxes), axis-parallel rays (1/0 = inf in the
    reference slab test, bvh.cpp:13-15), rays starting on a surface, degenerate (zero-area) triangles."""
    q = []
    for y in (0.0, 1.0, 2.0):
        a, b, c, d = [-2, y, -2], [2, y, -2], [2, y, 2], [-2, y, 2]
        q += [[a, c, b], [a, d, c]]
    q += [[[0, 0, 0], [0, 0, 0], [1, 1, 1]], [[1, 1, 1], [1, 1, 1], [1, 1, 1]]]  # degenerate
    for x in (-1.0, 1.0):
        a, b, c, d = [x, -1, -2], [x, 3, -2], [x, 3, 2], [x, -1, 2]
        q += [[a, b, c], [a, c, d]]
    tris = np.asarray(q, np.float32)
    rng = np.random.default_rng(seed)
    O = rng.uniform(-3, 3, (n_rays, 3)).astype(np.float32)
    D = np.zeros((n_rays, 3), np.float32)
    axis = rng.integers(0, 3, n_rays)
    sign = rng.choice([-1.0, 1.0], n_rays).astype(np.float32)
    D[np.arange(n_rays), axis] = sign
    # a quarter of the rays: random directions; another slice: origins exactly on the y=1 plane
    k = n_rays // 4
    d = rng.normal(size=(k, 3))
    D[:k] = scenes._normalize_ref((d / np.linalg.norm(d, axis=1, keepdims=True)).astype(np.float32))
    O[k:2 * k, 1] = 1.0
    return dict(name="axis", meshes=[tris], instances=[(0, None)], rays=(O, D))


def case_empty():
    O, D = scenes.random_rays(64, 2.0, seed=9)
    return dict(name="empty", meshes=[np.zeros((0, 3, 3), np.float32)], instances=[(0, None)], rays=(O, D))


def case_single_triangle():
    tri = np.asarray([[[-1, -1, 2], [1, -1, 2], [0, 1, 2]]], np.float32)
    O, D = scenes.random_rays(4096, 1.5, seed=10)
    O[:, 2] = -1.0
    D = scenes._normalize_ref((np.array([0, 0, 2], np.float32) + 0.9 * np.random.default_rng(2).uniform(-1, 1, (4096, 3)).astype(np.float32) - O))
    return dict(name="tri1", meshes=[tri], instances=[(0, None)], rays=(O, D))


SMALL_CASES = [case_soup, case_instanced, case_axis_aligned, case_single_triangle]


def load_oracle(case, kind="port"):
    from oracle import Oracle
    o = Oracle(kind)
    for i, m in enumerate(case["meshes"]):
        o.SetGeometry(i, m)
    for i, (mesh, mat) in enumerate(case["instances"]):
        o.SetInstance(i, mesh, mat)
    o.UpdateToplevel()
    return o


def load_core(core, case):
    for i, m in enumerate(case["meshes"]):
        core.SetGeometry(i, m)
    for i, (mesh, mat) in enumerate(case["instances"]):
        core.SetInstance(i, mesh, mat)
    core.SetInstance(len(case["instances"]), -1)
    core.UpdateToplevel()
    return core
