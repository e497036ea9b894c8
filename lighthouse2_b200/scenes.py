"""Procedural scenes and ray sets for BASELINE.json's configs C1..C5.

Pure numpy, deterministic for a given seed.  Geometry is returned triangle-expanded the way
RenderSystem hands it to a core: ``verts[T,3,3]`` float32 (vertex0/1/2 of each triangle), which is
what ``SetGeometry(meshIdx, float4 verts[3T], 3T, T, CoreTri[T])`` carries
(lib/RenderSystem/host_mesh.cpp:525-527, host_scene.cpp:211-213).

Scene scale note: the reference triangle test rejects ``|dot(e1, cross(D, e2))| < 1e-4``
(lib/RenderCore_Bart/common.h:29), i.e. every triangle whose doubled area is below 1e-4 is invisible
and larger ones vanish at grazing angles.  The generators therefore keep triangle edges around
0.05..1 world units and the whole scene inside roughly [-100,100]^3 so that ``t < 1/EPSILON = 1e4``
(common.h:43) never bites.
"""
from __future__ import annotations

import numpy as np

DEFAULT_SEED = 0x12345678  # lib/platform/system.cpp:38 seed value used by the reference RNG


def xor32(seed: int) -> int:
    """One step of the reference RNG (lib/platform/system.cpp:45)."""
    seed &= 0xFFFFFFFF
    seed ^= (seed << 13) & 0xFFFFFFFF
    seed ^= seed >> 17
    seed ^= (seed << 5) & 0xFFFFFFFF
    return seed


def grid_mesh(nx: int = 16, nz: int = 32) -> np.ndarray:
    """C1: height-field grid over [-1,1]^2, y = 0.15 sin(3x) cos(3z); 2*nx*nz triangles (1024)."""
    xs = np.linspace(-1.0, 1.0, nx + 1, dtype=np.float64)
    zs = np.linspace(-1.0, 1.0, nz + 1, dtype=np.float64)
    X, Z = np.meshgrid(xs, zs, indexing="ij")
    Y = 0.15 * np.sin(3.0 * X) * np.cos(3.0 * Z)
    P = np.stack([X, Y, Z], axis=-1).astype(np.float32)
    a, b, c, d = P[:-1, :-1], P[1:, :-1], P[1:, 1:], P[:-1, 1:]
    t0 = np.stack([a, c, b], axis=2)  # counter-clockwise seen from +y
    t1 = np.stack([a, d, c], axis=2)
    tris = np.stack([t0, t1], axis=2).reshape(-1, 3, 3)
    return np.ascontiguousarray(tris, dtype=np.float32)


def uv_sphere(slices: int, stacks: int, radius: float = 1.0, deform=None) -> np.ndarray:
    """UV sphere with 2*slices*(stacks-1) triangles.  ``deform(theta, phi) -> radius factor``."""
    th = np.linspace(0.0, np.pi, stacks + 1)[:, None]            # polar
    ph = (np.arange(slices + 1) * (2.0 * np.pi / slices))[None, :]  # azimuth
    r = radius * (deform(th, ph) if deform is not None else 1.0)
    X = r * np.sin(th) * np.cos(ph)
    Y = r * np.cos(th) * np.ones_like(ph)
    Z = r * np.sin(th) * np.sin(ph)
    P = np.stack([X, Y, Z], axis=-1)
    P[:, -1] = P[:, 0]  # close the seam exactly
    P[0, :] = P[0, 0]
    P[-1, :] = P[-1, 0]
    a, b, c, d = P[:-1, :-1], P[1:, :-1], P[1:, 1:], P[:-1, 1:]
    upper = np.stack([a, c, b], axis=2)  # a==d at the top pole: `lower` would be degenerate there
    lower = np.stack([a, d, c], axis=2)  # b==c at the bottom pole: `upper` would be degenerate there
    top = upper[0].reshape(-1, 3, 3)
    bottom = lower[-1].reshape(-1, 3, 3)
    body_u = upper[1:-1].reshape(-1, 3, 3)
    body_l = lower[1:-1].reshape(-1, 3, 3)
    tris = np.concatenate([top, body_u, body_l, bottom], axis=0)
    return np.ascontiguousarray(tris, dtype=np.float32)


def _sphere_tessellation(n_tris: int):
    """(slices, stacks) with 2*slices*(stacks-1) == n_tris when n_tris/2 factors, slices ~ 2*stacks."""
    half = max(2, n_tris // 2)
    target = (n_tris / 4.0) ** 0.5
    best = None
    for d in range(2, half + 1):
        if half % d == 0 and half // d >= 3:
            score = abs(np.log(d / target))
            if best is None or score < best[0]:
                best = (score, d)
        if d > 4 * target and best is not None:
            break
    d = best[1] if best else 2
    return half // d, d + 1


def sphere_field(n_tris: int = 1_000_000, seed: int = DEFAULT_SEED, spacing: float = 10.0,
                 radius: float = 4.0) -> np.ndarray:
    """C2: 5x4x5 lattice of jittered UV spheres baked into ONE mesh; exactly ``n_tris`` triangles
    when n_tris is 100 * 2*s*(k-1) (1M -> 100 spheres x 10 000 tris, slices=100, stacks=51)."""
    n_spheres = 100
    per = n_tris // n_spheres
    slices, stacks = _sphere_tessellation(per)
    base = uv_sphere(slices, stacks, radius).astype(np.float64)
    rng = np.random.default_rng(seed)
    out = []
    for i in range(n_spheres):
        ix, iy, iz = i % 5, (i // 5) % 4, i // 20
        c = np.array([(ix - 2.0) * spacing, (iy - 1.5) * spacing, (iz - 2.0) * spacing])
        c = c + rng.uniform(-1.0, 1.0, 3)
        out.append((base + c).astype(np.float32))
    return np.ascontiguousarray(np.concatenate(out, axis=0))


def triangle_soup(n_tris: int = 10_000_000, seed: int = DEFAULT_SEED, extent: float = 50.0,
                  edge: float = 0.5) -> np.ndarray:
    """C3: centres uniform in [-extent,extent]^3, edge vectors uniform in [-edge,edge]^3."""
    rng = np.random.default_rng(seed)
    c = rng.uniform(-extent, extent, (n_tris, 1, 3)).astype(np.float32)
    e = rng.uniform(-edge, edge, (n_tris, 2, 3)).astype(np.float32)
    tris = np.empty((n_tris, 3, 3), np.float32)
    tris[:, 0] = c[:, 0]
    tris[:, 1] = c[:, 0] + e[:, 0]
    tris[:, 2] = c[:, 0] + e[:, 1]
    return tris


def deforming_sphere(n_tris: int = 100_000, time: float = 0.0, radius: float = 3.0, k: float = 6.0) -> np.ndarray:
    """C4 BLAS: sphere with r * (1 + 0.05 sin(k*theta + time)); topology is time-independent."""
    slices, stacks = _sphere_tessellation(n_tris)
    return uv_sphere(slices, stacks, radius, deform=lambda th, ph: 1.0 + 0.05 * np.sin(k * th + time) * np.ones_like(ph))


def lattice_transforms(n: int = 1000, spacing: float = 8.0, time: float = 0.0, rotate: bool = False,
                       seed: int = DEFAULT_SEED, scale=None) -> np.ndarray:
    """C4 instances: n translation-only (bit-exact parity class) or randomly rotated transforms,
    row-major mat4 with translation in cells 3,7,11 (common_types.h:473,486).  scale = (lo, hi): every instance
    additionally gets a random non-uniform scale per axis (applied before the rotation)."""
    side = int(round(n ** (1.0 / 3.0)))
    while side ** 3 < n:
        side += 1
    rng = np.random.default_rng(seed)
    M = np.tile(np.eye(4, dtype=np.float32), (n, 1, 1))
    idx = np.arange(n)
    ix, iy, iz = idx % side, (idx // side) % side, idx // (side * side)
    off = (side - 1) / 2.0
    T = np.stack([(ix - off) * spacing, (iy - off) * spacing, (iz - off) * spacing], axis=1)
    T = T + 0.5 * np.sin(time + idx[:, None] * np.array([0.37, 0.61, 0.83]))
    M[:, :3, 3] = T.astype(np.float32)
    if rotate:
        ax = rng.normal(size=(n, 3))
        ax /= np.linalg.norm(ax, axis=1, keepdims=True)
        ang = rng.uniform(0, 2 * np.pi, n) + time
        c, s = np.cos(ang), np.sin(ang)
        x, y, z = ax[:, 0], ax[:, 1], ax[:, 2]
        R = np.empty((n, 3, 3))
        R[:, 0, 0] = c + x * x * (1 - c); R[:, 0, 1] = x * y * (1 - c) - z * s; R[:, 0, 2] = x * z * (1 - c) + y * s
        R[:, 1, 0] = y * x * (1 - c) + z * s; R[:, 1, 1] = c + y * y * (1 - c); R[:, 1, 2] = y * z * (1 - c) - x * s
        R[:, 2, 0] = z * x * (1 - c) - y * s; R[:, 2, 1] = z * y * (1 - c) + x * s; R[:, 2, 2] = c + z * z * (1 - c)
        M[:, :3, :3] = R.astype(np.float32)
    if scale is not None:
        S = rng.uniform(scale[0], scale[1], (n, 3)).astype(np.float32)
        M[:, :3, :3] = M[:, :3, :3] * S[:, None, :]
    return np.ascontiguousarray(M.reshape(n, 16))


# --------------------------------------------------------------------------------------------------
# cameras and ray sets
# --------------------------------------------------------------------------------------------------

def _normalize_ref(v: np.ndarray) -> np.ndarray:
    """Ray ctor normalisation in float32: v * (1.0f / sqrtf(x*x + y*y + z*z))
    (lib/RenderCore_Bart/common.h:9-11, common_types.h:87,349,363).  numpy float32 ops round once per
    operation like the reference's unfused scalar code."""
    v = v.astype(np.float32)
    d = v[..., 0] * v[..., 0]
    d = d + v[..., 1] * v[..., 1]
    d = d + v[..., 2] * v[..., 2]
    inv = np.float32(1.0) / np.sqrt(d, dtype=np.float32)
    return (v * inv[..., None]).astype(np.float32)


def view_pyramid(pos, target, fov_deg: float, width: int, height: int, focal_distance: float = 1.0):
    """pos/p1/p2/p3 as Camera::GetView builds them (lib/RenderSystem/camera.cpp:96-117): p1 = top-left,
    p2 = top-right, p3 = bottom-left of the virtual screen at ``focal_distance``."""
    pos = np.asarray(pos, np.float64)
    fwd = np.asarray(target, np.float64) - pos
    fwd /= np.linalg.norm(fwd)
    up0 = np.array([0.0, 1.0, 0.0])
    right = np.cross(up0, fwd)
    right /= np.linalg.norm(right)
    up = np.cross(fwd, right)
    aspect = width / height
    s = np.tan(np.radians(fov_deg) / 2.0)
    C = pos + focal_distance * fwd
    p1 = C - s * right * focal_distance * aspect + s * focal_distance * up
    p2 = C + s * right * focal_distance * aspect + s * focal_distance * up
    p3 = C - s * right * focal_distance * aspect - s * focal_distance * up
    f = lambda a: np.asarray(a, np.float32)
    return {"pos": f(pos), "p1": f(p1), "p2": f(p2), "p3": f(p3)}


def primary_rays(view, width: int, height: int, offset=(0.5, 0.5)):
    """Pinhole primary rays through pixel centres, P = p1 + u (p2-p1) + v (p3-p1), D = P - pos,
    normalised as the Ray ctor does.  Returns (O[n,3], D[n,3]) float32, row-major pixel order.
    (GPU cores: u=(x+r0)/w, v=(y+r1)/h, rendercore_optix7/optix/.optix.cu:66-123; jitter is fixed here
    so CPU and GPU consume the identical array.)"""
    f32 = np.float32
    u = ((np.arange(width, dtype=f32) + f32(offset[0])) / f32(width)).astype(f32)
    v = ((np.arange(height, dtype=f32) + f32(offset[1])) / f32(height)).astype(f32)
    right = (view["p2"] - view["p1"]).astype(f32)
    up = (view["p3"] - view["p1"]).astype(f32)
    P = view["p1"][None, None, :] + u[None, :, None] * right[None, None, :] + v[:, None, None] * up[None, None, :]
    D = _normalize_ref((P - view["pos"][None, None, :]).reshape(-1, 3))
    O = np.broadcast_to(view["pos"], D.shape).astype(f32).copy()
    return np.ascontiguousarray(O), np.ascontiguousarray(D)


def random_rays(n: int, extent: float, seed: int = DEFAULT_SEED):
    """Incoherent test rays: origins uniform in the box, directions uniform on the sphere."""
    rng = np.random.default_rng(seed)
    O = rng.uniform(-extent, extent, (n, 3)).astype(np.float32)
    d = rng.normal(size=(n, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    return O, _normalize_ref(d.astype(np.float32))


C1_CAMERA = dict(pos=(0.0, 1.5, -3.0), target=(0.0, 0.0, 0.0), fov_deg=40.0)
C2_CAMERA = dict(pos=(6.0, 9.0, -62.0), target=(0.0, 0.0, 0.0), fov_deg=45.0)
C3_CAMERA = dict(pos=(0.0, 0.0, -140.0), target=(0.0, 0.0, 0.0), fov_deg=40.0)
C4_CAMERA = dict(pos=(10.0, 14.0, -95.0), target=(0.0, 0.0, 0.0), fov_deg=45.0)
C2_LIGHT = (0.0, 60.0, -20.0)
