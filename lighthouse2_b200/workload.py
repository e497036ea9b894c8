"""Ray-set generator for the wave benchmark and the large parity tests (host side, numpy).

One "pass" over a scene is the sequence of waves a wavefront path tracer issues
(lib/rendercore_optix7/rendercore.cpp:692-748): a primary extend wave, one shadow (any-hit) wave towards a
point light from every primary hit, then ``bounces`` diffuse extend waves whose rays leave the previous
wave's hit points in cosine-weighted directions about the geometric normal
(DiffuseReflectionCosWeighted + Tangent2World, lib/CUDA/shared_kernel_code/tools_shared.h:250-256,211-220;
origin offset in the spirit of SafeOrigin, :279-292).

The generator is tracer-agnostic: ``tracer(O, D) -> (t, prim)`` is either the CUDA core (through the C ABI)
or, in bench.py's CPU legs, the CPU checker.  It produces the wire format of include/rendercore_b200.h:
float4 planes O4 = (O, tmin), D4 = (D, tmax).
"""
from __future__ import annotations

import numpy as np

from . import scenes

EPS = np.float32(1e-4)       # geometryEpsilon of the GPU cores (rendercore_optix7/rendercore.cpp:597-615)
MISS_T = np.float32(1e34)


def geometric_normals(verts: np.ndarray, prim: np.ndarray, D: np.ndarray) -> np.ndarray:
    """Unit geometric normal of triangle ``prim`` facing against the incoming direction D."""
    tri = verts[prim]
    n = np.cross(tri[:, 1] - tri[:, 0], tri[:, 2] - tri[:, 0]).astype(np.float64)
    ln = np.linalg.norm(n, axis=1, keepdims=True)
    n = n / np.maximum(ln, 1e-30)
    flip = np.einsum("ij,ij->i", n, D.astype(np.float64)) > 0
    n[flip] = -n[flip]
    return n.astype(np.float32)


def cosine_directions(N: np.ndarray, rng: np.random.Generator) -> np.ndarray:
    """Cosine-weighted hemisphere sample about N (tools_shared.h:250-256 + Tangent2World :211-220)."""
    n = N.shape[0]
    r0 = rng.random(n, dtype=np.float32)
    r1 = rng.random(n, dtype=np.float32)
    term1 = np.float32(2 * np.pi) * r0
    term2 = np.sqrt(np.float32(1) - r1)
    s, c = np.sin(term1), np.cos(term1)
    loc = np.stack([c * term2, s * term2, np.sqrt(r1)], axis=1).astype(np.float32)
    # tangent frame: T = normalize(cross(N, |Nx| > 0.99 ? (0,1,0) : (1,0,0))), B = cross(T, N)
    a = np.zeros_like(N)
    big = np.abs(N[:, 0]) > 0.99
    a[big, 1] = 1
    a[~big, 0] = 1
    T = np.cross(N, a)
    T /= np.maximum(np.linalg.norm(T, axis=1, keepdims=True), 1e-30)
    B = np.cross(T, N)
    d = loc[:, 0:1] * T + loc[:, 1:2] * B + loc[:, 2:3] * N
    return scenes._normalize_ref(d.astype(np.float32))


def pack(O, D, tmin, tmax):
    n = O.shape[0]
    O4 = np.empty((n, 4), np.float32)
    D4 = np.empty((n, 4), np.float32)
    O4[:, :3], D4[:, :3] = O, D
    O4[:, 3], D4[:, 3] = tmin, tmax
    return O4, D4


def make_pass(tracer, verts: np.ndarray, O: np.ndarray, D: np.ndarray, light, bounces: int = 4, seed: int = 1):
    """Returns a list of waves ``dict(cls='primary'|'shadow'|'diffuse', kind='closest'|'any', O4, D4, bounce)``.
    ``verts`` is the world-space triangle list of the (single-mesh, identity-instance) scene the rays hit."""
    rng = np.random.default_rng(seed)
    waves = []
    O4, D4 = pack(O, D, EPS, MISS_T)
    waves.append(dict(cls="primary", kind="closest", O4=O4, D4=D4, bounce=0))
    t, prim = tracer(O, D)
    curO, curD = O, D
    light = np.asarray(light, np.float32)
    for b in range(bounces + 1):
        hit = prim >= 0
        if not hit.any():
            break
        Oh, Dh, th, ph = curO[hit], curD[hit], t[hit], prim[hit]
        P = (Oh + Dh * th[:, None]).astype(np.float32)
        N = geometric_normals(verts, ph, Dh)
        Ps = (P + N * np.float32(2e-3)).astype(np.float32)  # leave the surface (scene scale ~ 100 units)
        if b == 0:
            L = light[None, :] - Ps
            dist = np.linalg.norm(L, axis=1).astype(np.float32)
            Ld = scenes._normalize_ref(L.astype(np.float32))
            front = np.einsum("ij,ij->i", Ld, N) > 0  # NEE only for surfaces facing the light (pathtracer.h:170-175)
            So4, Sd4 = pack(Ps[front], Ld[front], EPS, dist[front] - 2 * EPS)
            waves.append(dict(cls="shadow", kind="any", O4=So4, D4=Sd4, bounce=0))
        if b == bounces:
            break
        nd = cosine_directions(N, rng)
        curO, curD = Ps, nd
        O4, D4 = pack(curO, curD, EPS, MISS_T)
        waves.append(dict(cls="diffuse", kind="closest", O4=O4, D4=D4, bounce=b + 1))
        t, prim = tracer(curO, curD)
    return waves


def pass_ray_counts(waves):
    out = {}
    for w in waves:
        out[w["cls"]] = out.get(w["cls"], 0) + int(w["O4"].shape[0])
    return out
