"""CPU restatement of RenderCore_B200's wavefront frame (lighthouse2_b200/csrc/wavefront.cu) -- TEST INFRASTRUCTURE.

Same loop, same per-path random streams (WangHash + xor32, lib/CUDA/shared_kernel_code/tools_shared.h:60-62), but
every extend / shadow query goes through the CPU oracle (the reference's TopLevelBVH::Traverse), as SURVEY.md
Appendix A.9 prescribes for image-level parity.  numpy float32 throughout; transcendental functions differ from the
GPU's fast-math versions in the last bits, so images agree to an RMSE, not bit for bit.
"""
import numpy as np

f32 = np.float32
EPS = f32(1e-4)


def wang_hash(s):
    s = s.astype(np.uint32)
    s = (s ^ np.uint32(61)) ^ (s >> np.uint32(16))
    s = s * np.uint32(9)
    s = s ^ (s >> np.uint32(4))
    s = s * np.uint32(0x27d4eb2d)
    s = s ^ (s >> np.uint32(15))
    return s


def random_float(seed):
    seed ^= seed << np.uint32(13)
    seed ^= seed >> np.uint32(17)
    seed ^= seed << np.uint32(5)
    return seed, (seed.astype(f32) * f32(2.3283064365387e-10)).astype(f32)


def safe_origin(I, R, N, eps):
    parallel = f32(1) - np.abs(np.einsum("ij,ij->i", N, R)).astype(f32)
    v = (parallel * parallel)[:, None]
    return (I + R * eps * (1 - v) + N * eps * v).astype(f32)


def tangent2world(V, N):
    sign = np.copysign(f32(1), N[:, 2])
    a = f32(-1) / (sign + N[:, 2])
    b = N[:, 0] * N[:, 1] * a
    B = np.stack([1 + sign * N[:, 0] * N[:, 0] * a, sign * b, -sign * N[:, 0]], axis=1)
    T = np.stack([b, sign + N[:, 1] * N[:, 1] * a, -N[:, 1]], axis=1)
    return (V[:, 0:1] * T + V[:, 1:2] * B + V[:, 2:3] * N).astype(f32)


def render(oracle, meshes, instances, view, width, height, light=None, sky=(0, 0, 0), materials=None, max_path_length=5,
           spp=1, sample_offset=0, nthreads=8):
    """Returns the accumulator float32[h, w, 3] (sum over samples).  meshes: list of verts[T,3,3]; instances: list of
    (mesh, mat16 or None); light = (pos3, radiance3) or None.  Pixel-centre rays (jitter off)."""
    npx = width * height
    acc = np.zeros((npx, 3), np.float64)
    pos, p1 = view["pos"].astype(f32), view["p1"].astype(f32)
    right, up = (view["p2"] - view["p1"]).astype(f32), (view["p3"] - view["p1"]).astype(f32)
    normals = []
    for v in meshes:
        n = np.cross(v[:, 1] - v[:, 0], v[:, 2] - v[:, 0]).astype(f32)
        l2 = np.einsum("ij,ij->i", n, n)
        il = np.where(l2 > 0, f32(1) / np.sqrt(np.maximum(l2, 1e-38)), 0).astype(f32)
        normals.append(n * il[:, None])
    mats = [np.eye(4, dtype=f32) if m is None else np.asarray(m, f32).reshape(4, 4) for _, m in instances]
    inst_mesh = [mesh for mesh, _ in instances]
    sky = np.asarray(sky, f32)
    for s in range(spp):
        sample = sample_offset + s
        R0 = np.uint32(7907)
        pix = np.arange(npx, dtype=np.uint32)
        path_idx = (pix + np.uint32(sample) * np.uint32(npx)).astype(np.uint32)
        x, y = (pix % width).astype(f32), (pix // width).astype(f32)
        u, v = (x + f32(0.5)) / f32(width), (y + f32(0.5)) / f32(height)
        D = (p1[None] + u[:, None] * right[None] + v[:, None] * up[None] - pos[None]).astype(f32)
        D = (D * (f32(1) / np.sqrt(np.einsum("ij,ij->i", D, D)))[:, None]).astype(f32)
        O = np.broadcast_to(pos, D.shape).astype(f32).copy()
        T = np.ones((npx, 3), f32)
        for length in range(1, max_path_length + 1):
            if len(O) == 0:
                break
            t, prim, inst = oracle.trace(O, D, nthreads=nthreads)
            px = (path_idx % np.uint32(npx)).astype(np.int64)
            miss = prim < 0
            np.add.at(acc, px[miss], (T[miss] * sky[None]).astype(np.float64))
            h = ~miss
            if not h.any():
                break
            O, D, T, t, prim, inst, path_idx, px = O[h], D[h], T[h], t[h], prim[h], inst[h], path_idx[h], px[h]
            N = np.empty_like(O)
            color = np.full_like(O, f32(0.8))
            for i in np.unique(inst):
                sel = inst == i
                n_obj = normals[inst_mesh[i]][prim[sel]]
                n_w = (n_obj @ mats[i][:3, :3].T).astype(f32)
                N[sel] = n_w * (f32(1) / np.sqrt(np.einsum("ij,ij->i", n_w, n_w)))[:, None]
            if materials is not None:
                raise NotImplementedError
            flip = np.einsum("ij,ij->i", D, N) > 0
            N[flip] = -N[flip]
            I = (O + t[:, None] * D).astype(f32)
            if light is not None:
                lp, lr = np.asarray(light[0], f32), np.asarray(light[1], f32)
                L = (lp[None] - I).astype(f32)
                d2 = np.einsum("ij,ij->i", L, L).astype(f32)
                dist = np.sqrt(d2)
                L = (L * (f32(1) / dist)[:, None]).astype(f32)
                ndl = np.einsum("ij,ij->i", L, N).astype(f32)
                ok = (ndl > 0) & (dist > 4 * EPS)
                if ok.any():
                    sc = (ndl[ok] * f32(0.31830988618379067154) / d2[ok]).astype(f32)
                    E = (T[ok] * color[ok] * lr[None] * sc[:, None]).astype(f32)
                    So = safe_origin(I[ok], L[ok], N[ok], EPS)
                    tmax = dist[ok] - 2 * EPS
                    ts, _, _ = oracle.trace(So, L[ok], nthreads=nthreads)
                    vis = ~(ts < tmax)
                    np.add.at(acc, px[ok][vis], E[vis].astype(np.float64))
            if length == max_path_length:
                break
            seed = wang_hash(path_idx * np.uint32(17) + R0 + np.uint32(length) * np.uint32(91771))
            seed, r0 = random_float(seed)
            seed, r1 = random_float(seed)
            term1, term2 = f32(6.28318530717958647692) * r0, np.sqrt(f32(1) - r1)
            V = np.stack([np.cos(term1) * term2, np.sin(term1) * term2, np.sqrt(r1)], axis=1).astype(f32)
            R = tangent2world(V, N)
            R = (R * (f32(1) / np.sqrt(np.einsum("ij,ij->i", R, R)))[:, None]).astype(f32)
            O = safe_origin(I, R, N, EPS)
            D = R
            T = (T * color).astype(f32)
    return acc.reshape(height, width, 3).astype(f32)
