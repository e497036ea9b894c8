"""CPU restatement of RenderCore_B200's wavefront frame (lighthouse2_b200/csrc/wavefront.cu) -- TEST INFRASTRUCTURE.

Same loop, same per-path random streams (WangHash + xor32, lib/CUDA/shared_kernel_code/tools_shared.h:60-62), but
every extend / shadow query goes through the CPU oracle (the reference's TopLevelBVH::Traverse), as SURVEY.md
Appendix A.9 prescribes for image-level parity.  numpy float32 throughout; transcendental functions differ from the
GPU's fast-math versions in the last bits, so images agree to an RMSE, not bit for bit.

The shade stage restates lib/rendercore_optix7/kernels/pathtracer.h:54-240 with the Lambert BSDF of
lib/sharedBSDFs/lambert.h (no transmission) and the light sampling of lib/CUDA/shared_kernel_code/lights_shared.h, with
the same simplifications as the CUDA core (geometric normals, no textures, WangHash/xor32 instead of blue noise).
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


def sincos_det(x):
    """wavefront.cu::SinCosDet operation by operation (float32, every operation rounded on its own): the device frames and this
    restatement then agree on every bounce direction to the bit, where the hardware __sincosf of the reference cannot be restated."""
    x = np.asarray(x, f32)
    kf = np.rint(x * f32(0.636619772)).astype(f32)
    k = kf.astype(np.int64)
    r = (x - kf * f32(1.5707963705062866)).astype(f32)
    r = (r - kf * f32(-4.371139000186243e-8)).astype(f32)
    r2 = (r * r).astype(f32)
    sp = (r + (r * r2) * (f32(-0.16666667) + r2 * (f32(0.0083333310) + r2 * (f32(-0.00019840874) + r2 * f32(2.7525562e-6))))).astype(f32)
    cp = (f32(1.0) + r2 * (f32(-0.5) + r2 * (f32(0.041666638) + r2 * (f32(-0.0013888378) + r2 * f32(2.4760495e-5))))).astype(f32)
    odd = (k & 1) != 0
    ss, cc = np.where(odd, cp, sp), np.where(odd, sp, cp)
    return np.where((k & 2) != 0, -ss, ss).astype(f32), np.where(((k + 1) & 2) != 0, -cc, cc).astype(f32)


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


INVPI = f32(0.31830988618379067154)
TWOPI = f32(6.28318530717958647692)
S_SPECULAR, S_BOUNCED, S_VIASPECULAR, S_BOUNCEDTWICE = 1, 2, 4, 8


def dot(a, b):
    return (a[:, 0] * b[:, 0] + a[:, 1] * b[:, 1] + a[:, 2] * b[:, 2]).astype(f32)


def normalize(a):
    return (a * (f32(1) / np.sqrt(dot(a, a)))[:, None]).astype(f32)


class Lights:
    """The four light arrays of SetLights as float32 records (lighthouse2_b200/lights.py)."""

    def __init__(self, area=None, point=None, spot=None, directional=None):
        self.area = np.zeros((0, 24), f32) if area is None else np.asarray(area, f32).reshape(-1, 24)
        self.point = np.zeros((0, 8), f32) if point is None else np.asarray(point, f32).reshape(-1, 8)
        self.spot = np.zeros((0, 12), f32) if spot is None else np.asarray(spot, f32).reshape(-1, 12)
        self.dir = np.zeros((0, 8), f32) if directional is None else np.asarray(directional, f32).reshape(-1, 8)

    @property
    def count(self):
        return len(self.area) + len(self.point) + len(self.spot) + len(self.dir)

    def potentials(self, I, N, P=None, bary=None):
        """[n, lights] potential contributions (lights_shared.h:33-94).  bary given: area lights sampled at the
        barycentric point; else (implicit connection) seen at point P."""
        cols = []
        with np.errstate(all="ignore"):
            for a in self.area:
                if bary is not None:
                    L = (bary[:, 0:1] * a[12:15][None] + bary[:, 1:2] * a[16:19][None] + bary[:, 2:3] * a[20:23][None]).astype(f32)
                else:
                    L = P
                L = (L - I).astype(f32)
                att = f32(1) / dot(L, L)
                L = normalize(L)
                ln = np.broadcast_to(a[4:7], L.shape)
                cols.append(a[3] * np.maximum(f32(0), -dot(ln, L)) * np.maximum(f32(0), dot(N, L)) * att)
            for p in self.point:
                L = (p[0:3][None] - I).astype(f32)
                cols.append(p[3] * np.maximum(f32(0), dot(N, L)) * (f32(1) / dot(L, L)))
            for s in self.spot:
                L = (s[0:3][None] - I).astype(f32)
                att = f32(1) / dot(L, L)
                L = normalize(L)
                d = (np.maximum(f32(0), -dot(L, np.broadcast_to(s[8:11], L.shape))) - s[7]) / (s[3] - s[7])
                cols.append((s[4] + s[5] + s[6]) * np.maximum(f32(0), np.minimum(f32(1), d)) * np.maximum(f32(0), dot(N, L)) * att)
            for d in self.dir:
                cols.append(d[3] * np.maximum(f32(0), -dot(np.broadcast_to(d[0:3], N.shape), N)))
        return np.stack(cols, axis=1).astype(f32) if cols else np.zeros((len(I), 0), f32)


def random_barycentrics(r0):
    """lights_shared.h:122-142."""
    uf = np.minimum(r0.astype(np.float64) * 4294967296.0, 4294967295.0).astype(np.uint32)  # float -> uint saturates on the GPU
    n = len(r0)
    A = np.tile(np.array([1, 0], f32), (n, 1)); B = np.tile(np.array([0, 1], f32), (n, 1)); C = np.zeros((n, 2), f32)
    h = f32(0.5)
    for i in range(16):
        d = ((uf >> np.uint32(2 * (15 - i))) & np.uint32(3))[:, None]
        An = np.where(d == 0, (B + C) * h, np.where(d == 1, A, np.where(d == 2, (B + A) * h, (C + A) * h)))
        Bn = np.where(d == 0, (A + C) * h, np.where(d == 1, (A + B) * h, np.where(d == 2, B, (C + B) * h)))
        Cn = np.where(d == 0, (A + B) * h, np.where(d == 1, (A + C) * h, np.where(d == 2, (B + C) * h, C)))
        A, B, C = An.astype(f32), Bn.astype(f32), Cn.astype(f32)
    r = ((A + B + C) * f32(0.3333333)).astype(f32)
    return np.stack([r[:, 0], r[:, 1], f32(1) - r[:, 0] - r[:, 1]], axis=1).astype(f32)


def random_point_on_light(lights, r0, r1, I, N):
    """lights_shared.h:144-234.  Returns P, pickProb, lightPdf, lightColor."""
    n = len(I)
    bary = random_barycentrics(r0)
    pot = lights.potentials(I, N, bary=bary)
    total = np.zeros(n, f32)
    run = []
    for k in range(pot.shape[1]):
        total = (total + pot[:, k]).astype(f32)
        run.append(total)
    run = np.stack(run, axis=1)
    s = run[:, -1]
    ok = s > 0
    r1s = (r1 * s).astype(f32)
    ge = run >= r1s[:, None]
    k = np.where(ge.any(axis=1), ge.argmax(axis=1), 0)
    with np.errstate(all="ignore"):
        pick = (pot[np.arange(n), k] / s).astype(f32)
    P = np.ones((n, 3), f32)
    pdf = np.zeros(n, f32)
    col = np.zeros((n, 3), f32)
    na, npt, nsp = len(lights.area), len(lights.point), len(lights.spot)
    with np.errstate(all="ignore"):
        for i in range(lights.count):
            sel = ok & (k == i)
            if not sel.any():
                continue
            Is, Ns = I[sel], N[sel]
            if i < na:
                a = lights.area[i]
                b = bary[sel]
                Pp = (b[:, 0:1] * a[12:15][None] + b[:, 1:2] * a[16:19][None] + b[:, 2:3] * a[20:23][None]).astype(f32)
                L = (Is - Pp).astype(f32)
                sq = dot(L, L)
                L = normalize(L)
                lndl = dot(L, np.broadcast_to(a[4:7], L.shape))
                P[sel], col[sel] = Pp, a[8:11]
                pdf[sel] = np.where((lndl > 0) & (dot(L, Ns) < 0), sq / (a[7] * lndl), f32(0))
            elif i < na + npt:
                p = lights.point[i - na]
                L = (Is - p[0:3][None]).astype(f32)
                P[sel], col[sel] = p[0:3], p[4:7]
                pdf[sel] = np.where(dot(L, Ns) < 0, dot(L, L), f32(0))
            elif i < na + npt + nsp:
                sp = lights.spot[i - na - npt]
                L = (Is - sp[0:3][None]).astype(f32)
                sq = dot(L, L)
                L = normalize(L)
                d = (np.maximum(f32(0), dot(L, np.broadcast_to(sp[8:11], L.shape))) - sp[7]) / (sp[3] - sp[7])
                lndl = np.minimum(f32(1), d)
                P[sel], col[sel] = sp[0:3], sp[4:7]
                pdf[sel] = np.where((lndl > 0) & (dot(L, Ns) < 0), sq / lndl, f32(0))
            else:
                dl = lights.dir[i - na - npt - nsp]
                P[sel], col[sel] = (Is - dl[0:3][None] * f32(1000)).astype(f32), dl[4:7]
                pdf[sel] = np.where(dot(np.broadcast_to(dl[0:3], Ns.shape), Ns) < 0, f32(1), f32(0))
    pick = np.where(ok, pick, f32(0)).astype(f32)
    return P, pick, pdf.astype(f32), col


def lambert_sample(color, absorption, rough, transmission, eta, fN, gF, face, D, t, r3, r4):
    """SampleBSDF of lib/sharedBSDFs/lambert.h:71-123 over arrays, as the shade stage calls it (pathtracer.h:219).  fN / gF: shading and
    geometric normal on the side the ray came from (iN * flip, N * flip of the reference), face: +1 outside, -1 inside the medium, D: ray
    direction (wo = -D), t: distance travelled (Beer absorption).  Returns R, pdf, specular, bsdf value, dot(fN, R).  One deviation,
    shared with the device code: on total internal reflection the reference returns a zero value with pdf 1 (the path lives on with a zero
    throughput); here the pdf is 0 and the path ends.  Diffuse directions come from sincos_det, the reference's from __sincosf."""
    p_reflect = (f32(1) - rough).astype(f32)
    ddotn = dot(D, fN)
    R_refl = (D - fN * (f32(2) * ddotn)[:, None]).astype(f32)
    trans = r4 < transmission
    with np.errstate(all="ignore"):
        # dielectric branch (lambert.h:84-103)
        eio = np.where(face < 0, f32(1) / eta, eta).astype(f32)
        c = -ddotn
        e2 = np.where(c < 0, f32(1) / eio, eio).astype(f32)
        c = np.abs(c)
        sinT2 = (e2 * e2 * (f32(1) - c * c)).astype(f32)
        ldn = np.sqrt(np.maximum(f32(0), f32(1) - sinT2)).astype(f32)
        ra, rb = (c - e2 * ldn) / (c + e2 * ldn), (ldn - e2 * c) / (ldn + e2 * c)
        F = np.where(sinT2 <= 1, f32(0.5) * (ra * ra + rb * rb), f32(1)).astype(f32)
        beer = np.exp(-absorption * (t * f32(2))[:, None]).astype(f32)
        cosI = np.abs(ddotn)
        sin2T = (eio * eio * np.maximum(f32(0), f32(1) - cosI * cosI)).astype(f32)
        tir = sin2T >= 1
        R_refr = (D * eio[:, None] + fN * (eio * cosI - np.sqrt(np.maximum(f32(0), f32(1) - sin2T)))[:, None]).astype(f32)
        reflect_t = trans & (r3 < F)
        refract_t = trans & ~(r3 < F) & ~tir
        dead_tir = trans & ~(r3 < F) & tir
        # opaque branch
        spec_o = ~trans & (r3 < p_reflect)
        r5 = ((r3 - p_reflect) / (f32(1) - p_reflect)).astype(f32)
        r6 = ((r4 - transmission) / (f32(1) - transmission)).astype(f32)
        term1, term2 = TWOPI * r5, np.sqrt(f32(1) - r6)
        sn_, cs_ = sincos_det(term1)
        V = np.stack([cs_ * term2, sn_ * term2, np.sqrt(r6)], axis=1).astype(f32)
        R_diff = normalize(tangent2world(V, fN))
        diffuse = ~trans & ~spec_o
        R = np.where((reflect_t | spec_o)[:, None], R_refl, np.where(refract_t[:, None], R_refr, R_diff)).astype(f32)
        cosr = dot(fN, R)
        inv_cos = (f32(1) / np.abs(cosr))[:, None]
        bsdf = np.where(trans[:, None], color * beer * inv_cos, np.where(spec_o[:, None], color * inv_cos, color * INVPI)).astype(f32)
        new_pdf = np.where(diffuse, np.maximum(f32(0), cosr) * INVPI, f32(1)).astype(f32)
    new_pdf = np.where(dead_tir, f32(0), new_pdf)
    new_pdf = np.where(~refract_t & (dot(gF, R) <= 0), f32(0), new_pdf)
    return R, new_pdf, ~diffuse, bsdf, cosr


def light_pick_prob(lights, ltri, O, N, I):
    """LightPickProb (lights_shared.h:122-138): the probability that next-event estimation at (O, N) would have picked area light `ltri`,
    seen at point I; 0 where ltri < 0.  (The reference sums into a fixed float[8] and overruns it with more lights; nothing is stored here.)"""
    n = len(O)
    with np.errstate(all="ignore"):
        pot = lights.potentials(O, N, P=I)
        ssum = np.zeros(n, f32)
        for k in range(pot.shape[1]):
            ssum = (ssum + pot[:, k]).astype(f32)
        mine = pot[np.arange(n), np.maximum(ltri, 0)] if pot.shape[1] else np.zeros(n, f32)
        return np.where((ltri >= 0) & (ssum > 0), mine / ssum, f32(0)).astype(f32)


def sample_skydome(D, sky, sky_image):
    """tools_shared.h:185-192."""
    if sky_image is None:
        return np.broadcast_to(np.asarray(sky, f32), D.shape).astype(f32)
    h, w = sky_image.shape[:2]
    u = (f32(w) * f32(0.5) * (f32(1) + np.arctan2(D[:, 0], -D[:, 2]).astype(f32) * INVPI)).astype(f32).astype(np.int64)
    v = (f32(h) * np.arccos(np.clip(D[:, 1], -1, 1)).astype(f32) * INVPI).astype(f32).astype(np.int64)
    idx = u + v * w
    flat = sky_image.reshape(-1, 3).astype(f32)
    ok = (idx >= 0) & (idx < w * h)
    out = np.zeros(D.shape, f32)
    out[ok] = flat[idx[ok]]
    return out


def decode_texels(tex):
    """One texture (width, height, texels, mip_levels[, storage]) as float32[n, 4]: ARGB32 / NRM32 words through
    __uchar4_to_float4 (sampling_shared.h:29-33: bytes / 256), ARGB128 (storage 1) texels as they are (:55-59)."""
    if len(tex) > 4 and tex[4] == 1:
        return np.asarray(tex[2], f32).reshape(-1, 4)
    t = np.asarray(tex[2], np.uint32)
    return np.stack([(t & 255), (t >> 8) & 255, (t >> 16) & 255, t >> 24], axis=1).astype(f32) * f32(1.0 / 256.0)


def fetch_texel(texels, tu, tv, o, w, h):
    """sampling_shared.h:36-76 (BILINEAR); o, w, h are arrays (per ray); texels = float32[N, 4], see decode_texels."""
    tcx = (np.maximum(tu + f32(1000), f32(0)) * w.astype(f32) - f32(0.5)).astype(f32)
    tcy = (np.maximum(tv + f32(1000), f32(0)) * h.astype(f32) - f32(0.5)).astype(f32)
    iu, iv = tcx.astype(np.int64) % w, tcy.astype(np.int64) % h
    fu, fv = (tcx - np.floor(tcx)).astype(f32), (tcy - np.floor(tcy)).astype(f32)
    w0, w1, w2 = (1 - fu) * (1 - fv), fu * (1 - fv), (1 - fu) * fv
    w3 = (1 - (w0 + w1 + w2)).astype(f32)
    iu1, iv1 = (iu + 1) % w, (iv + 1) % h

    def px(x, y):
        return texels[o + x + y * w]
    return (px(iu, iv) * w0[:, None] + px(iu1, iv) * w1[:, None] + px(iu, iv1) * w2[:, None] + px(iu1, iv1) * w3[:, None]).astype(f32)


def fetch_texel_trilinear(texels, lam, tu, tv, offset, width, height, levels):
    """sampling_shared.h:78-93 with the texture's own MIP count."""
    level0 = np.minimum(levels - 1, np.trunc(lam).astype(np.int64))
    level1 = np.minimum(levels - 1, level0 + 1)
    fl = (lam - np.floor(lam)).astype(f32)

    def level(lv):
        o, w, h = offset.astype(np.int64).copy(), width.astype(np.int64).copy(), height.astype(np.int64).copy()
        for i in range(int(max(lv.max(), 0))):
            go = lv > i
            o = np.where(go, o + w * h, o); w = np.where(go, w >> 1, w); h = np.where(go, h >> 1, h)
        return o, w, h
    p0 = fetch_texel(texels, tu, tv, *level(level0))
    p1 = fetch_texel(texels, tu, tv, *level(level1))
    return ((1 - fl)[:, None] * p0 + fl[:, None] * p1).astype(f32)


def blue_noise_sampler(table, x, y, sample_index, dim):
    """blueNoiseSampler (tools_shared.h:332-347) over arrays of pixel coordinates; table = uint32[5 * 65536]."""
    x, y = np.asarray(x, np.int64) & 127, np.asarray(y, np.int64) & 127
    si, dim = int(sample_index) & 255, int(dim) & 255
    ranked = (si ^ table[dim + (x + y * 128) * 8 + 65536 * 3].astype(np.int64)) & 255
    value = table[dim + ranked * 256].astype(np.int64)
    value = value ^ table[(dim & 7) + (x + y * 128) * 8 + 65536].astype(np.int64)
    return ((f32(0.5) + value.astype(f32)) * f32(1.0 / 256.0)).astype(f32)


def disney_tint(color):
    """tint4 of GetShadingData (material_shared.h:19-35, 68-69): the base colour at unit luminance, and CIE Y."""
    c = np.asarray(color, f32)
    M = np.array([[0.412453, 0.357580, 0.180423], [0.212671, 0.715160, 0.072169], [0.019334, 0.119193, 0.950227]], f32)
    Mi = np.array([[3.240479, -1.537150, -0.498535], [-0.969256, 1.875992, 0.041556], [0.055648, -0.204043, 1.057311]], f32)
    xyz = np.stack([np.maximum(f32(0), ((M[k, 0] * c[:, 0] + M[k, 1] * c[:, 1]).astype(f32) + M[k, 2] * c[:, 2]).astype(f32)) for k in range(3)], axis=1)
    Y = xyz[:, 1]
    with np.errstate(all="ignore"):
        s = (xyz * (f32(1) / Y)[:, None]).astype(f32)
    rgb = np.stack([np.maximum(f32(0), ((Mi[k, 0] * s[:, 0] + Mi[k, 1] * s[:, 1]).astype(f32) + Mi[k, 2] * s[:, 2]).astype(f32)) for k in range(3)], axis=1)
    rgb = np.where((Y > 0)[:, None], rgb, f32(1)).astype(f32)
    return np.concatenate([rgb, Y[:, None]], axis=1).astype(f32)


def clamp_and_fix(c, clamp):
    m = np.max(c, axis=1)
    with np.errstate(all="ignore"):
        scale = np.where(m > clamp, clamp / m, f32(1)).astype(f32)
    c = (c * scale[:, None]).astype(f32)
    bad = ~np.isfinite(c.sum(axis=1))
    c[bad] = 0
    return c


def render(oracle, meshes, instances, view, width, height, light=None, lights=None, sky=(0, 0, 0), sky_image=None,
           materials=None, prim_materials=None, vertex_normals=None, uvs=None, lods=None, textures=None, spread_angle=0.01,
           max_path_length=5, spp=1, sample_offset=0, clamp=10.0, bounce_cap=0, roulette=True, nthreads=8, stats=None,
           disney=None, material_words=None, tangents=None, alphas=None, blue_noise=None, material_maps=None, bitangents=None,
           filter_out=None, filter_jitter=(0.0, 0.0), blue_slot=0, salt=7907):
    """Returns the accumulator float32[h, w, 4] (rgb sums over samples, w = depth sum of the primary hits).
    meshes: list of verts[T,3,3]; instances: list of (mesh, mat16 or None); light = (pos3, radiance3) shortcut for one
    point light, or lights = Lights(...); materials = float32[n, 4] (albedo rgb, roughness) or the float32[n, 16] table of
    b200_set_material_table; prim_materials = per mesh uint32[T] material index (default 0); vertex_normals = per mesh
    float32[T, 3, 3] or None; uvs = per mesh float32[T, 3, 2] or None, lods = per mesh float32[T] or None; textures = list of
    (width, height, texels uint32[], mip_levels[, storage]) (storage 1: HDR, texels float32[n, 4]); sky_image = float32[h, w, 3] skydome.  Pixel-centre rays (jitter off).
    disney = oracle.DisneyRef() switches the BSDF to the reference's Disney one (Setting bsdf = 1): material_words =
    uint32[n, 4] (core.disney_words), tangents = per mesh float32[T, 3] (CoreTri::T), alphas = per mesh float32[T, 3]
    (CoreTri::alpha; consistent normal interpolation, material_shared.h:87-97).  material_maps = per material None or a dict
    with "detail" / "normals" / "detail_normals" / "roughness_map" -> dict(texture, scale, uvscale, uvoffset)
    (material_shared.h:120-169), bitangents = per mesh float32[T, 3] (CoreTri::B, for the normal maps).
    filter_out = {}: the frame as the reference's filtering core renders it (lib/RenderCore_Optix7Filter/kernels/pathtracer.h;
    pass bounce_cap=1, roulette=False): rays through the pixel CORNERS of the view shifted by filter_jitter pixels, no lens; light that
    arrives after the diffuse bounce goes to a second accumulator; the first sample writes the feature buffers.  Filled with
    "indirect" float32[h, w, 4], "features" uint32[h, w, 4] (history bits zero), "world_pos" float32[h, w, 4], "delta_depth" float32[h, w, 4].
    blue_slot is added to the sample index of the shade stage's blue-noise lookups (the filtering core rotates it per frame); salt = R0."""
    npx = width * height
    acc = np.zeros((npx, 4), np.float64)
    filt = filter_out is not None
    acc2 = np.zeros((npx, 4), np.float64)
    feat = np.zeros((npx, 4), np.uint32)
    wpos, ddep = np.zeros((npx, 4), f32), np.zeros((npx, 4), f32)
    pos, p1 = view["pos"].astype(f32), view["p1"].astype(f32)
    right, up = (view["p2"] - view["p1"]).astype(f32), (view["p3"] - view["p1"]).astype(f32)
    if filt:
        p1 = (p1 + ((f32(filter_jitter[0]) / f32(width)) * right + (f32(filter_jitter[1]) / f32(height)) * up)).astype(f32)

    def pack_normal2(n):
        q = [np.minimum(((n[:, k] + f32(1)) * f32(511)).astype(np.int64), 1023).astype(np.uint32) for k in range(3)]
        return (q[0] << np.uint32(2)) + (q[1] << np.uint32(12)) + (q[2] << np.uint32(22))

    def hdr_to_rgb32(c):
        c = np.minimum(f32(1), c.astype(f32))
        return ((f32(1023) * c[:, 0]).astype(np.uint32) << np.uint32(22)) + ((f32(2047) * c[:, 1]).astype(np.uint32) << np.uint32(11)) + (f32(2047) * c[:, 2]).astype(np.uint32)

    def rgb32_to_hdr(c):
        return np.stack([(c >> np.uint32(22)).astype(f32) * f32(1.0 / 1023.0), ((c >> np.uint32(11)) & np.uint32(2047)).astype(f32) * f32(1.0 / 2047.0),
                         (c & np.uint32(2047)).astype(f32) * f32(1.0 / 2047.0)], axis=1).astype(f32)

    def store_features(px, albedo, pn, t, via, mat_id, Iw=None, Nw=None):
        feat[px, 0], feat[px, 1], feat[px, 2], feat[px, 3] = albedo, pn, t.astype(f32).view(np.uint32), (via << np.uint32(4)) + (np.asarray(mat_id).astype(np.uint32) << np.uint32(6))
        if Iw is not None:
            wpos[px, :3], wpos[px, 3] = Iw, pn.view(f32)
        if Nw is not None:  # depth derivatives: the neighbouring pixels' rays on the plane of this surface (pathtracer.h:48-58)
            fx, fy = (px % width).astype(f32), (px // width).astype(f32)
            plane = dot((Iw - pos[None]).astype(f32), Nw)
            dx = normalize(((p1[None] + right[None] * ((fx + f32(1.5)) * f32(1.0 / width))[:, None]).astype(f32) + up[None] * ((fy + f32(0.5)) * f32(1.0 / height))[:, None] - pos[None]).astype(f32))
            dy = normalize(((p1[None] + right[None] * ((fx + f32(0.5)) * f32(1.0 / width))[:, None]).astype(f32) + up[None] * ((fy + f32(1.5)) * f32(1.0 / height))[:, None] - pos[None]).astype(f32))
            with np.errstate(all="ignore"):
                ddep[px, 2], ddep[px, 3] = (plane / dot(Nw, dx) - t).astype(f32), (plane / dot(Nw, dy) - t).astype(f32)
    clamp = f32(clamp)
    if lights is None:
        lights = Lights()
        if light is not None:
            lp, lr = np.asarray(light[0], f32), np.asarray(light[1], f32)
            lights = Lights(point=np.array([[lp[0], lp[1], lp[2], lr[0] + lr[1] + lr[2], lr[0], lr[1], lr[2], 0]], f32))
    normals = []
    for v in meshes:
        n = np.cross(v[:, 1] - v[:, 0], v[:, 2] - v[:, 0]).astype(f32)
        l2 = dot(n, n)
        il = np.where(l2 > 0, f32(1) / np.sqrt(np.maximum(l2, 1e-38)), 0).astype(f32)
        normals.append(n * il[:, None])
    mats = [np.eye(4, dtype=f32) if m is None else np.asarray(m, f32).reshape(4, 4) for _, m in instances]
    inst_mesh = [mesh for mesh, _ in instances]
    if materials is not None:
        materials = np.asarray(materials, f32)
        if materials.shape[-1] == 4:
            full = np.zeros((len(materials), 16), f32)
            full[:, :4], full[:, 5], full[:, 11], full[:, 12:14] = materials, 1, -1, 1
            materials = full
    tex_pool, tex_rec = np.zeros((1, 4), f32), []
    if textures:
        off = 0
        chunks = []
        for tex in textures:
            tw, th, tx, mips = tex[:4]
            tex_rec.append((off, tw, th, mips))
            chunks.append(decode_texels(tex))
            off += len(chunks[-1])
        tex_pool = np.concatenate(chunks)
    invs = [np.linalg.inv(m.astype(np.float64)).astype(f32) for m in mats]

    def add(px, c, indirect=None):
        if indirect is None or not filt:
            np.add.at(acc, (px, slice(0, 3)), c.astype(np.float64))
        else:
            np.add.at(acc, (px[~indirect], slice(0, 3)), c[~indirect].astype(np.float64))
            np.add.at(acc2, (px[indirect], slice(0, 3)), c[indirect].astype(np.float64))

    for s in range(spp):
        sample = sample_offset + s
        R0 = np.uint32(salt)
        pix = np.arange(npx, dtype=np.uint32)
        path_idx = (pix + np.uint32(sample) * np.uint32(npx)).astype(np.uint32)
        x, y = (pix % width).astype(f32), (pix // width).astype(f32)
        u, v = (x + f32(0.5)) / f32(width), (y + f32(0.5)) / f32(height)
        O = np.broadcast_to(pos, (npx, 3)).astype(f32).copy()
        aperture, distortion = f32(view.get("aperture", 0.0)), f32(view.get("distortion", 0.0))
        if filt:  # Optix7Filter/optix/.optix.cu:84-88
            u, v, aperture, distortion = x / f32(width), y / f32(height), f32(0), f32(0)
            if s == 0:
                feat[:, 2] = f32(1e34).view(np.uint32)
        if aperture > 0:  # RandomPointOnLens (rendercore_optix7/optix/.optix.cu:51-64); jitter is off, so the lens takes the first two numbers
            seed = wang_hash(path_idx + R0)
            if blue_noise is not None and sample < 256:
                l0 = blue_noise_sampler(blue_noise, pix % width, pix // width, sample, 2)
                l1 = blue_noise_sampler(blue_noise, pix % width, pix // width, sample, 3)
            else:
                seed, l0 = random_float(seed)
                seed, l1 = random_float(seed)
            blade = (l0 * f32(9)).astype(np.int32).astype(f32)
            l2 = ((l0 - blade * f32(1.0 / 9.0)) * f32(9)).astype(f32)
            a1, a2 = blade * f32(np.pi) / f32(4.5), (blade + f32(1)) * f32(np.pi) / f32(4.5)
            (x1, y1), (x2, y2) = sincos_det(a1), sincos_det(a2)
            swap = (l1 + l2) > 1
            l1, l2 = np.where(swap, f32(1) - l1, l1).astype(f32), np.where(swap, f32(1) - l2, l2).astype(f32)
            xr, yr = (x1 * l1 + x2 * l2).astype(f32), (y1 * l1 + y2 * l2).astype(f32)
            O = (O + aperture * (right[None] * xr[:, None] + up[None] * yr[:, None])).astype(f32)
        if distortion != 0:  # barrel distortion (.optix.cu:88-97)
            tx, ty = x / f32(width) - f32(0.5), y / f32(height) - f32(0.5)
            rr = tx * tx + ty * ty
            rq = np.sqrt(rr) * (f32(1) + distortion * rr + distortion * rr * rr)
            theta = np.arctan2(tx, ty)
            bx, by = (np.sin(theta) * rq + f32(0.5)) * f32(width), (np.cos(theta) * rq + f32(0.5)) * f32(height)
            u, v = ((bx + f32(0.5)) / f32(width)).astype(f32), ((by + f32(0.5)) / f32(height)).astype(f32)
        D = (p1[None] + u[:, None] * right[None] + v[:, None] * up[None] - O).astype(f32)
        D = (D * (f32(1) / np.sqrt(dot(D, D)))[:, None]).astype(f32)
        T = np.ones((npx, 3), f32)
        pdf_prev = np.ones(npx, f32)
        flags = np.zeros(npx, np.uint32)
        lastN = np.zeros((npx, 3), f32)
        for length in range(1, max_path_length + 1):
            if len(O) == 0:
                break
            if stats is not None:
                stats.setdefault("ext", []).append(len(O))
            t, prim, inst = oracle.trace(O, D, nthreads=nthreads)
            px = (path_idx % np.uint32(npx)).astype(np.int64)
            miss = prim < 0
            if length == 1:
                np.add.at(acc, (px, 3), np.where(miss, 10000.0, t.astype(np.float64)))
            bounced = (flags & S_BOUNCED) != 0      # as the wave starts: which accumulator plane its light goes to (pathtracer.h:93)
            first = np.full(len(O), filt and s == 0) & ~bounced
            via = ((flags & S_VIASPECULAR) != 0).astype(np.uint32)
            with np.errstate(all="ignore"):
                sky_c = clamp_and_fix((T[miss] * sample_skydome(D[miss], sky, sky_image) * (f32(1) / pdf_prev[miss])[:, None]).astype(f32), clamp)
                add(px[miss], sky_c, bounced[miss])
            fm = first[miss]
            if fm.any():  # the path ends before a diffuse vertex (:115-123)
                Dm, Om = D[miss][fm], O[miss][fm]
                store_features(px[miss][fm], hdr_to_rgb32(sky_c[fm]), pack_normal2((Dm * f32(-1)).astype(f32)) + via[miss][fm], np.full(int(fm.sum()), 1e34, f32), via[miss][fm], 0,
                               (Om + f32(50000) * Dm).astype(f32))
                ddep[px[miss][fm]] = 0
            h = ~miss
            if not h.any():
                break
            O, D, T, t, prim, inst, path_idx, px = O[h], D[h], T[h], t[h], prim[h], inst[h], path_idx[h], px[h]
            pdf_prev, flags, lastN = pdf_prev[h], flags[h], lastN[h]
            bounced, first, via = bounced[h], first[h], via[h]
            n = len(O)
            N = np.empty_like(O)
            sN = np.empty_like(O)
            mat = np.zeros(n, np.int64)

            def to_world(m, v):
                return np.stack([m[0, 0] * v[:, 0] + m[0, 1] * v[:, 1] + m[0, 2] * v[:, 2],
                                 m[1, 0] * v[:, 0] + m[1, 1] * v[:, 1] + m[1, 2] * v[:, 2],
                                 m[2, 0] * v[:, 0] + m[2, 1] * v[:, 1] + m[2, 2] * v[:, 2]], axis=1).astype(f32)
            for i in np.unique(inst):
                sel = inst == i
                mesh = inst_mesh[i]
                N[sel] = normalize(to_world(mats[i], normals[mesh][prim[sel]]))
                if prim_materials is not None:
                    mat[sel] = prim_materials[mesh][prim[sel]]
            color = np.full((n, 3), f32(0.8))
            rough = np.ones(n, f32)
            transmission, eta, smooth = np.zeros(n, f32), np.ones(n, f32), np.zeros(n, bool)
            absorption = np.zeros((n, 3), f32)
            if materials is not None:
                known = mat < len(materials)
                mk = materials[mat[known]]
                color[known], rough[known], transmission[known], eta[known] = mk[:, :3], mk[:, 3], mk[:, 4], mk[:, 5]
                mflags = np.zeros(n, np.int64)
                mflags[known] = mk[:, 6].astype(np.int64)
                smooth[known], absorption[known] = (mflags[known] & 1) != 0, mk[:, 8:11]
                tex_id = np.full(n, -1, np.int64)
                tex_id[known] = mk[:, 11].astype(np.int64)
                uvso = np.zeros((n, 4), f32)
                uvso[known] = mk[:, 12:16]
            words = np.tile(np.array([0x80800000, 0, 0, f32(1).view(np.uint32)], np.uint32), (n, 1))
            base_color = color.copy()
            if disney is not None and materials is not None:
                words[known] = np.asarray(material_words, np.uint32)[mat[known]]
                byte = lambda w, sh: (((w >> np.uint32(sh)) & np.uint32(255)).astype(f32) * f32(1.0 / 255.0)).astype(f32)
                rough = np.maximum(f32(0.001), byte(words[:, 0], 24))
                transmission = byte(words[:, 2], 16)
            tang = np.zeros((n, 3), f32)
            # interpolated normals for SMOOTH materials: w N0 + u N1 + v N2 with the 16-bit barycentrics of the hit record
            sN[:] = N
            I_all = (O + D * t[:, None]).astype(f32)
            passthrough = np.zeros(n, bool)
            mapped = np.zeros(n, bool)
            if materials is not None and textures:
                mapped = (tex_id >= 0) & (tex_id < len(tex_rec))
            has_maps = np.zeros(n, bool)
            if material_maps is not None and textures:
                for k, mm in enumerate(material_maps):
                    if mm is not None:
                        has_maps |= mat == k
            if vertex_normals is not None or mapped.any() or disney is not None or has_maps.any():
                for i in np.unique(inst):
                    mesh = inst_mesh[i]
                    has_vn = vertex_normals is not None and vertex_normals[mesh] is not None
                    sel = np.nonzero((inst == i) & ((smooth & has_vn) | mapped | has_maps | (disney is not None)))[0]
                    if len(sel) == 0:
                        continue
                    tri = meshes[mesh][prim[sel]]
                    iv = invs[i]
                    Oo = (to_world(iv, O[sel]) + iv[:3, 3][None]).astype(f32)
                    Do = normalize(to_world(iv, D[sel]))
                    e1, e2 = (tri[:, 1] - tri[:, 0]).astype(f32), (tri[:, 2] - tri[:, 0]).astype(f32)
                    hh = np.cross(Do, e2).astype(f32)
                    ff = (f32(1) / dot(e1, hh)).astype(f32)
                    ss = (Oo - tri[:, 0]).astype(f32)
                    bu = (ff * dot(ss, hh)).astype(f32)
                    bv = (ff * dot(Do, np.cross(ss, e1).astype(f32))).astype(f32)
                    bu = np.floor(f32(65535) * bu).astype(f32) * f32(1.0 / 65535.0)
                    bv = np.floor(f32(65535) * bv).astype(f32) * f32(1.0 / 65535.0)
                    bw = f32(1) - (bu + bv)
                    if has_vn:
                        sm = smooth[sel]
                        vn = vertex_normals[mesh][prim[sel]].astype(f32)
                        on = (vn[:, 0] * bw[:, None] + vn[:, 1] * bu[:, None] + vn[:, 2] * bv[:, None]).astype(f32)
                        ok = sm & (dot(on, on) > 0)
                        sN[sel[ok]] = normalize(to_world(mats[i], on[ok]))
                    if disney is not None:
                        if alphas is not None:  # consistent normal interpolation (material_shared.h:87-97)
                            al = alphas[mesh][prim[sel]].astype(f32) if alphas[mesh] is not None else np.zeros((len(sel), 3), f32)
                            alpha = ((bw * al[:, 0] + bu * al[:, 1]).astype(f32) + bv * al[:, 2]).astype(f32)
                            back = dot(D[sel], N[sel]) > 0
                            sg = np.where(back, f32(-1), f32(1))[:, None]
                            sN[sel] = (disney.consistent_normal((D[sel] * f32(-1)).astype(f32), (sN[sel] * sg).astype(f32), alpha) * sg).astype(f32)
                        if tangents is not None and tangents[mesh] is not None:
                            tang[sel] = tangents[mesh][prim[sel]]
                    mp = mapped[sel]
                    maps_here = bool(has_maps[sel].any())
                    if mp.any() or maps_here:
                        uv_all = uvs[mesh][prim[sel]].astype(f32) if (uvs is not None and uvs[mesh] is not None) else np.zeros((len(sel), 3, 2), f32)
                        tu_all = (bw * uv_all[:, 0, 0] + bu * uv_all[:, 1, 0] + bv * uv_all[:, 2, 0]).astype(f32)
                        tv_all = (bw * uv_all[:, 0, 1] + bu * uv_all[:, 1, 1] + bv * uv_all[:, 2, 1]).astype(f32)
                    if mp.any():
                        # diffuse map (material_shared.h:99-131): trilinear fetch at the ray-cone LOD, alpha below 0.5 = pass through
                        q = sel[mp]
                        ld = lods[mesh][prim[q]].astype(f32) if (lods is not None and lods[mesh] is not None) else np.zeros(len(q), f32)
                        tu, tv = tu_all[mp], tv_all[mp]
                        with np.errstate(all="ignore"):
                            lam = (ld + np.log2(f32(spread_angle) * t[q] * (f32(1) / np.abs(dot(D[q], N[q]))))).astype(f32)
                        rec = np.array([tex_rec[k] for k in tex_id[q]], np.int64)
                        texel = fetch_texel_trilinear(tex_pool, lam, (uvso[q, 0] * (uvso[q, 2] + tu)).astype(f32), (uvso[q, 1] * (uvso[q, 3] + tv)).astype(f32),
                                                      rec[:, 0], rec[:, 1], rec[:, 2], rec[:, 3])
                        hole = ((mflags[q] & 2) != 0) & (texel[:, 3] < f32(0.5))
                        passthrough[q[hole]] = True
                        color[q[~hole]] = (color[q[~hole]] * texel[~hole, :3]).astype(f32)
                    if maps_here:
                        def fetch_map(mm, idx):
                            o, tw, th, _ = tex_rec[mm["texture"]]
                            us, uo = mm.get("uvscale", (1.0, 1.0)), mm.get("uvoffset", (0.0, 0.0))
                            return fetch_texel(tex_pool, (f32(us[0]) * (f32(uo[0]) + tu_all[idx])).astype(f32), (f32(us[1]) * (f32(uo[1]) + tv_all[idx])).astype(f32),
                                               np.full(len(idx), o, np.int64), np.full(len(idx), tw, np.int64), np.full(len(idx), th, np.int64))
                        for k in np.unique(mat[sel]):
                            mm = material_maps[k] if k < len(material_maps) else None
                            if mm is None:
                                continue
                            idx = np.nonzero((mat[sel] == k) & ~passthrough[sel])[0]
                            if len(idx) == 0:
                                continue
                            g = sel[idx]
                            if "detail" in mm and mapped[g].all():  # additive around 0.5, needs the base map (:120-130)
                                color[g] = ((color[g] + fetch_map(mm["detail"], idx)[:, :3]).astype(f32) - f32(0.5)).astype(f32)
                            if "normals" in mm:  # :132-159, with the map's scale field for the xy scale
                                sn = ((fetch_map(mm["normals"], idx)[:, :3] - f32(0.5)) * f32(2)).astype(f32)
                                sn[:, :2] = (sn[:, :2] * f32(mm["normals"].get("scale", 1.0))).astype(f32)
                                if "detail_normals" in mm:
                                    l1 = ((fetch_map(mm["detail_normals"], idx)[:, :3] - f32(0.5)) * f32(2)).astype(f32)
                                    l1[:, :2] = (l1[:, :2] * f32(mm["detail_normals"].get("scale", 1.0))).astype(f32)
                                    sn = (sn + l1).astype(f32)
                                sn = normalize(sn)
                                Tt = tangents[mesh][prim[g]].astype(f32)
                                Bt = bitangents[mesh][prim[g]].astype(f32)
                                sN[g] = normalize(((Tt * sn[:, 0:1]).astype(f32) + (Bt * sn[:, 1:2]).astype(f32)).astype(f32) + (sN[g] * sn[:, 2:3]).astype(f32))
                            if "roughness_map" in mm:  # :161-169: the red channel replaces the roughness byte
                                rb = (fetch_map(mm["roughness_map"], idx)[:, 0] * f32(255)).astype(f32).astype(np.uint32)
                                rough[g] = np.maximum(f32(0.001), rb.astype(f32) * f32(1.0 / 255.0))
                                words[g, 0] = (words[g, 0] & np.uint32(0x00ffffff)) | (rb << np.uint32(24))
            if disney is not None:  # a triangle without a tangent gets an arbitrary one (the reference would divide by zero)
                notan = ~(dot(tang, tang) > 0)
                tang[notan] = np.where((np.abs(sN[notan, 0]) > f32(0.9))[:, None], np.array([0, 1, 0], f32), np.array([1, 0, 0], f32))
            I = I_all
            # alpha holes (pathtracer.h:104-116): the path continues from just behind the surface with its state unchanged
            carry = None
            if passthrough.any() and length < max_path_length:
                pt = passthrough
                Tpt = T[pt].copy()
                Tpt[~np.isfinite(Tpt.sum(axis=1))] = 0
                carry = ((I[pt] + D[pt] * EPS).astype(f32), D[pt], Tpt, pdf_prev[pt], flags[pt], lastN[pt], path_idx[pt])
            if filt and passthrough.any() and length == max_path_length:  # it ends here: store something sensible (:143-154)
                q = np.nonzero(passthrough & first)[0]
                store_features(px[q], 0, pack_normal2(N[q]) + via[q], t[q], via[q], 0, I[q])
            if filt:  # FILTERINGCORE (material_shared.h:172-177)
                color = np.maximum(f32(0.05), color).astype(f32)
            if passthrough.any():
                kp = ~passthrough
                bounced, first, via, mat = bounced[kp], first[kp], via[kp], mat[kp]
                O, D, T, t, prim, inst, path_idx, px, N, sN, I, color, rough = (a[kp] for a in (O, D, T, t, prim, inst, path_idx, px, N, sN, I, color, rough))
                words, base_color, tang = words[kp], base_color[kp], tang[kp]
                pdf_prev, flags, lastN, transmission, eta, absorption = (a[kp] for a in (pdf_prev, flags, lastN, transmission, eta, absorption))
                n = len(O)
            # ---- paths that end on a light (pathtracer.h:120-148)
            emissive = (color > 1).any(axis=1)
            if emissive.any():
                e = np.nonzero(emissive)[0]
                ddn = -dot(D[e], N[e])
                c = (T[e] * color[e]).astype(f32)
                mis = (length > 1) & ((flags[e] & S_SPECULAR) == 0)
                if mis.any():
                    ltri = np.full(len(e), -1)
                    area = np.zeros(len(e), f32)
                    for li, a in enumerate(lights.area):
                        match = (prim[e] == a[15:16].view(np.int32)[0]) & (inst[e] == a[19:20].view(np.int32)[0])
                        ltri[match], area[match] = li, a[7]
                    with np.errstate(all="ignore"):
                        light_pdf = np.where(ltri >= 0, (t[e] * t[e]) / (ddn * area), f32(0)).astype(f32)
                        pick = light_pick_prob(lights, ltri, O[e], lastN[e], I[e])
                        w = (pdf_prev[e] + light_pdf * pick).astype(f32)
                        cm = np.where((w > 0)[:, None], c * (f32(1) / w)[:, None], f32(0)).astype(f32)
                    c = np.where(mis[:, None], cm, c)
                c = clamp_and_fix(c, clamp)
                front = ddn > 0
                add(px[e][front], c[front], bounced[e][front])
                q = e[first[e]]
                if len(q):  # the path ends on a light (:193-201)
                    store_features(px[q], hdr_to_rgb32(color[q]), pack_normal2(N[q]) + via[q], t[q], via[q], mat[q], I[q], N[q])
            keep = ~emissive
            if not keep.any():
                if carry is None:
                    break
                O, D, T, pdf_prev, flags, lastN, path_idx = carry
                continue
            O, D, T, t, N, I, color, rough, path_idx, px = (a[keep] for a in (O, D, T, t, N, I, color, rough, path_idx, px))
            sN, transmission, eta, absorption = sN[keep], transmission[keep], eta[keep], absorption[keep]
            words, base_color, tang = words[keep], base_color[keep], tang[keep]
            pdf_prev, flags, lastN = pdf_prev[keep], flags[keep], lastN[keep]
            bounced, first, via, mat = bounced[keep], first[keep], via[keep], mat[keep]
            n = len(O)
            flags = np.where((rough <= f32(0.001)) | (transmission > f32(0.999)), flags | S_SPECULAR, flags & ~np.uint32(S_SPECULAR)).astype(np.uint32)
            if first.any():  # :211-232
                spec_now = (flags & S_SPECULAR) != 0
                q = np.nonzero(first & spec_now)[0]
                feat[px[q], 0] = hdr_to_rgb32(color[q])
                q = np.nonzero(first & ~spec_now)[0]
                albedo = color[q].copy()
                vq = via[q] != 0
                albedo[vq] = (albedo[vq] * rgb32_to_hdr(feat[px[q][vq], 0])).astype(f32)
                store_features(px[q], hdr_to_rgb32(albedo), pack_normal2((sN[q] * f32(-1)).astype(f32)) + via[q], t[q], via[q], mat[q], I[q], N[q])
            seed = wang_hash(path_idx * np.uint32(17) + R0 + np.uint32(length) * np.uint32(91771))
            face = np.where(dot(D, N) > 0, f32(-1), f32(1))
            gF = (N * face[:, None]).astype(f32)   # geometric normal on the side the ray came from
            fN = (sN * face[:, None]).astype(f32)  # shading normal, same side
            absorption = np.where((face == 1)[:, None], f32(0), absorption).astype(f32)
            with np.errstate(all="ignore"):
                T = (T * (f32(1) / pdf_prev)[:, None]).astype(f32)
            # ---- next event estimation (pathtracer.h:163-199)
            nee = (flags & S_SPECULAR) == 0
            if lights.count > 0 and nee.any():
                if blue_noise is not None and sample < 2:  # pathtracer.h:168-173
                    r0 = blue_noise_sampler(blue_noise, px % width, px // width, sample + blue_slot, 4 + 4 * length)
                    r1 = blue_noise_sampler(blue_noise, px % width, px // width, sample + blue_slot, 5 + 4 * length)
                else:
                    seed_n, r0 = random_float(seed.copy())
                    seed_n, r1 = random_float(seed_n)
                    seed = np.where(nee, seed_n, seed)
                q = np.nonzero(nee)[0]
                P, pick, lpdf, lcol = random_point_on_light(lights, r0[q], r1[q], I[q], fN[q])
                L = (P - I[q]).astype(f32)
                dist = np.sqrt(dot(L, L))
                with np.errstate(all="ignore"):
                    L = (L * (f32(1) / dist)[:, None]).astype(f32)
                    ndl = dot(L, fN[q])
                    if disney is None:
                        bpdf = (np.abs(ndl) * INVPI).astype(f32)
                        ok = (ndl > 0) & (lpdf > 0) & (bpdf > 0)
                        sc = (INVPI * rough[q] * (ndl / (pick * lpdf + bpdf))).astype(f32)
                        E = clamp_and_fix((T[q] * color[q] * lcol * sc[:, None]).astype(f32), clamp)
                    else:
                        # EvaluateBSDF with the unflipped shading normal (pathtracer.h:186)
                        val, bpdf = disney.evaluate(color[q], absorption[q], disney_tint(base_color[q]), words[q], sN[q], tang[q], (D[q] * f32(-1)).astype(f32), L)
                        ok = (ndl > 0) & (lpdf > 0) & (bpdf > 0)
                        E = clamp_and_fix((T[q] * val * lcol * (ndl / (pick * lpdf + bpdf))[:, None]).astype(f32), clamp)
                if ok.any():
                    So = safe_origin(I[q][ok], L[ok], gF[q][ok], EPS)
                    tmax = dist[ok] - 2 * EPS
                    if stats is not None:
                        stats.setdefault("shadow", []).append(int(ok.sum()))
                    ts, _, _ = oracle.trace(So, L[ok], nthreads=nthreads)
                    vis = ~(ts < tmax)
                    add(px[q][ok][vis], E[ok][vis], bounced[q][ok][vis])
            # ---- extension (pathtracer.h:201-239, SampleBSDF of lambert.h with TRANSMISSION = 0)
            enough = ((bounce_cap == 1) & ((flags & S_BOUNCED) != 0)) | ((bounce_cap == 2) & ((flags & S_BOUNCEDTWICE) != 0))
            if length == max_path_length:
                q = np.nonzero(first & ~enough)[0]
                if len(q):  # the depth cap ends the path before the feature record was finished (:289-298)
                    store_features(px[q], 0, pack_normal2(N[q]) + via[q], t[q], via[q], 0, I[q])
                break
            if blue_noise is not None and sample < 256:  # pathtracer.h:212-217
                r3 = blue_noise_sampler(blue_noise, px % width, px // width, sample + blue_slot, 6 + 4 * length)
                r4 = blue_noise_sampler(blue_noise, px % width, px // width, sample + blue_slot, 7 + 4 * length)
            else:
                seed, r3 = random_float(seed)
                seed, r4 = random_float(seed)
            R, new_pdf, spec, bsdf, cosr = lambert_sample(color, absorption, rough, transmission, eta, fN, gF, face, D, t, r3, r4)
            if disney is not None:
                # SampleBSDF( shadingData, fN, N, T, -D, HIT_T, r3, r4, ... ) (pathtracer.h:219); total internal reflection ends the path
                R, new_pdf, spec, bsdf = disney.sample(color, absorption, disney_tint(base_color), words, sN, N, tang, (D * f32(-1)).astype(f32), t, r3, r4)
                new_pdf = np.where(spec & (np.abs(R).sum(axis=1) == 0), f32(0), new_pdf).astype(f32)
                with np.errstate(all="ignore"):
                    cosr = dot(fN, R)
            alive = ~enough & (new_pdf >= f32(1e-4))
            flags = np.where(alive & spec, flags | S_SPECULAR, flags).astype(np.uint32)
            p = np.where(((flags & S_SPECULAR) != 0) | ((flags & S_BOUNCED) == 0) | (not roulette), f32(1), np.minimum(f32(1), bsdf.max(axis=1))).astype(f32)
            seed, rr = random_float(seed)
            alive &= ~(p < rr)
            with np.errstate(all="ignore"):
                T = (T * (f32(1) / p)[:, None]).astype(f32)
            is_spec = (flags & S_SPECULAR) != 0
            flags = np.where(is_spec, flags | S_VIASPECULAR, flags | np.where((flags & S_BOUNCED) != 0, S_BOUNCEDTWICE, S_BOUNCED)).astype(np.uint32)
            T[~np.isfinite(T.sum(axis=1))] = 0
            newO = safe_origin(I, R, gF, EPS)
            newT = (T * bsdf * np.abs(cosr)[:, None]).astype(f32)
            O, D, T, pdf_prev, flags, lastN, path_idx = newO[alive], R[alive], newT[alive], new_pdf[alive], flags[alive], fN[alive], path_idx[alive]
            if carry is not None:
                O, D, T, pdf_prev, flags, lastN, path_idx = (np.concatenate([a, b]) for a, b in zip((O, D, T, pdf_prev, flags, lastN, path_idx), carry))
    if filt:
        filter_out.update(indirect=acc2.reshape(height, width, 4).astype(f32), features=feat.reshape(height, width, 4), world_pos=wpos.reshape(height, width, 4),
                          delta_depth=ddep.reshape(height, width, 4))
    return acc.reshape(height, width, 4).astype(f32)
