"""The SVGF-style filter and TAA (lighthouse2_b200/csrc/filter.cu) against the reference's own kernels
(lib/CUDA/shared_kernel_code/finalize_shared.h:102-483, compiled in place by nvcc: oracle/ref_filter_driver.cu), stage by stage on the
same device buffers: prepare (demodulation, reprojection, moments), the three reconstruction passes, TAA, unsharp mask, the no-TAA finish.

Bars.  The packed outputs (5:11 fixed point light, history counters) are compared as integers: equal but for an ulp-of-the-packing step
on a few pixels, since the two builds may contract a different multiply-add here and there (both are nvcc builds with FMA on; the code
shapes differ).  Float outputs: 1e-4 of the value range.  Every stage also has a bound on the share of pixels that may differ at all by
more than that (a threshold of the filter -- a normal within 0.95, a history tap in or out -- can fall on the other side): 0.2 %."""

import numpy as np
import pytest
import torch

from lighthouse2_b200 import core

pytestmark = pytest.mark.gpu

W, H = 176, 104  # not multiples of the 32x8 / 32x2 blocks


def pack_normal(n):
    q = np.clip(((n + 1.0) * 511).astype(np.int64), 0, 1023).astype(np.uint32)
    return (q[..., 0] << 2) + (q[..., 1] << 12) + (q[..., 2] << 22)


def pack_albedo(c):
    c = np.minimum(c, 1.0)
    return ((1023.0 * c[..., 0]).astype(np.uint32) << 22) + ((2047.0 * c[..., 1]).astype(np.uint32) << 11) + (2047.0 * c[..., 2]).astype(np.uint32)


def pack_light(a, b):
    q = lambda v: (np.minimum(v, 31.999) * 2048.0).astype(np.uint32)
    out = np.empty(a.shape[:-1] + (4,), np.uint32)
    out[..., 0], out[..., 1] = (q(a[..., 0]) << 16) + q(a[..., 1]), q(a[..., 2])
    out[..., 2], out[..., 3] = (q(b[..., 0]) << 16) + q(b[..., 1]), q(b[..., 2])
    return out.view(np.float32)


VIEW = dict(pos=(0.0, 1.5, -6.0), p1=(-2.0, 2.625, -3.0), p2=(2.0, 2.625, -3.0), p3=(-2.0, 0.375, -3.0))


def scene(seed, moved):
    """A plausible G-buffer: a floor and a back wall seen through VIEW, a band of mirror pixels, three materials; the previous frame is
    the same surface from a camera shifted by `moved`."""
    rng = np.random.default_rng(seed)
    ys, xs = np.mgrid[0:H, 0:W].astype(np.float32)
    pos, p1, p2, p3 = (np.array(VIEW[k], np.float32) for k in ("pos", "p1", "p2", "p3"))
    target = p1 + ((xs + 0.5) / W)[..., None] * (p2 - p1) + ((ys + 0.5) / H)[..., None] * (p3 - p1)
    D = target - pos
    D /= np.linalg.norm(D, axis=2, keepdims=True)
    t_floor = np.where(D[..., 1] < -1e-3, (0.0 - pos[1]) / np.minimum(D[..., 1], -1e-3), 1e9)
    t_wall = (4.0 - pos[2]) / D[..., 2]
    floor = t_floor < t_wall
    t = np.where(floor, t_floor, t_wall).astype(np.float32)
    P = (pos + D * t[..., None]).astype(np.float32)
    N = np.where(floor[..., None], np.array([0, 1, 0], np.float32), np.array([0, 0, -1], np.float32))
    N = N + rng.normal(0, 0.02, N.shape).astype(np.float32)  # a little bump
    N /= np.linalg.norm(N, axis=2, keepdims=True)
    specular = ((xs > W * 0.55) & (xs < W * 0.7)).astype(np.uint32)  # a mirror band: reprojected by the diamond search
    material = (1 + (xs // 40).astype(np.uint32) % 3) * floor + 7 * (~floor)
    albedo = np.stack([0.3 + 0.5 * ((xs // 8 + ys // 8) % 2), 0.5 + 0.3 * np.sin(xs / 9.0), 0.4 + 0.4 * np.cos(ys / 7.0)], axis=2).astype(np.float32)
    packed_n = pack_normal(N) + specular
    features = np.stack([pack_albedo(albedo), packed_n, t.view(np.uint32), rng.integers(0, 16, (H, W)).astype(np.uint32) + (specular << 4) + (material.astype(np.uint32) << 6)], axis=2)
    world = np.concatenate([P, packed_n.view(np.float32)[..., None]], axis=2)
    prev_world = world.copy()
    prev_world[..., :3] += rng.normal(0, 0.002, P.shape).astype(np.float32)
    prev_world = np.roll(prev_world, 2 if moved else 0, axis=1)  # the previous frame saw everything two pixels to the right
    dd = np.zeros((H, W, 4), np.float32)
    dd[..., 2], dd[..., 3] = np.gradient(t, axis=1), np.gradient(t, axis=0)
    acc = np.zeros((2, H, W, 4), np.float32)
    acc[0, ..., :3] = rng.gamma(2.0, 0.4, (H, W, 3)) * albedo * 4  # 4 spp
    acc[1, ..., :3] = rng.gamma(1.5, 0.2, (H, W, 3)) * albedo * 4
    acc[0, 5:9, 5:9, :3] = 500.0  # fireflies: the clamps
    prev_moments = np.abs(rng.normal(0.8, 0.3, (H, W, 4))).astype(np.float32)
    prev_moments[..., 1] = prev_moments[..., 0] ** 2 + np.abs(rng.normal(0, 0.05, (H, W)))
    prev_moments[..., 3] = prev_moments[..., 2] ** 2 + np.abs(rng.normal(0, 0.05, (H, W)))
    history = pack_light(rng.gamma(2.0, 0.4, (H, W, 3)).astype(np.float32), rng.gamma(1.5, 0.2, (H, W, 3)).astype(np.float32))
    prev_view = np.array([VIEW["pos"], VIEW["p1"], VIEW["p2"], VIEW["p3"]], np.float32)
    if moved:
        prev_view = prev_view + np.array([0.05, 0.0, 0.0], np.float32)
    return dict(features=features, world=world.astype(np.float32), prev_world=prev_world.astype(np.float32), dd=dd, acc=acc, prev_moments=prev_moments.astype(np.float32),
                history=history, prev_view=prev_view)


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a).view(np.int32 if a.dtype == np.uint32 else a.dtype)).cuda()


def host(t, dtype=np.float32):
    return t.cpu().numpy().view(dtype)


def close(name, got, want, scale, share=0.002, atol_rel=1e-4):
    bad = np.abs(got.astype(np.float64) - want.astype(np.float64)) > atol_rel * scale
    bad = bad.reshape(bad.shape[0], bad.shape[1], -1).any(axis=2)
    print("%s: %d of %d pixels differ by more than %.0e of %.3g" % (name, int(bad.sum()), bad.size, atol_rel, scale))
    assert bad.mean() <= share, name


def packed_close(name, got, want, share=0.002):
    """5:11 fixed-point light: the six channels as integers, equal or one step apart."""
    g, w = got.view(np.uint32), want.view(np.uint32)
    chans = lambda v: np.stack([v[..., 0] >> 16, v[..., 0] & 65535, v[..., 1], v[..., 2] >> 16, v[..., 2] & 65535, v[..., 3]], axis=-1).astype(np.int64)
    d = np.abs(chans(g) - chans(w))
    print("%s: channels equal %.4f, one step apart %.4f, more %.5f" % (name, (d == 0).mean(), (d == 1).mean(), (d > 1).mean()))
    assert (d > 1).any(axis=-1).mean() <= share, name
    assert (d == 0).mean() > 0.98, name


@pytest.fixture(scope="module")
def ref():
    import oracle
    return oracle.FilterRef()


@pytest.fixture(scope="module")
def lib():
    return core.load_library()


@pytest.mark.parametrize("moved", [False, True])
def test_prepare_matches_the_reference_kernel(ref, lib, moved):
    s = scene(1, moved)
    p = lambda t: t.data_ptr()
    common = dict(world=dev(s["world"]), prev_world=dev(s["prev_world"]), dd=dev(s["dd"]), acc=dev(s["acc"]), prev_moments=dev(s["prev_moments"]))
    out = {}
    for side in ("ref", "b200"):
        feat = dev(s["features"])
        shading, motion, moments = torch.zeros(H, W, 4, device="cuda"), torch.zeros(H, W, 2, device="cuda"), torch.zeros(H, W, 4, device="cuda")
        j, pj = (0.3, 0.2), (0.7, 0.8)
        if side == "ref":
            ref.prepare(p(common["acc"]), p(feat), p(common["world"]), p(common["prev_world"]), p(shading), p(motion), p(moments), p(common["prev_moments"]), p(common["dd"]),
                        s["prev_view"], j, pj, W, H, 4, 8.0, 6.0, not moved)
        else:
            view = np.ascontiguousarray(s["prev_view"], np.float32).reshape(12)
            code = lib.b200_filter_prepare(p(common["acc"]), p(feat), p(common["world"]), p(common["prev_world"]), p(shading), p(motion), p(moments), p(common["prev_moments"]),
                                           p(common["dd"]), view.ctypes.data, j[0], j[1], pj[0], pj[1], W, H, 4, 8.0, 6.0, int(not moved))
            assert code == 0, lib.b200_last_error().decode()
        out[side] = dict(feat=host(feat, np.uint32), shading=host(shading), motion=host(motion), moments=host(moments))
    a, b = out["b200"], out["ref"]
    packed_close("shading", a["shading"], b["shading"])
    close("motion", a["motion"], b["motion"], float(W))
    close("moments", a["moments"], b["moments"], float(np.abs(b["moments"]).max()))
    same_count = (a["feat"][..., 3] == b["feat"][..., 3]).mean()
    print("history counters equal:", same_count)
    assert same_count > 0.998 and np.array_equal(a["feat"][..., :3], b["feat"][..., :3])
    # the stage did something: histories were found and counted up, fireflies were clamped
    assert ((b["feat"][..., 3] & 15) != (s["features"][..., 3] & 15)).mean() > 0.3
    assert (b["motion"] != 0).all() and np.isfinite(b["moments"]).all()
    if moved:  # the mirror band was searched: its motion vectors moved away from the pixel centres
        band = (np.arange(W) > W * 0.55) & (np.arange(W) < W * 0.7)
        centre = np.stack(np.meshgrid(np.arange(W) + 0.5, np.arange(H) + 0.5), axis=2)
        assert (np.abs(b["motion"] - centre)[:, band].max(axis=2) > 0.5).mean() > 0.3


def _prepared(ref, s, moved):
    """Inputs of the reconstruction passes, made by the reference's prepare."""
    p = lambda t: t.data_ptr()
    d = dict(world=dev(s["world"]), prev_world=dev(s["prev_world"]), dd=dev(s["dd"]), acc=dev(s["acc"]), prev_moments=dev(s["prev_moments"]), feat=dev(s["features"]),
             shading=torch.zeros(H, W, 4, device="cuda"), motion=torch.zeros(H, W, 2, device="cuda"), moments=torch.zeros(H, W, 4, device="cuda"), history=dev(s["history"]))
    ref.prepare(p(d["acc"]), p(d["feat"]), p(d["world"]), p(d["prev_world"]), p(d["shading"]), p(d["motion"]), p(d["moments"]), p(d["prev_moments"]), p(d["dd"]),
                s["prev_view"], (0.3, 0.2), (0.7, 0.8), W, H, 4, 8.0, 6.0, not moved)
    return d


@pytest.mark.parametrize("moved", [False, True])
def test_reconstruction_passes_match_the_reference_kernel(ref, lib, moved):
    s = scene(2, moved)
    d = _prepared(ref, s, moved)
    p = lambda t: t.data_ptr()
    A = d["shading"]
    for phase, last in ((1, 0), (2, 0), (3, 1)):
        want, got = torch.zeros(H, W, 4, device="cuda"), torch.zeros(H, W, 4, device="cuda")
        ref.apply(p(d["feat"]), p(d["prev_world"]), p(d["world"]), p(d["dd"]), p(d["motion"]), p(d["moments"]), p(A), p(d["history"]), p(want), W, H, phase, last)
        code = lib.b200_filter_reconstruct(p(d["feat"]), p(d["prev_world"]), p(d["world"]), p(d["dd"]), p(d["motion"]), p(d["moments"]), p(A), p(d["history"]), p(got), W, H, phase, last)
        assert code == 0, lib.b200_last_error().decode()
        if last:
            close("reconstruct phase %d (remodulated)" % phase, host(got), host(want), float(host(want)[..., :3].max()))
            assert host(want)[..., :3].std() > 0.05
        else:
            packed_close("reconstruct phase %d" % phase, host(got), host(want))
        assert not torch.equal(want, A)  # the pass filtered
        A = want  # both sides go on from the reference's output: differences do not pile up


def test_taa_and_the_two_finishes_match_the_reference_kernels(ref, lib):
    s = scene(3, True)
    d = _prepared(ref, s, True)
    p = lambda t: t.data_ptr()
    rng = np.random.default_rng(9)
    ys, xs = np.mgrid[0:H, 0:W]
    img = np.zeros((H, W, 4), np.float32)
    img[..., :3] = np.sqrt(0.1 + 0.6 * (((xs // 6 + ys // 6) % 2)[..., None]) + rng.random((H, W, 3)) * 0.2)  # gamma-corrected, edges + noise
    prev = np.roll(img, 2, axis=1) * np.float32(0.9)
    prev[3, 7, 0] = np.nan  # a poisoned history texel: the NaN guard
    motion = host(d["motion"])
    hist = dev(prev)
    # The reference blends IN PLACE: a pixel's 3x3 clamp window sees whichever neighbours other threads have already overwritten, so its
    # output changes from run to run wherever the variance clamp binds.  To hold the whole kernel -- clamp included -- to the strict bar,
    # only every third pixel in x and y is blended at a time (the others get a motion vector that points off the screen and pass
    # through): no blended pixel then has a blended neighbour, and nine such runs cover the image.
    blended = np.zeros((H, W), bool)
    for phase in range(9):
        on = (xs % 3 == phase % 3) & (ys % 3 == phase // 3)
        mv = np.where(on[..., None], motion, np.float32(-10.0)).astype(np.float32)
        want, got, src, dmv = dev(img), torch.zeros(H, W, 4, device="cuda"), dev(img), dev(mv)  # (named: a temporary's memory is reused at once)
        ref.taa(p(want), p(hist), p(d["world"]), p(d["prev_world"]), p(dmv), W, H)
        assert lib.b200_filter_taa(p(src), p(hist), p(dmv), p(got), W, H) == 0, lib.b200_last_error().decode()
        g, w = host(got), host(want)
        bad = (np.abs(g - w) > 1e-4).any(axis=2)
        assert np.isfinite(w).all() and bad.mean() <= 0.002, (phase, int(bad.sum()))
        assert np.array_equal(g[~on], img[~on] * np.float32([1, 1, 1, 0]))  # the others pass through (w = 0)
        blended |= on & (np.abs(w[..., :3] - img[..., :3]).max(axis=2) > 1e-3)
    assert blended.mean() > 0.8
    # the full frame at once, for the record: the race-free result against one outcome of the racy one
    want, got, src, dmv = dev(img), torch.zeros(H, W, 4, device="cuda"), dev(img), dev(motion)
    ref.taa(p(want), p(hist), p(d["world"]), p(d["prev_world"]), p(dmv), W, H)
    assert lib.b200_filter_taa(p(src), p(hist), p(dmv), p(got), W, H) == 0, lib.b200_last_error().decode()
    diff = np.abs(host(got) - host(want)).max(axis=2)
    print("TAA, whole frame at once: pixels equal to 1e-4: %.4f, largest difference %.4f, rmse %.5f" % ((diff <= 1e-4).mean(), diff.max(), np.sqrt((diff ** 2).mean())))
    assert (diff <= 1e-4).mean() > 0.8 and np.sqrt((diff ** 2).mean()) < 0.02
    for name, ref_fn, fn in (("unsharpen", ref.unsharpen, lib.b200_filter_unsharpen), ("finish without TAA", ref.finalize_no_taa, lib.b200_filter_finish_no_taa)):
        out_ref, out = torch.zeros(H, W, 4, device="cuda"), torch.zeros(H, W, 4, device="cuda")
        ref_fn(p(want), p(out_ref), W, H)
        assert fn(p(want), p(out), W, H) == 0, lib.b200_last_error().decode()
        close(name, host(out), host(out_ref), float(host(out_ref).max()), share=0.0, atol_rel=1e-6)
        assert (host(out_ref)[0] == 0).all() and (host(out_ref)[:, 0] == 0).all() and host(out_ref)[1:-1, 1:-1, :3].min() > 0  # the border stays unwritten


# ---- the filtered frame through Render() ---------------------------------------------------------------------------------------------
def _filter_scene_core(w, h, spp, **settings):
    import test_wavefront_gpu as twg
    sc = twg._lit_scene()
    rc = twg._lit_core(sc, w, h, spp=spp, filterCore=1, **settings)
    return sc, rc


def test_filter_mode_frame_inputs_match_the_cpu_restatement():
    """What the shade stage hands to the filter (lib/RenderCore_Optix7Filter/kernels/pathtracer.h): direct and indirect accumulator planes
    and the feature buffers of the first sample, against tests/wavefront_ref.py in filter mode (one diffuse bounce, no roulette, rays through
    the pixel corners of the jittered view)."""
    import cases
    import wavefront_ref
    from lighthouse2_b200 import scenes
    w, h, spp = 160, 90, 2
    sc, rc = _filter_scene_core(w, h, spp, TAA=1)
    view = scenes.view_pyramid(width=w, height=h, **scenes.C1_CAMERA)
    rc.Render(view, converge=0)
    st = rc.GetCoreStats()
    direct, indirect = rc.ReadAccumulator(), rc.ReadFilterBuffer(3)
    feat, wpos, dd = rc.ReadFilterBuffer(0), rc.ReadFilterBuffer(1), rc.ReadFilterBuffer(2)
    frame = rc.ReadFrame()
    rc.Shutdown()
    port = cases.load_oracle(dict(sc["case"], rays=None), "port")
    out, stats = {}, {}
    ref = wavefront_ref.render(port, sc["case"]["meshes"], sc["case"]["instances"], view, w, h,
                               lights=wavefront_ref.Lights(sc["area"], sc["point"], sc["spot"], sc["dir"]), sky_image=sc["sky"],
                               materials=sc["materials"], prim_materials=sc["prim_materials"], vertex_normals=sc["vertex_normals"],
                               uvs=sc["uvs"], lods=sc["lods"], textures=sc["textures"], max_path_length=5, spp=spp, bounce_cap=1, roulette=False, stats=stats,
                               filter_out=out, filter_jitter=(0.3, 0.2))  # the first frame's TAA jitter (rendercore.cpp:733-736)
    mean = float(ref[..., :3].mean() + out["indirect"][..., :3].mean())
    for name, got, want in (("direct", direct, ref), ("indirect", indirect, out["indirect"])):
        rmse = float(np.sqrt(np.mean((got[..., :3] - want[..., :3]) ** 2)))
        print(name, "plane: mean", float(want[..., :3].mean()), "rmse", rmse)
        assert rmse <= 1e-4 * mean, (name, rmse, mean)
    assert out["indirect"][..., :3].mean() > 0.02 * mean and np.allclose(direct[..., 3], ref[..., 3], rtol=1e-5)
    assert st["totalExtensionRays"] == sum(stats["ext"]) and st["totalShadowRays"] == sum(stats["shadow"])
    # albedo, packed normal, specular flags and material: as integers; the depth: as a float (the primary rays of the two sides differ in the last bit
    # here and there); the history length (low four bits of w) belongs to the filter's prepare stage, which has already run
    same = (feat[..., :2] == out["features"][..., :2]).all(axis=2) & ((feat[..., 3] >> 4) == (out["features"][..., 3] >> 4))
    print("feature records equal: %.5f" % same.mean(), "kinds:", np.unique(feat[..., 3] >> 6).tolist(), "specular-first pixels:", int(((feat[..., 3] >> 4) & 3 != 0).sum()))
    assert same.mean() > 0.999  # (a quantised normal or albedo component may fall on the other side of a step)
    depth_ok = np.isclose(feat[..., 2].view(np.float32), out["features"][..., 2].view(np.float32), rtol=1e-5)
    print("depths equal to 1e-5: %.5f" % depth_ok.mean())
    assert depth_ok.mean() > 0.999
    same &= depth_ok
    assert ((feat[..., 3] >> 4) & 3 != 0).sum() > 50 and len(np.unique(feat[..., 3] >> 6)) >= 5
    assert np.array_equal(wpos[..., 3].view(np.uint32)[same], out["world_pos"][..., 3].view(np.uint32)[same])
    a, b = wpos[..., :3][same], out["world_pos"][..., :3][same]
    assert (np.abs(a - b).max(axis=1) <= 1e-5 * np.linalg.norm(b, axis=1) + 1e-3).all()  # (a miss stores a point 50000 units away)
    ok = same & np.isfinite(out["delta_depth"]).all(axis=2)
    dd_ok = np.isclose(dd[..., 2:], out["delta_depth"][..., 2:], rtol=1e-3, atol=1e-4).all(axis=2)[ok]
    print("depth derivatives equal to 1e-3: %.5f" % dd_ok.mean(), "non-zero: %.3f" % (dd[..., 2] != 0).mean())
    assert dd_ok.mean() > 0.995 and (dd[..., 2] != 0).mean() > 0.1  # (most of this view is sky)
    assert np.isfinite(frame[1:-1, 1:-1, :3]).all() and frame[1:-1, 1:-1, :3].mean() > 0.01 and st["filterTime"] > 0


@pytest.mark.parametrize("taa", [1, 0])
def test_filtered_frames_follow_the_reference_pipeline(ref, taa):
    """Three frames with a moving camera through Render() with the filter on; after each, the reference's host sequence
    (lib/RenderCore_Optix7Filter/rendercore.cpp:827-861: prepare, three passes, TAA + unsharp mask or the plain finish, the swaps) is replayed
    with the reference's kernels on the frame's own inputs (accumulator planes and feature buffers read back from the core) and must give the
    frame the core produced."""
    from lighthouse2_b200 import scenes
    w, h, spp = 160, 96, 2
    sc, rc = _filter_scene_core(w, h, spp, TAA=taa, clampDirect=2.5, clampIndirect=15.0)
    p = lambda t: t.data_ptr()
    z4 = lambda: torch.zeros(h, w, 4, device="cuda")
    prev_world, prev_moments, history, prev_pixels = z4(), z4(), z4(), z4()
    hx, hy = (0.3, 0.7, 0.2, 0.8), (0.2, 0.8, 0.7, 0.3)
    prev_view, prev_j = None, (0.0, 0.0)
    for k in range(3):
        cam = dict(scenes.C1_CAMERA)
        cam["pos"] = tuple(np.asarray(cam["pos"], np.float32) + np.array([0.04 * k, 0.01 * k, 0.0], np.float32))
        view = scenes.view_pyramid(width=w, height=h, **cam)
        rc.Render(view, converge=1 if k == 0 else 0)
        frame = rc.ReadFrame()
        acc = np.stack([rc.ReadAccumulator(), rc.ReadFilterBuffer(3)])
        feat_after, wpos, dd = rc.ReadFilterBuffer(0), rc.ReadFilterBuffer(1), rc.ReadFilterBuffer(2)
        vnow = np.array([view["pos"], view["p1"], view["p2"], view["p3"]], np.float32)
        if prev_view is None:
            prev_view = vnow
        j = (hx[k], hy[k]) if taa else (0.0, 0.0)
        # the feature buffer as it was BEFORE prepare: the history length is what the reference's own prepare left last frame
        feat_in = feat_after.copy()
        feat_in[..., 3] = (feat_in[..., 3] & ~np.uint32(15)) | (hist_len if k else np.zeros((h, w), np.uint32))
        d_feat, d_acc, d_world, d_dd = dev(feat_in), dev(acc), dev(wpos), dev(dd)
        shading, motion, moments, f_out, f_in2 = z4(), torch.zeros(h, w, 2, device="cuda"), z4(), z4(), z4()
        ref.prepare(p(d_acc), p(d_feat), p(d_world), p(prev_world), p(shading), p(motion), p(moments), p(prev_moments), p(d_dd), prev_view, j, prev_j, w, h, spp, 2.5, 15.0, k != 0)
        hist_len = host(d_feat, np.uint32)[..., 3] & np.uint32(15)
        assert np.array_equal(hist_len, feat_after[..., 3] & np.uint32(15)) or (hist_len == (feat_after[..., 3] & 15)).mean() > 0.998
        ref.apply(p(d_feat), p(prev_world), p(d_world), p(d_dd), p(motion), p(moments), p(shading), p(history), p(f_out), w, h, 1, 0)
        ref.apply(p(d_feat), p(prev_world), p(d_world), p(d_dd), p(motion), p(moments), p(f_out), p(history), p(f_in2), w, h, 2, 0)
        ref.apply(p(d_feat), p(prev_world), p(d_world), p(d_dd), p(motion), p(moments), p(f_in2), p(history), p(shading), w, h, 3, 1)
        want = z4()
        if taa:
            # (the reference blends in place, with a race on its 3x3 window: its result is one of many; the core's own blend is checked against
            # it in test_taa_...; here the core's TAA history stands in so that the pipeline around it is held to the strict bar)
            blended = dev(rc.ReadFilterBuffer(7))
            ref.unsharpen(p(blended), p(want), w, h)
            prev_pixels = blended
        else:
            ref.finalize_no_taa(p(shading), p(want), w, h)
            prev_pixels = shading
        close("frame %d (TAA %d)" % (k, taa), frame, host(want), float(max(host(want)[..., :3].max(), 1e-3)), share=0.004)
        packed_close("first reconstruction pass %d" % k, rc.ReadFilterBuffer(6), host(f_out), share=0.004)
        if taa and k:
            # the blend itself: the core's history image against the reference's racy in-place result on the same inputs
            racy = shading.clone()
            ref.taa(p(racy), p(last_blended), p(d_world), p(prev_world), p(motion), w, h)
            diff = np.abs(host(racy) - host(blended)).max(axis=2)
            print("TAA blend, frame %d: pixels equal to 1e-4: %.4f" % (k, (diff <= 1e-4).mean()))
            assert (diff <= 1e-4).mean() > 0.8
        last_blended = prev_pixels
        prev_world, prev_moments, history, prev_view, prev_j = d_world, moments, f_out, vnow, j
        assert frame[1:-1, 1:-1, :3].mean() > 0.01
    rc.Shutdown()


def test_filtered_frames_do_not_repeat_their_noise():
    """Every filtered frame starts from sample 0 with the same seed (lib/RenderCore_Optix7Filter/rendercore.cpp:687); in the reference the frames
    differ through the blue-noise slot that moves on by one per frame (:860, pathtracer.h:248-249, 306-307).  With a table set, the second
    frame of a still camera must be the restatement's frame with blue_slot = 1; without a table the salt changes instead."""
    import cases
    import wavefront_ref
    from lighthouse2_b200 import scenes
    w, h, spp = 96, 54, 1
    view = scenes.view_pyramid(width=w, height=h, **scenes.C1_CAMERA)
    table = np.random.default_rng(78).integers(0, 256, 5 * 65536).astype(np.uint32)
    for with_table in (True, False):
        sc, rc = _filter_scene_core(w, h, spp, TAA=1)  # (with the TAA jitter: corner rays of the unshifted view run along the mesh's edges in the middle column)
        if with_table:
            rc.SetBlueNoise(table)
        planes = []
        for k in range(2):
            rc.Render(view, converge=0)
            planes.append((rc.ReadAccumulator(), rc.ReadFilterBuffer(3)))
        rc.Shutdown()
        assert np.abs(planes[0][0][..., :3] - planes[1][0][..., :3]).mean() > 1e-3  # the noise moved
        port = cases.load_oracle(dict(sc["case"], rays=None), "port")
        out = {}
        kw = dict(blue_noise=table, blue_slot=1) if with_table else dict(salt=2 * 7907)
        ref = wavefront_ref.render(port, sc["case"]["meshes"], sc["case"]["instances"], view, w, h,
                                   lights=wavefront_ref.Lights(sc["area"], sc["point"], sc["spot"], sc["dir"]), sky_image=sc["sky"],
                                   materials=sc["materials"], prim_materials=sc["prim_materials"], vertex_normals=sc["vertex_normals"],
                                   uvs=sc["uvs"], lods=sc["lods"], textures=sc["textures"], max_path_length=5, spp=spp, bounce_cap=1, roulette=False,
                                   filter_out=out, filter_jitter=(0.7, 0.8), **kw)  # the second frame's jitter (rendercore.cpp:733-736)
        mean = float(ref[..., :3].mean() + out["indirect"][..., :3].mean())
        for name, got, want in (("direct", planes[1][0], ref), ("indirect", planes[1][1], out["indirect"])):
            rmse = float(np.sqrt(np.mean((got[..., :3] - want[..., :3]) ** 2)))
            print("second frame,", "table" if with_table else "no table", name, "rmse", rmse, "of mean", mean)
            assert rmse <= 1e-4 * mean
