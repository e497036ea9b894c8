"""GPU tests of the full wavefront frame (Render through the C ABI): image vs the CPU restatement that uses the oracle's
traversal, tile sharding == single core, fused connect == any-hit + add, CoreStats bookkeeping."""
import numpy as np
import pytest

import cases
import wavefront_ref
from lighthouse2_b200 import core, lights, scenes

pytestmark = pytest.mark.gpu


def _scene():
    grid = scenes.grid_mesh(16, 32)
    sph = scenes.uv_sphere(24, 13, 0.35)
    M = np.tile(np.eye(4, dtype=np.float32).reshape(16), (4, 1))
    M[1, [3, 7, 11]] = (-0.5, 0.6, 0.2)
    M[2, [3, 7, 11]] = (0.4, 0.5, -0.3)
    M[3, [3, 7, 11]] = (0.0, 0.9, 0.5)
    inst = [(0, M[0]), (1, M[1]), (1, M[2]), (1, M[3])]
    return dict(name="wf", meshes=[grid, sph], instances=inst)


LIGHT = ((0.5, 3.0, -1.0), (40.0, 36.0, 30.0))
SKY = (0.3, 0.4, 0.6)


def _core(case, w, h, spp=1, max_len=4, **settings):
    rc = cases.load_core(core.RenderCoreB200(0), dict(case, rays=None))
    rc.SetTarget(w, h, spp)
    rc.SetPointLight(*LIGHT)
    rc.SetSky(SKY)
    rc.Setting("maxPathLength", max_len)
    rc.Setting("jitter", 0)
    for k, v in settings.items():
        rc.Setting(k, v)
    return rc


def test_frame_matches_cpu_restatement():
    """Image RMSE vs the CPU wavefront loop that uses the reference's Traverse for every query.
    Stated bar: RMSE <= 1e-4 -- asserted here as 1e-4 of the mean radiance (the mean is below 1),
    with NO pixel off by more than 1 %.  Both sides draw every bounce direction from the same separately-rounded sin / cos
    (wavefront.cu::SinCosDet == wavefront_ref.sincos_det), so no path forks; what is left is summation order (measured: 7e-9)."""
    case = _scene()
    w, h = 192, 108
    view = scenes.view_pyramid(width=w, height=h, **scenes.C1_CAMERA)
    rc = _core(case, w, h)
    rc.Render(view, converge=1)
    acc = rc.ReadAccumulator()[..., :3]
    port = cases.load_oracle(dict(case, rays=None), "port")
    ref4 = wavefront_ref.render(port, case["meshes"], case["instances"], view, w, h, light=LIGHT, sky=SKY, max_path_length=4)
    ref = ref4[..., :3]
    depth = rc.ReadAccumulator()[..., 3]
    assert np.allclose(depth, ref4[..., 3], rtol=1e-5)  # primary hit distance (10000 on a miss), pathtracer.h:74
    mean = float(ref.mean())
    rmse = float(np.sqrt(np.mean((acc - ref) ** 2)))
    big = np.abs(acc - ref).max(axis=2) > 0.01 * mean
    print("wavefront image: mean", mean, "rmse", rmse, "pixels off by >1%:", int(big.sum()), "of", w * h)
    assert rmse <= 1e-4 * mean, (rmse, mean)
    assert big.sum() == 0
    st = rc.GetCoreStats()
    assert st["primaryRayCount"] == w * h
    assert st["totalExtensionRays"] >= st["primaryRayCount"] + st["bounce1RayCount"]
    assert st["totalShadowRays"] > 0 and st["renderTime"] > 0
    frame = rc.ReadFrame()
    assert np.allclose(frame[..., :3], acc, rtol=1e-6)
    rc.Shutdown()


@pytest.mark.parametrize("shards", [2, 3])
def test_tile_sharding_adds_up_bit_exact(shards):
    """Interleaved-tile shards rendered by separate cores sum to the single-core accumulator bit for bit
    (disjoint pixels, per-path random streams independent of the sharding) -- the multi-GPU gather contract."""
    case = _scene()
    w, h = 200, 100  # not a multiple of the 32x8 tile: partial tiles are clipped
    view = scenes.view_pyramid(width=w, height=h, **scenes.C1_CAMERA)
    rc = _core(case, w, h)
    rc.Render(view, converge=1)
    full = rc.ReadAccumulator()
    rc.Shutdown()
    total = np.zeros_like(full)
    owned = np.zeros((h, w), np.int32)
    for i in range(shards):
        r = _core(case, w, h, shardCount=shards, shardIndex=i)
        r.Render(view, converge=1)
        part = r.ReadAccumulator()
        owned += (np.abs(part[..., :3]).sum(axis=2) > 0)
        total += part
        r.Shutdown()
    assert owned.max() <= 1, "shards overlap"
    assert np.array_equal(total.view(np.uint32), full.view(np.uint32))


def test_sample_sharding_and_convergence():
    """spp = 2 in one frame == two 1-spp cores with sampleOffset 0 and 1 (what the spp-sharded multi-GPU mode sums)."""
    case = _scene()
    w, h = 128, 64
    view = scenes.view_pyramid(width=w, height=h, **scenes.C1_CAMERA)
    rc = _core(case, w, h, spp=2)
    rc.Render(view, converge=1)
    both = rc.ReadAccumulator()[..., :3]
    rc.Shutdown()
    parts = []
    for off in (0, 1):
        r = _core(case, w, h, spp=1, sampleOffset=off)
        r.Render(view, converge=1)
        parts.append(r.ReadAccumulator()[..., :3])
        r.Shutdown()
    assert np.allclose(parts[0] + parts[1], both, rtol=1e-5, atol=1e-6)  # atomic float adds of 2 samples: order may differ
    # Converge (0) keeps accumulating, Restart (1) clears (common_classes.h:38-42)
    r = _core(case, w, h, spp=1)
    r.Render(view, converge=1)
    a1 = r.ReadAccumulator()[..., :3].copy()
    r.Render(view, converge=0)
    a2 = r.ReadAccumulator()[..., :3].copy()
    assert a2.sum() > 1.5 * a1.sum()
    r.Render(view, converge=1)
    assert np.array_equal(r.ReadAccumulator()[..., :3], a1)
    r.Shutdown()


def test_fused_connect_equals_any_hit_plus_add():
    import torch
    case = cases.case_soup(n_tris=20000, n_rays=50000, seed=61)
    O, D = case["rays"]
    rc = cases.load_core(core.RenderCoreB200(0), case)
    rng = np.random.default_rng(3)
    tmax = rng.uniform(0.5, 8.0, len(O)).astype(np.float32)
    O4, D4 = core.pack_rays(O, D, tmin=1e-4, tmax=tmax)
    occ = rc.TraceAny(O4, D4)
    npx = 4096
    pix = rng.integers(0, npx, len(O)).astype(np.uint32)
    E = np.empty((len(O), 4), np.float32)
    E[:, :3] = rng.integers(1, 8, (len(O), 3)).astype(np.float32)  # small integers: float adds are exact in any order
    E[:, 3] = pix.view(np.float32)
    dev = torch.device("cuda", 0)
    dO, dD, dE = (torch.from_numpy(a).to(dev) for a in (O4, D4, E))
    acc = torch.zeros((npx, 4), dtype=torch.float32, device=dev)
    rc.SetStream(torch.cuda.current_stream().cuda_stream)
    rc.TraceConnectDevice(dO.data_ptr(), dD.data_ptr(), len(O), dE.data_ptr(), acc.data_ptr())
    torch.cuda.synchronize()
    expect = np.zeros((npx, 3), np.float64)
    np.add.at(expect, pix[~occ], E[~occ, :3].astype(np.float64))
    assert np.array_equal(acc.cpu().numpy()[:, :3], expect.astype(np.float32))
    rc.Shutdown()


def test_probe_reports_primary_hit():
    case = _scene()
    w, h = 160, 90
    view = scenes.view_pyramid(width=w, height=h, **scenes.C1_CAMERA)
    rc = _core(case, w, h)
    rc._chk(rc.lib.b200_set_probe(rc.h, w // 2, h // 2))
    rc.Render(view, converge=1)
    st = rc.GetCoreStats()
    O, D = scenes.primary_rays(view, w, h)
    i = (h // 2) * w + w // 2
    port = cases.load_oracle(dict(case, rays=None), "port")
    t, prim, inst = port.trace(O[i:i + 1], D[i:i + 1])
    assert st["probedTriid"] == prim[0] and st["probedInstid"] == inst[0]
    assert abs(st["probedDist"] - t[0]) <= 1e-5 * t[0]
    rc.Shutdown()


def _lit_scene():
    """_scene() plus an emissive quad above it; materials: 0 grey diffuse, 1 red glossy, 2 mirror, 3 emitter."""
    case = _scene()
    quad = np.array([[[-0.4, 0, -0.3], [0.4, 0, -0.3], [0.4, 0, 0.3]], [[-0.4, 0, -0.3], [0.4, 0, 0.3], [-0.4, 0, 0.3]]], np.float32)
    Mq = np.eye(4, dtype=np.float32).reshape(16)
    Mq[[3, 7, 11]] = (0.1, 1.6, 0.0)
    case["meshes"] = case["meshes"] + [quad]
    case["instances"] = case["instances"] + [(2, Mq)]
    grid, sph = case["meshes"][0], case["meshes"][1]
    pm = [np.where((np.arange(len(grid)) // 2) % 5 == 0, 1, 6).astype(np.uint32),  # floor: textured with alpha holes, every fifth quad glossy red
          np.choose((np.arange(len(sph)) // 7) % 4, [0, 2, 4, 5]).astype(np.uint32),  # bands on the spheres: diffuse, mirror, glass, smooth frosted
          np.full(len(quad), 3, np.uint32)]
    # material table (b200_set_material_table): albedo rgb, roughness | transmission, eta, flags (1 SMOOTH + 2 HASALPHA), - |
    # extinction rgb, diffuse texture or -1 | uv scale, uv offset
    mats = np.array([[0.8, 0.8, 0.8, 1.0, 0, 1, 0, 0, 0, 0, 0, -1, 1, 1, 0, 0],             # 0 grey diffuse
                     [0.9, 0.2, 0.15, 0.3, 0, 1, 0, 0, 0, 0, 0, -1, 1, 1, 0, 0],            # 1 red glossy
                     [0.95, 0.95, 0.95, 0.0, 0, 1, 0, 0, 0, 0, 0, -1, 1, 1, 0, 0],          # 2 mirror
                     [12.0, 11.0, 10.0, 1.0, 0, 1, 0, 0, 0, 0, 0, -1, 1, 1, 0, 0],          # 3 emitter
                     [0.95, 0.97, 1.0, 1.0, 1.0, 1.0 / 1.5, 1, 0, 0.75, 0.25, 0.125, -1, 1, 1, 0, 0],   # 4 glass (smooth normals, absorbing)
                     [0.7, 0.8, 0.6, 0.8, 0.4, 1.0 / 1.3, 1, 0, 0.25, 0.25, 0.25, -1, 1, 1, 0, 0],      # 5 frosted: 40% transmission, smooth
                     [1.0, 0.9, 0.8, 1.0, 0, 1, 2, 0, 0, 0, 0, 0, 1.5, 2.0, 0.25, 0.125]],              # 6 mapped floor, HASALPHA
                    np.float32)
    # texture 0: 16x16 ARGB32 with 5 MIP levels; a fifth of the base texels are transparent
    rng_t = np.random.default_rng(5)
    levels = []
    for k in range(5):
        side = 16 >> k
        rgba = rng_t.integers(60, 256, (side * side, 4)).astype(np.uint32)
        rgba[:, 3] = np.where(rng_t.random(side * side) < 0.2, 0, 255)
        levels.append(rgba[:, 0] | (rgba[:, 1] << 8) | (rgba[:, 2] << 16) | (rgba[:, 3] << 24))
    textures = [(16, 16, np.concatenate(levels).astype(np.uint32), 5)]
    guv = np.stack([grid[:, :, 0] * 0.7 + 0.3, grid[:, :, 2] * 0.9 - 0.1], axis=2).astype(np.float32)  # uv from the floor's x, z
    uvs = [guv, None, None]
    lods = [(0.3 + 0.2 * np.sin(np.arange(len(grid)))).astype(np.float32) + 6.0, None, None]  # CoreTri::LOD (+6: the cone is narrow at this scale)
    sphn = (sph / np.linalg.norm(sph, axis=2, keepdims=True)).astype(np.float32)  # the sphere is centred on its origin
    vnormals = [None, sphn, None]
    world = quad + np.array([0.1, 1.6, 0.0], np.float32)
    n = np.cross(world[0, 1] - world[0, 0], world[0, 2] - world[0, 0])
    assert n[1] < 0, "the emitter must face down"
    area = lights.stack([lights.area_light(world[i, 0], world[i, 1], world[i, 2], mats[3, :3], i, 4) for i in range(2)], 24)
    point = lights.stack([lights.point_light((0.5, 3.0, -1.0), (20.0, 18.0, 15.0))], 8)
    spot = lights.stack([lights.spot_light((-1.5, 2.0, -1.0), (0.6, -0.8, 0.4), 0.95, 0.8, (30.0, 30.0, 40.0))], 12)
    direc = lights.stack([lights.directional_light((0.3, -1.0, 0.2), (0.5, 0.45, 0.4))], 8)
    rng = np.random.default_rng(11)
    sky = (rng.random((8, 16, 3)) * np.array([0.3, 0.4, 0.6])).astype(np.float32)
    return dict(case=case, prim_materials=pm, materials=mats, vertex_normals=vnormals, uvs=uvs, lods=lods, textures=textures, area=area, point=point, spot=spot, dir=direc, sky=sky)


def _lit_core(sc, w, h, spp=1, max_len=5, **settings):
    case = sc["case"]
    rc = core.RenderCoreB200(0)
    for i, m in enumerate(case["meshes"]):
        rc.SetGeometry(i, m, core.core_tris(m, sc["prim_materials"][i], sc["vertex_normals"][i], sc["uvs"][i], sc["lods"][i]))
    for i, (mesh, mat) in enumerate(case["instances"]):
        rc.SetInstance(i, mesh, mat)
    rc.SetInstance(len(case["instances"]), -1)
    rc.UpdateToplevel()
    rc.SetTarget(w, h, spp)
    rc.SetTextures(sc["textures"])
    rc.SetMaterialTable(sc["materials"])
    rc.SetLights(sc["area"], sc["point"], sc["spot"], sc["dir"])
    rc.SetSkyData(sc["sky"])
    rc.Setting("maxPathLength", max_len)
    rc.Setting("jitter", 0)
    for k, v in settings.items():
        rc.Setting(k, v)
    return rc


@pytest.mark.parametrize("cap,roulette", [(0, 1), (1, 1), (0, 0)])
def test_lit_frame_matches_cpu_restatement(cap, roulette):
    """The full shade stage: area + point + spot + directional lights picked by potential contribution,
    MIS on emissive hits, glossy / mirror / diffuse Lambert materials, Russian roulette, skydome image, intensity clamp.
    Image vs the numpy restatement that traces through the CPU oracle.  cap = 1 is the reference's ENOUGH_BOUNCES.
    Bar: RMSE <= 1% of the mean radiance and < 2% of pixels off by more than 2% (discrete choices -- light pick,
    roulette, specular/diffuse lobe -- fork on last-bit differences of the GPU's fast intrinsics)."""
    sc = _lit_scene()
    w, h = 160, 90
    view = scenes.view_pyramid(width=w, height=h, **scenes.C1_CAMERA)
    rc = _lit_core(sc, w, h, spp=2, diffuseBounceCap=cap, russianRoulette=roulette)
    rc.Render(view, converge=1)
    acc = rc.ReadAccumulator()
    st = rc.GetCoreStats()
    port = cases.load_oracle(dict(sc["case"], rays=None), "port")
    stats = {}
    ref = wavefront_ref.render(port, sc["case"]["meshes"], sc["case"]["instances"], view, w, h,
                               lights=wavefront_ref.Lights(sc["area"], sc["point"], sc["spot"], sc["dir"]), sky_image=sc["sky"],
                               materials=sc["materials"], prim_materials=sc["prim_materials"], vertex_normals=sc["vertex_normals"],
                               uvs=sc["uvs"], lods=sc["lods"], textures=sc["textures"], max_path_length=5, spp=2,
                               bounce_cap=cap, roulette=bool(roulette), stats=stats)
    mean = float(ref[..., :3].mean())
    rmse = float(np.sqrt(np.mean((acc[..., :3] - ref[..., :3]) ** 2)))
    big = np.abs(acc[..., :3] - ref[..., :3]).max(axis=2) > 0.02 * mean
    print("lit image: mean", mean, "rmse", rmse, "pixels off by >2%:", int(big.sum()), "of", w * h,
          "rays gpu", st["totalExtensionRays"], st["totalShadowRays"], "ref", sum(stats["ext"]), sum(stats["shadow"]))
    assert np.isfinite(acc).all() and mean > 0.05
    assert rmse <= 1e-4 * mean, (rmse, mean)  # the stated image bar; measured 9e-8
    assert big.sum() == 0
    assert np.allclose(acc[..., 3], ref[..., 3], rtol=1e-5)
    # exactly the same number of rays: no path forks
    assert st["totalExtensionRays"] == sum(stats["ext"]) and st["totalShadowRays"] == sum(stats["shadow"])
    rc.Shutdown()


def test_hdr_texture_matches_cpu_restatement():
    """ARGB128 (HDR) storage (common_classes.h:240-269; SyncStorageType, rendercore_optix7/rendercore.cpp:434-505): the floor's diffuse
    map as float4 texels, values above 1 and an alpha channel that cuts holes, beside an ARGB32 texture in the same pool (so that
    the 16-byte alignment of the HDR texels is exercised).  The sampler is FetchTexel's ARGB128 branch (sampling_shared.h:55-59)."""
    sc = _lit_scene()
    rng = np.random.default_rng(5)
    hdr = np.empty((16 * 16, 4), np.float32)
    hdr[:, :3] = rng.uniform(0.05, 2.5, (256, 3))
    hdr[:, 3] = np.where(rng.uniform(size=256) < 0.2, 0.0, 1.0)
    # texture 0: three odd ARGB32 texels (never sampled: keeps the HDR texture off a 16-byte boundary before padding); texture 1: the HDR map
    sc["textures"] = [(3, 1, np.array([0xff102030, 0xff405060, 0xff708090], np.uint32), 1), (16, 16, hdr, 1, 1)]
    sc["materials"] = sc["materials"].copy()
    sc["materials"][sc["materials"][:, 11] == 0, 11] = 1
    w, h = 160, 90
    view = scenes.view_pyramid(width=w, height=h, **scenes.C1_CAMERA)
    rc = _lit_core(sc, w, h, spp=2)
    st0 = rc.GetCoreStats()
    assert (st0["argb32TexelCount"], st0["argb128TexelCount"], st0["nrm32TexelCount"]) == (3, 256, 0)
    rc.Render(view, converge=1)
    acc, st = rc.ReadAccumulator(), rc.GetCoreStats()
    rc.Shutdown()
    port = cases.load_oracle(dict(sc["case"], rays=None), "port")
    stats = {}
    ref = wavefront_ref.render(port, sc["case"]["meshes"], sc["case"]["instances"], view, w, h,
                               lights=wavefront_ref.Lights(sc["area"], sc["point"], sc["spot"], sc["dir"]), sky_image=sc["sky"],
                               materials=sc["materials"], prim_materials=sc["prim_materials"], vertex_normals=sc["vertex_normals"],
                               uvs=sc["uvs"], lods=sc["lods"], textures=sc["textures"], max_path_length=5, spp=2, stats=stats)
    mean = float(ref[..., :3].mean())
    rmse = float(np.sqrt(np.mean((acc[..., :3] - ref[..., :3]) ** 2)))
    big = np.abs(acc[..., :3] - ref[..., :3]).max(axis=2) > 0.02 * mean
    print("hdr-mapped image: mean", mean, "rmse", rmse, "pixels off by >2%:", int(big.sum()), "rays", st["totalExtensionRays"], sum(stats["ext"]))
    assert rmse <= 1e-4 * mean and big.sum() == 0
    assert st["totalExtensionRays"] == sum(stats["ext"]) and st["totalShadowRays"] == sum(stats["shadow"])
    # and the map is really sampled: the ARGB32 version of the scene gives another image
    base = _lit_scene()
    rc = _lit_core(base, w, h, spp=2)
    rc.Render(view, converge=1)
    ldr = rc.ReadAccumulator()
    rc.Shutdown()
    assert np.abs(ldr[..., :3] - acc[..., :3]).mean() > 0.01 * mean


def test_blue_noise_sampling_matches_cpu_restatement():
    """BLUENOISE of the reference's cores (pathtracer.h:168-222, .optix.cu:66-79): with a table set, the first 256 samples
    draw the lens point and the BSDF sample from it and the first two the light sample.  The table here is synthetic (the
    reference compiles Heitz' tables in; this core takes them from the host): the check is the indexing, sample 0..2 so
    that both the `< 2` and the `< 256` rule are crossed."""
    sc = _lit_scene()
    w, h, spp = 160, 90, 3
    view = scenes.view_pyramid(width=w, height=h, **scenes.C1_CAMERA)
    view["aperture"] = 0.02
    table = np.random.default_rng(77).integers(0, 256, 5 * 65536).astype(np.uint32)
    rc = _lit_core(sc, w, h, spp=spp)
    rc.SetBlueNoise(table)
    rc.Render(view, converge=1)
    acc, st = rc.ReadAccumulator(), rc.GetCoreStats()
    rc.SetBlueNoise(None)
    rc.Render(view, converge=1)
    plain = rc.ReadAccumulator()
    rc.Shutdown()
    port = cases.load_oracle(dict(sc["case"], rays=None), "port")
    stats = {}
    ref = wavefront_ref.render(port, sc["case"]["meshes"], sc["case"]["instances"], view, w, h,
                               lights=wavefront_ref.Lights(sc["area"], sc["point"], sc["spot"], sc["dir"]), sky_image=sc["sky"],
                               materials=sc["materials"], prim_materials=sc["prim_materials"], vertex_normals=sc["vertex_normals"],
                               uvs=sc["uvs"], lods=sc["lods"], textures=sc["textures"], max_path_length=5, spp=spp, stats=stats, blue_noise=table)
    mean = float(ref[..., :3].mean())
    rmse = float(np.sqrt(np.mean((acc[..., :3] - ref[..., :3]) ** 2)))
    big = np.abs(acc[..., :3] - ref[..., :3]).max(axis=2) > 0.02 * mean
    print("blue noise image: mean", mean, "rmse", rmse, "pixels off by >2%:", int(big.sum()), "rays", st["totalExtensionRays"], sum(stats["ext"]))
    assert rmse <= 1e-4 * mean and big.sum() == 0  # measured 4e-6
    assert st["totalExtensionRays"] == sum(stats["ext"])
    assert np.abs(plain[..., :3] - acc[..., :3]).mean() > 0.01 * mean  # the table is really used


def test_emitter_seen_directly_and_one_sided():
    """A camera ray that hits the emitter's front adds its radiance (clamped by clampValue); its back is black."""
    sc = _lit_scene()
    w, h = 64, 64
    for eye_y, expect_lit in ((1.0, True), (2.5, False)):  # from below (front side) / from above (back side)
        view = scenes.view_pyramid(width=w, height=h, pos=(0.1, eye_y, -0.1), target=(0.1, 1.6, 0.0), fov_deg=30.0)
        rc = _lit_core(sc, w, h, clampValue=1e6)
        rc.Render(view, converge=1)
        centre = rc.ReadAccumulator()[h // 2, w // 2, :3]
        if expect_lit:
            assert np.allclose(centre, sc["materials"][3, :3], rtol=1e-6)
        else:
            assert np.all(centre == 0)
        rc.Shutdown()


def test_lens_and_distortion_match_cpu_restatement():
    """Thin-lens sampling on the 9-blade aperture and barrel distortion of the primary rays (.optix.cu:51-97), checked
    through the primary-hit depth channel and the image.  Depth differs where a ray passes a silhouette within the
    last bits of sin/cos, hence the small allowance."""
    case = _scene()
    w, h = 160, 90
    view = dict(scenes.view_pyramid(width=w, height=h, **scenes.C1_CAMERA), aperture=0.02, distortion=0.05)
    rc = _core(case, w, h, spp=2, max_len=2)
    rc.Render(view, converge=1)
    acc = rc.ReadAccumulator()
    port = cases.load_oracle(dict(case, rays=None), "port")
    ref = wavefront_ref.render(port, case["meshes"], case["instances"], view, w, h, light=LIGHT, sky=SKY, max_path_length=2, spp=2)
    pin = wavefront_ref.render(port, case["meshes"], case["instances"], dict(view, aperture=0.0, distortion=0.0), w, h, light=LIGHT, sky=SKY,
                               max_path_length=2, spp=2)
    off = ~np.isclose(acc[..., 3], ref[..., 3], rtol=1e-4)
    print("lens/distortion: depth mismatches", int(off.sum()), "of", w * h, "; differs from pinhole in",
          int((~np.isclose(pin[..., 3], ref[..., 3], rtol=1e-4)).sum()))
    assert off.mean() < 2e-3
    assert (~np.isclose(pin[..., 3], ref[..., 3], rtol=1e-4)).mean() > 0.2  # the lens and the distortion do move the rays (most of the rest is sky)
    mean = float(ref[..., :3].mean())
    assert float(np.sqrt(np.mean((acc[..., :3] - ref[..., :3]) ** 2))) <= 2e-2 * mean
    rc.Shutdown()


def test_async_frame_readback_equals_the_blocking_one():
    """b200_read_frame_async / b200_wait_frame: the frame leaves beside the next Render().  Two frames in a row into two page-locked
    buffers, the second asked for while the first may still be on the bus; both equal what the blocking call returns."""
    import torch
    rc = core.RenderCoreB200(0)
    rc.SetGeometry(0, scenes.sphere_field(20_000)); rc.SetInstance(0, 0); rc.SetInstance(1, -1); rc.UpdateToplevel()
    w, h = 320, 180
    rc.SetTarget(w, h, 2); rc.SetPointLight((0.5, 30.0, -10.0), (4e3, 3.6e3, 3e3)); rc.SetSky((0.3, 0.4, 0.6)); rc.Setting("maxPathLength", 3)
    views = [scenes.view_pyramid(width=w, height=h, **scenes.C2_CAMERA), scenes.view_pyramid(width=w, height=h, **dict(scenes.C2_CAMERA, fov_deg=35.0))]
    want = []
    for v in views:
        rc.Render(v, converge=1)
        want.append(rc.ReadFrame())
    bufs = [torch.empty((h, w, 4), dtype=torch.float32).pin_memory() for _ in views]
    for v, b in zip(views, bufs):
        rc.Render(v, converge=1)
        rc.ReadFrameAsync(b.data_ptr())
    rc.WaitFrame()
    rc.WaitFrame()  # (nothing pending: returns at once)
    for b, wnt in zip(bufs, want):
        assert np.array_equal(b.numpy().view(np.uint32), wnt.view(np.uint32))
    assert not np.array_equal(want[0], want[1]) and want[0][..., :3].mean() > 0.01
    rc.Shutdown()
