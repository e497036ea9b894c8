"""The drop-in boundary end to end: a C++ host (tests/cpp/plugin_host.cpp) loads libRenderCore_B200.so exactly as the
reference's loader does (dlopen + CreateCore, lib/RenderSystem/core_api_base.cpp:97-132), drives it through the CoreAPI_Base
vtable in RenderSystem's call order with the reference's own structs (CoreTri, CoreMaterial, CoreLightTri, ...), and must
produce the same frame, bit for bit, as the flat C ABI driven from Python."""
import os
import struct
import subprocess

import numpy as np
import pytest

from lighthouse2_b200 import core, scenes
from test_wavefront_gpu import _lit_scene

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def _write_scene(path, sc, view, w, h, spp, frames, probe):
    case = sc["case"]
    with open(path, "wb") as f:
        f.write(struct.pack("<6i", w, h, spp, frames, probe[0], probe[1]))
        vp = np.zeros(17, np.float32)
        vp[0:3], vp[3:6], vp[6:9], vp[9:12] = view["pos"], view["p1"], view["p2"], view["p3"]
        vp[12:17] = (0.0, 0.01, 0.01, 1.0, 0.0)
        f.write(vp.tobytes())
        sky = sc["sky"]
        f.write(struct.pack("<2i", sky.shape[1], sky.shape[0]))
        f.write(np.ascontiguousarray(sky, np.float32).tobytes())
        f.write(struct.pack("<i", len(sc["textures"])))
        for (tw, th, texels, mips) in sc["textures"]:
            f.write(struct.pack("<4i", tw, th, len(texels), mips))
            f.write(np.ascontiguousarray(texels, np.uint32).tobytes())
        mats = np.ascontiguousarray(sc["materials"], np.float32)  # [n, 16]: see _lit_scene
        f.write(struct.pack("<i", len(mats)))
        f.write(mats.tobytes())
        f.write(struct.pack("<i", len(case["meshes"])))
        for i, m in enumerate(case["meshes"]):
            v4 = np.ones((len(m) * 3, 4), np.float32)
            v4[:, :3] = m.reshape(-1, 3)
            f.write(struct.pack("<i", len(m)))
            f.write(v4.tobytes())
            f.write(core.core_tris(m, sc["prim_materials"][i], sc["vertex_normals"][i], sc["uvs"][i], sc["lods"][i]).tobytes())
        f.write(struct.pack("<i", len(case["instances"])))
        for mesh, mat in case["instances"]:
            f.write(struct.pack("<i", mesh))
            f.write(np.ascontiguousarray(np.eye(4, dtype=np.float32).reshape(16) if mat is None else mat, np.float32).tobytes())
        f.write(struct.pack("<4i", len(sc["area"]), len(sc["point"]), len(sc["spot"]), len(sc["dir"])))
        for k in ("area", "point", "spot", "dir"):
            f.write(np.ascontiguousarray(sc[k], np.float32).tobytes())


def _run_host_and_compare(tmp_path, exe, env=None):
    sc = _lit_scene()
    w, h, spp, frames = 160, 90, 1, 2
    probe = (w // 2, h // 2)
    view = scenes.view_pyramid(width=w, height=h, **scenes.C1_CAMERA)
    scene_bin, out_bin = str(tmp_path / "scene.bin"), str(tmp_path / "out.bin")
    _write_scene(scene_bin, sc, view, w, h, spp, frames, probe)
    r = subprocess.run([exe, core.LIB_PATH, scene_bin, out_bin], capture_output=True, text=True, timeout=300, env=env)
    assert r.returncode == 0, r.stderr + r.stdout
    raw = open(out_bin, "rb").read()
    n = w * h * 4
    frame = np.frombuffer(raw, np.float32, n).reshape(h, w, 4)
    acc = np.frombuffer(raw, np.float32, n, offset=n * 4).reshape(h, w, 4)
    stats = raw[2 * n * 4:2 * n * 4 + 104]
    name = raw[2 * n * 4 + 104:].split(b"\0")[0].decode()
    # CoreStats (core_api_base.h:30-61): char* @0, SMcount..VRAM @8, texel counts @24, bvhBuildTime @36, totalRays @40, ... probed @92
    sm, cc_major = struct.unpack_from("<2I", stats, 8)
    total, ext, shadow = struct.unpack_from("<3I", stats, 40)
    render_time, primary = struct.unpack_from("<fI", stats, 52)
    probed_inst, probed_tri, probed_dist = struct.unpack_from("<iif", stats, 92)
    assert sm == 148 and cc_major == 10 and "B200" in name
    assert primary == w * h and total == ext + shadow and 0 < render_time < 1.0  # seconds

    # the same frames through the flat C ABI
    from test_wavefront_gpu import _lit_core
    rc = _lit_core(sc, w, h, spp=spp, max_len=4)
    rc._chk(rc.lib.b200_set_probe(rc.h, *probe))
    rc.Render(view, converge=1)
    rc.Render(view, converge=0)
    ref_acc, ref_frame, st = rc.ReadAccumulator(), rc.ReadFrame(), rc.GetCoreStats()
    rc.Shutdown()
    assert np.array_equal(acc.view(np.uint32), ref_acc.view(np.uint32))
    assert np.array_equal(frame.view(np.uint32), ref_frame.view(np.uint32))
    assert np.allclose(frame[..., :3] * frames * spp, acc[..., :3], rtol=1e-6)
    assert (total, ext, shadow) == (st["totalRays"], st["totalExtensionRays"], st["totalShadowRays"])
    assert (probed_inst, probed_tri) == (st["probedInstid"], st["probedTriid"]) and probed_tri >= 0
    assert probed_dist == pytest.approx(st["probedDist"], rel=1e-6)


def test_cpp_host_through_the_vtable_equals_the_c_abi(tmp_path):
    exe = str(tmp_path / "plugin_host")
    subprocess.run(["g++", "-std=c++17", "-O1", "-o", exe, os.path.join(HERE, "cpp", "plugin_host.cpp"), "-ldl"], check=True)
    _run_host_and_compare(tmp_path, exe)


def test_host_built_from_the_reference_headers_and_loader(tmp_path):
    """The same host compiled against the REFERENCE'S OWN declarations (lib/RenderSystem/rendersystem.h) and linked with the
    reference's own loader (lib/RenderSystem/core_api_base.cpp, compiled in place by oracle/Makefile where /root/reference exists;
    the binary travels to the GPU box): CoreAPI_Base::CreateCoreAPI( "RenderCore_B200" ) finds libRenderCore_B200.so on the library
    path, and every virtual call, struct and the by-value CoreStats go through the reference's compiler view of the types."""
    exe = os.path.join(os.path.dirname(HERE), "oracle", "_ref", "plugin_host_ref")
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/plugin_host_ref not built (needs /root/reference: make -C oracle)")
    env = dict(os.environ)
    env["LD_LIBRARY_PATH"] = os.path.dirname(core.LIB_PATH) + os.pathsep + env.get("LD_LIBRARY_PATH", "")
    _run_host_and_compare(tmp_path, exe, env)
