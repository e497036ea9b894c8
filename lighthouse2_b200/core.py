"""ctypes driver over the C ABI of libRenderCore_B200.so (include/rendercore_b200.h).

This is the test/bench-side host layer: method names follow CoreAPI_Base
(lib/RenderSystem/core_api_base.h:84-113) so the parity tests read like calls into a reference core.
There is no CPU fallback anywhere: if the shared library is missing or no sm_100 GPU is visible the
constructor raises.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("B200_LIB") or os.path.join(HERE, "lib", "libRenderCore_B200.so")  # B200_LIB: tuning variants (tools/build_variant.sh)
CSRC = os.path.join(HERE, "csrc")

# every symbol include/rendercore_b200.h declares (checked by tests/test_abi.py without a GPU)
C_ABI_SYMBOLS = [
    "b200_create", "b200_destroy", "b200_last_error", "b200_set_stream", "b200_set_geometry", "b200_set_instance",
    "b200_update_toplevel", "b200_refit", "b200_refit_device", "b200_setting", "b200_trace_closest", "b200_trace_any",
    "b200_trace_closest_device", "b200_trace_any_device", "b200_read_step_counters", "b200_read_tree", "b200_set_target",
    "b200_set_point_light", "b200_set_sky", "b200_set_lights", "b200_set_sky_data", "b200_render", "b200_read_accumulator", "b200_read_frame", "b200_read_frame_async", "b200_wait_frame",
    "b200_device_accumulator", "b200_get_stats", "b200_set_probe", "b200_plugin_core", "b200_host_alloc",
    "b200_host_free", "CreateCore", "DestroyCore", "b200_set_materials", "b200_set_material_colors", "b200_set_material_table", "b200_set_textures",
    "b200_disney_probe", "b200_set_blue_noise",
    "b200_set_accumulator_buffer", "b200_finalize", "b200_trace_connect_device",
    "b200_trace_closest_async", "b200_trace_any_async", "b200_wait",
    "b200_trace_closest_noalloc",
    "b200_peer_handle", "b200_open_peers", "b200_close_peers", "b200_gather_finalize",
    "b200_read_filter_buffer", "b200_filter_prepare", "b200_filter_reconstruct", "b200_filter_taa", "b200_filter_unsharpen", "b200_filter_finish_no_taa",
]


class Stats(C.Structure):
    _fields_ = [
        ("smCount", C.c_uint32), ("ccMajor", C.c_uint32), ("ccMinor", C.c_uint32), ("vramMB", C.c_uint32),
        ("bvhBuildTime", C.c_float), ("refitTime", C.c_float), ("tlasTime", C.c_float),
        ("totalRays", C.c_uint32), ("totalExtensionRays", C.c_uint32), ("totalShadowRays", C.c_uint32),
        ("renderTime", C.c_float),
        ("primaryRayCount", C.c_uint32), ("traceTime0", C.c_float),
        ("bounce1RayCount", C.c_uint32), ("traceTime1", C.c_float),
        ("deepRayCount", C.c_uint32), ("traceTimeX", C.c_float),
        ("shadowTraceTime", C.c_float), ("shadeTime", C.c_float),
        ("probedInstid", C.c_int32), ("probedTriid", C.c_int32), ("probedDist", C.c_float),
        ("wideNodeCount", C.c_uint32), ("triangleCount", C.c_uint32), ("instanceCount", C.c_uint32), ("tlasNodeCount", C.c_uint32),
        ("kernelLaunches", C.c_uint32), ("lastTraceMs", C.c_float), ("bvhCacheHits", C.c_uint32),
        ("gatherKernelMs", C.c_float), ("frameGraphLaunches", C.c_uint32), ("frameChains", C.c_uint32), ("framePasses", C.c_uint32),
        ("hostEnqueueMs", C.c_float), ("hostWaitMs", C.c_float), ("hostTotalMs", C.c_float), ("gatherMs", C.c_float),
        ("argb32TexelCount", C.c_uint32), ("argb128TexelCount", C.c_uint32), ("nrm32TexelCount", C.c_uint32), ("filterTime", C.c_float),
    ]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


class ViewPyramid(C.Structure):
    """lib/RenderSystem/common_classes.h:362-385 (68 bytes)."""
    _fields_ = [("pos", C.c_float * 3), ("p1", C.c_float * 3), ("p2", C.c_float * 3), ("p3", C.c_float * 3),
                ("aperture", C.c_float), ("spreadAngle", C.c_float), ("imagePlane", C.c_float),
                ("focalDistance", C.c_float), ("distortion", C.c_float)]


def build_library() -> str:
    """nvcc build of the shared library for sm_100a (works without a GPU)."""
    subprocess.run(["make", "-s", "-j8", "-C", CSRC], check=True)
    return LIB_PATH


_lib = None


def load_library() -> C.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(RenderCore_B200 has no CPU fallback)")
    lib = C.CDLL(LIB_PATH)
    vp, ip, fp, up = C.c_void_p, C.c_int, C.POINTER(C.c_float), C.POINTER(C.c_uint32)
    sig = {
        "b200_create": ([C.POINTER(vp), ip], ip), "b200_destroy": ([vp], None), "b200_last_error": ([], C.c_char_p),
        "b200_set_stream": ([vp, vp], ip),
        "b200_set_geometry": ([vp, ip, vp, ip, ip, vp, vp], ip), "b200_set_instance": ([vp, ip, ip, vp], ip),
        "b200_update_toplevel": ([vp], ip), "b200_refit": ([vp, ip, vp, ip], ip), "b200_refit_device": ([vp, ip, vp, ip], ip),
        "b200_setting": ([vp, C.c_char_p, C.c_float], ip),
        "b200_trace_closest": ([vp, vp, vp, ip, vp], ip), "b200_trace_any": ([vp, vp, vp, ip, vp], ip),
        "b200_trace_closest_device": ([vp, vp, vp, ip, vp], ip), "b200_trace_any_device": ([vp, vp, vp, ip, vp], ip),
        "b200_read_step_counters": ([vp, ip, vp, vp], ip), "b200_read_tree": ([vp, ip, vp, C.c_size_t], ip),
        "b200_set_target": ([vp, ip, ip, ip], ip), "b200_set_point_light": ([vp, vp, vp], ip), "b200_set_sky": ([vp, vp], ip),
        "b200_set_lights": ([vp, vp, ip, vp, ip, vp, ip, vp, ip], ip), "b200_set_sky_data": ([vp, vp, ip, ip], ip),
        "b200_render": ([vp, vp, ip], ip), "b200_read_accumulator": ([vp, vp], ip), "b200_read_frame": ([vp, vp], ip), "b200_read_frame_async": ([vp, vp], ip), "b200_wait_frame": ([vp], ip),
        "b200_device_accumulator": ([vp], vp), "b200_get_stats": ([vp, C.POINTER(Stats)], ip), "b200_set_probe": ([vp, ip, ip], ip),
        "b200_plugin_core": ([], vp), "b200_host_alloc": ([C.c_uint64], vp), "b200_host_free": ([vp], None),
        "b200_set_materials": ([vp, vp, ip], ip), "b200_set_material_colors": ([vp, vp, ip], ip), "b200_set_material_table": ([vp, vp, ip], ip), "b200_set_textures": ([vp, vp, ip], ip),
        "b200_disney_probe": ([vp, ip, ip, vp, vp], ip), "b200_set_blue_noise": ([vp, vp, ip], ip),
        "b200_set_accumulator_buffer": ([vp, vp], ip), "b200_finalize": ([vp, C.c_uint32], ip),
        "b200_trace_connect_device": ([vp, vp, vp, ip, vp, vp], ip),
        "b200_trace_closest_async": ([vp, vp, vp, ip, vp], ip), "b200_trace_any_async": ([vp, vp, vp, ip, vp], ip), "b200_wait": ([vp], ip),
        "b200_trace_closest_noalloc": ([vp, vp, vp, ip, vp], ip),
        "b200_peer_handle": ([vp, vp], ip), "b200_open_peers": ([vp, vp, ip, ip], ip), "b200_close_peers": ([vp], ip),
        "b200_gather_finalize": ([vp, C.c_uint32], ip),
        "b200_read_filter_buffer": ([vp, ip, vp], ip),
        "b200_filter_prepare": ([vp] * 9 + [vp] + [C.c_float] * 4 + [ip, ip, C.c_uint32, C.c_float, C.c_float, ip], ip),
        "b200_filter_reconstruct": ([vp] * 9 + [ip, ip, ip, ip], ip),
        "b200_filter_taa": ([vp, vp, vp, vp, ip, ip], ip), "b200_filter_unsharpen": ([vp, vp, ip, ip], ip), "b200_filter_finish_no_taa": ([vp, vp, ip, ip], ip),
    }
    for name, (args, res) in sig.items():
        fn = getattr(lib, name)
        fn.argtypes, fn.restype = args, res
    _lib = lib
    return lib


class B200Error(RuntimeError):
    pass


def pack_rays(O, D, tmin=0.0, tmax=1e34):
    """(O[n,3], D[n,3]) -> the two float4 planes of the wire format: O4=(O,tmin), D4=(D,tmax)."""
    O = np.asarray(O, np.float32).reshape(-1, 3)
    D = np.asarray(D, np.float32).reshape(-1, 3)
    n = O.shape[0]
    O4 = np.empty((n, 4), np.float32)
    D4 = np.empty((n, 4), np.float32)
    O4[:, :3], D4[:, :3] = O, D
    O4[:, 3], D4[:, 3] = tmin, tmax
    return O4, D4


def unpack_hits(hits: np.ndarray):
    """float4 hit records -> dict(t, prim, inst, u, v) following the optix7 decode macros
    (lib/rendercore_optix7/kernels/pathtracer.h:36-40)."""
    h = np.ascontiguousarray(hits, np.float32).reshape(-1, 4)
    bits = h.view(np.uint32)
    return {
        "t": h[:, 3].copy(), "prim": bits[:, 2].view(np.int32).copy(), "inst": bits[:, 1].view(np.int32).copy(),
        "u": (bits[:, 0] & 0xFFFF).astype(np.float32) * np.float32(1.0 / 65535.0),
        "v": (bits[:, 0] >> 16).astype(np.float32) * np.float32(1.0 / 65535.0),
    }


def core_tris(verts: np.ndarray, materials=None, vertex_normals=None, uvs=None, lod=None, tangents=None, alphas=None) -> np.ndarray:
    """CoreTri[T] (176 bytes each, lib/RenderSystem/common_classes.h:60-82) for verts[T,3,3]: material (offset 28),
    geometric normal Nx/Ny/Nz (44/60/76), area (92) and vertex0/1/2 (128/144/160) filled, the rest zero; ltriIdx = -1.
    vertex_normals[T,3,3] fills vN0/vN1/vN2 (offsets 32/48/64) for SMOOTH materials; uvs[T,3,2] the texture coordinates
    u0..2 / v0..2 (offsets 0 / 16), lod[T] CoreTri::LOD (124), tangents[T,3] CoreTri::T (80; the Disney BSDF's anisotropy
    axis), alphas[T,3] the per-vertex angles of the consistent-normal interpolation (CoreTri::alpha, 112)."""
    v = np.asarray(verts, np.float32)
    t = v.shape[0]
    out = np.zeros((t, 44), np.float32)
    n = np.cross(v[:, 1] - v[:, 0], v[:, 2] - v[:, 0]).astype(np.float32)
    ln = np.sqrt(np.einsum("ij,ij->i", n, n)).astype(np.float32)
    nn = n / np.maximum(ln, np.float32(1e-30))[:, None]
    out[:, 11], out[:, 15], out[:, 19] = nn[:, 0], nn[:, 1], nn[:, 2]
    out[:, 23] = ln * np.float32(0.5)
    out[:, 32:35], out[:, 36:39], out[:, 40:43] = v[:, 0], v[:, 1], v[:, 2]
    if vertex_normals is not None:
        vn = np.asarray(vertex_normals, np.float32)
        out[:, 8:11], out[:, 12:15], out[:, 16:19] = vn[:, 0], vn[:, 1], vn[:, 2]
    if uvs is not None:
        uv = np.asarray(uvs, np.float32)
        out[:, 0:3], out[:, 4:7] = uv[:, :, 0], uv[:, :, 1]
    if lod is not None:
        out[:, 31] = np.asarray(lod, np.float32)
    if tangents is not None:
        out[:, 20:23] = np.asarray(tangents, np.float32)
    if alphas is not None:
        out[:, 28:31] = np.asarray(alphas, np.float32)
    iv = out.view(np.int32)
    iv[:, 3] = -1
    iv[:, 7] = 0 if materials is None else np.asarray(materials, np.int32)
    return out


# CoreMaterial (688 bytes, lib/RenderSystem/common_classes.h:177-238): byte offset of each value; Vec3Value = value float3,
# textureID, scale, uvscale, uvoffset (40 B); ScalarValue = value, textureID, component, scale, uvscale, uvoffset (32 B)
CORE_MATERIAL_SCALARS = dict(metallic=208, subsurface=240, specular=272, roughness=304, specularTint=336, anisotropic=368, sheen=400,
                             sheenTint=432, clearcoat=464, clearcoatGloss=496, transmission=528, eta=560)
CORE_MATERIAL_DEFAULTS = dict(metallic=0.0, subsurface=0.0, specular=0.5, roughness=0.5, specularTint=0.0, anisotropic=0.0, sheen=0.0,
                              sheenTint=0.0, clearcoat=0.0, clearcoatGloss=1.0, transmission=0.0, eta=1.0)  # host_material.h


# byte offsets of (textureID, scale, uvscale, uvoffset) of detailColor @40, normals @80, detailNormals @120 (Vec3Value: textureID +12,
# scale +16, uvscale +24, uvoffset +32 -- float2 is 8-byte aligned in the reference, common_types.h:62) and roughness @304 (ScalarValue:
# textureID +4, scale +12, uvscale +16, uvoffset +24); tests/cpp/abi_vs_reference.cpp checks them against the reference header
CORE_MATERIAL_MAPS = dict(detail=(52, 56, 64, 72), normals=(92, 96, 104, 112), detail_normals=(132, 136, 144, 152), roughness_map=(308, 316, 320, 328))


def core_materials(materials) -> np.ndarray:
    """CoreMaterial[n] as bytes [n, 688] from a list of dicts: color (rgb), absorption (rgb, default 0), flags (1 SMOOTH,
    2 HASALPHA), texture (diffuse map index, default -1), uvscale / uvoffset, and the Disney scalars of
    CORE_MATERIAL_SCALARS (HostMaterial defaults where missing).  Every textureID that is not given is -1."""
    out = np.zeros((len(materials), 688), np.uint8)
    f, i = out.view(np.float32), out.view(np.int32)
    for k, m in enumerate(materials):
        for off in (0, 40, 80, 120, 168):  # color, detailColor, normals, detailNormals, absorption
            i[k, off // 4 + 3] = -1
            f[k, off // 4 + 4] = 1.0
            f[k, off // 4 + 5:off // 4 + 7] = 1.0
        for off in list(CORE_MATERIAL_SCALARS.values()) + [592, 624, 656]:
            i[k, off // 4 + 1] = -1
            f[k, off // 4 + 3] = 1.0
            f[k, off // 4 + 4:off // 4 + 6] = 1.0
        f[k, 0:3] = m.get("color", (0.8, 0.8, 0.8))
        i[k, 3] = m.get("texture", -1)
        f[k, 6:8] = m.get("uvscale", (1.0, 1.0))
        f[k, 8:10] = m.get("uvoffset", (0.0, 0.0))
        i[k, 40] = m.get("flags", 0)
        f[k, 42:45] = m.get("absorption", (0.0, 0.0, 0.0))
        for name, off in CORE_MATERIAL_SCALARS.items():
            f[k, off // 4] = m.get(name, CORE_MATERIAL_DEFAULTS[name])
        # the other maps: {"detail" | "normals" | "detail_normals" | "roughness_map": dict(texture=, scale=, uvscale=, uvoffset=)}
        for name, (o_tex, o_scale, o_uvs, o_uvo) in CORE_MATERIAL_MAPS.items():
            mm = m.get(name)
            if mm is not None:
                i[k, o_tex // 4] = mm.get("texture", -1)
                f[k, o_scale // 4] = mm.get("scale", 1.0)
                f[k, o_uvs // 4:o_uvs // 4 + 2] = mm.get("uvscale", (1.0, 1.0))
                f[k, o_uvo // 4:o_uvo // 4 + 2] = mm.get("uvoffset", (0.0, 0.0))
    return out


def disney_words(materials) -> np.ndarray:
    """The three 0.8 fixed-point parameter words per material as the cores pack them (TOUINT4,
    lib/rendercore_optix7/rendercore.cpp:511-512, 532-534) plus eta as float bits: uint32[n, 4]."""
    order = ["metallic", "subsurface", "specular", "roughness", "specularTint", "anisotropic", "sheen", "sheenTint", "clearcoat",
             "clearcoatGloss", "transmission"]
    out = np.zeros((len(materials), 4), np.uint32)
    for k, m in enumerate(materials):
        for j, name in enumerate(order):
            v = np.float32(m.get(name, CORE_MATERIAL_DEFAULTS[name]))
            out[k, j // 4] += np.uint32(int(np.float32(v * np.float32(255.0)))) << np.uint32(8 * (j & 3))
        out[k, 3] = np.float32(m.get("eta", CORE_MATERIAL_DEFAULTS["eta"])).view(np.uint32)
    return out


class RenderCoreB200:
    """One core instance on one GPU."""

    def __init__(self, device: int = 0):
        self.lib = load_library()
        h = C.c_void_p()
        if self.lib.b200_create(C.byref(h), device):
            raise B200Error(self.lib.b200_last_error().decode())
        self.h = h
        self._keep = []

    def _chk(self, rc):
        if rc:
            raise B200Error(self.lib.b200_last_error().decode())

    def Shutdown(self):
        if getattr(self, "h", None):
            self.lib.b200_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.Shutdown()
        except Exception:
            pass

    # -- scene ------------------------------------------------------------------------------------
    def SetGeometry(self, mesh_idx: int, verts: np.ndarray, core_tris: np.ndarray | None = None):
        """verts[T,3,3] (or float4[3T]) -> float4[3T] vertexData as RenderSystem sends it (w = 1)."""
        v = np.asarray(verts, np.float32)
        if v.ndim == 3 and v.shape[1:] == (3, 3):
            v4 = np.ones((v.shape[0] * 3, 4), np.float32)
            v4[:, :3] = v.reshape(-1, 3)
        else:
            v4 = np.ascontiguousarray(v.reshape(-1, 4))
        t = v4.shape[0] // 3
        tri_ptr = core_tris.ctypes.data if core_tris is not None else None
        self._chk(self.lib.b200_set_geometry(self.h, mesh_idx, v4.ctypes.data, v4.shape[0], t, tri_ptr, None))

    def SetInstance(self, inst_idx: int, mesh_idx: int, transform=None):
        m = None
        if transform is not None:
            m = np.ascontiguousarray(transform, np.float32).reshape(16)
        self._chk(self.lib.b200_set_instance(self.h, inst_idx, mesh_idx, m.ctypes.data if m is not None else None))

    def UpdateToplevel(self):
        self._chk(self.lib.b200_update_toplevel(self.h))

    def Refit(self, mesh_idx: int, verts: np.ndarray):
        v = np.asarray(verts, np.float32)
        v4 = np.ones((v.shape[0] * 3, 4), np.float32)
        v4[:, :3] = v.reshape(-1, 3)
        self._chk(self.lib.b200_refit(self.h, mesh_idx, v4.ctypes.data, v4.shape[0]))

    def RefitDevice(self, mesh_idx: int, dptr: int, vertex_count: int):
        self._chk(self.lib.b200_refit_device(self.h, mesh_idx, dptr, vertex_count))

    def Setting(self, name: str, value: float):
        self._chk(self.lib.b200_setting(self.h, name.encode(), float(value)))

    def SetStream(self, stream_handle: int):
        self._chk(self.lib.b200_set_stream(self.h, stream_handle))

    # -- raw waves ----------------------------------------------------------------------------------
    def TraceClosest(self, O4: np.ndarray, D4: np.ndarray) -> np.ndarray:
        O4 = np.ascontiguousarray(O4, np.float32).reshape(-1, 4)
        D4 = np.ascontiguousarray(D4, np.float32).reshape(-1, 4)
        hits = np.empty((O4.shape[0], 4), np.float32)
        self._chk(self.lib.b200_trace_closest(self.h, O4.ctypes.data, D4.ctypes.data, O4.shape[0], hits.ctypes.data))
        return hits

    def TraceAny(self, O4: np.ndarray, D4: np.ndarray) -> np.ndarray:
        """Returns bool[n]: True = occluded."""
        O4 = np.ascontiguousarray(O4, np.float32).reshape(-1, 4)
        D4 = np.ascontiguousarray(D4, np.float32).reshape(-1, 4)
        n = O4.shape[0]
        bits = np.zeros((n + 31) // 32, np.uint32)
        self._chk(self.lib.b200_trace_any(self.h, O4.ctypes.data, D4.ctypes.data, n, bits.ctypes.data))
        return np.unpackbits(bits.view(np.uint8), bitorder="little")[:n].astype(bool)

    def TraceQueued(self, waves):
        """Queue several host-buffer waves back to back (b200_trace_*_async) and wait for all of them (b200_wait).
        waves: [(kind, O4, D4)] with kind "closest" | "any"; returns the result arrays in order.  Buffers are staged in pinned memory."""
        out, keep = [], []
        for kind, O4, D4 in waves:
            O4 = np.ascontiguousarray(O4, np.float32).reshape(-1, 4)
            D4 = np.ascontiguousarray(D4, np.float32).reshape(-1, 4)
            n = O4.shape[0]
            pin = []
            for a in (O4, D4):
                p = self.lib.b200_host_alloc(max(a.nbytes, 16))
                if not p:
                    raise RuntimeError("b200_host_alloc failed")
                C.memmove(p, a.ctypes.data, a.nbytes)
                pin.append(p)
            nbytes = n * 16 if kind == "closest" else ((n + 31) // 32) * 4
            pr = self.lib.b200_host_alloc(max(nbytes, 16))
            fn = self.lib.b200_trace_closest_async if kind == "closest" else self.lib.b200_trace_any_async
            keep.append((kind, n, nbytes, pin, pr))
            self._chk(fn(self.h, pin[0], pin[1], n, pr))
        self._chk(self.lib.b200_wait(self.h))
        for kind, n, nbytes, pin, pr in keep:
            raw = np.frombuffer(C.string_at(pr, nbytes), np.uint8).copy()
            if kind == "closest":
                out.append(raw.view(np.float32).reshape(n, 4))
            else:
                out.append(np.unpackbits(raw, bitorder="little")[:n].astype(bool))
            for p in pin + [pr]:
                self.lib.b200_host_free(p)
        return out

    def TraceClosestDevice(self, dO4: int, dD4: int, n: int, dHits: int):
        self._chk(self.lib.b200_trace_closest_device(self.h, dO4, dD4, n, dHits))

    def TraceAnyDevice(self, dO4: int, dD4: int, n: int, dBits: int):
        self._chk(self.lib.b200_trace_any_device(self.h, dO4, dD4, n, dBits))

    def ReadTree(self):
        """(wide nodes uint32[n, 20], triangle records uint32[t, 12]) of the mesh trees as they lie in HBM."""
        st = self.GetCoreStats()
        nodes = np.empty((st["wideNodeCount"], 20), np.uint32)
        tris = np.empty((st["triangleCount"], 12), np.uint32)
        self._chk(self.lib.b200_read_tree(self.h, 0, nodes.ctypes.data, nodes.nbytes))
        self._chk(self.lib.b200_read_tree(self.h, 1, tris.ctypes.data, tris.nbytes))
        return nodes, tris

    def StepCounters(self, n: int):
        a = np.empty(n, np.uint32)
        b = np.empty(n, np.uint32)
        self._chk(self.lib.b200_read_step_counters(self.h, n, a.ctypes.data, b.ctypes.data))
        return a, b

    # -- frame --------------------------------------------------------------------------------------
    def SetTarget(self, width: int, height: int, spp: int = 1):
        self.width, self.height = width, height
        self._chk(self.lib.b200_set_target(self.h, width, height, spp))

    def SetPointLight(self, pos, radiance):
        p = np.asarray(pos, np.float32)
        r = np.asarray(radiance, np.float32)
        self._chk(self.lib.b200_set_point_light(self.h, p.ctypes.data, r.ctypes.data))

    def SetSky(self, radiance):
        r = np.asarray(radiance, np.float32)
        self._chk(self.lib.b200_set_sky(self.h, r.ctypes.data))

    def SetLights(self, area=None, point=None, spot=None, directional=None):
        """CoreAPI_Base::SetLights.  Each argument is a float32 array in the reference's record layout
        (common_classes.h:275-345): area [n, 24], point [n, 8], spot [n, 12], directional [n, 8] (see lights.py)."""
        arrs = []
        for a, width in ((area, 24), (point, 8), (spot, 12), (directional, 8)):
            a = np.zeros((0, width), np.float32) if a is None else np.ascontiguousarray(a, np.float32).reshape(-1, width)
            arrs.append(a)
        args = []
        for a in arrs:
            args += [a.ctypes.data if len(a) else None, len(a)]
        self._chk(self.lib.b200_set_lights(self.h, *args))

    def SetSkyData(self, rgb):
        """CoreAPI_Base::SetSkyData: equirectangular float32 image [height, width, 3]."""
        px = np.ascontiguousarray(rgb, np.float32)
        self._chk(self.lib.b200_set_sky_data(self.h, px.ctypes.data, px.shape[1], px.shape[0]))

    def SetMaterialColors(self, rgba):
        c = np.ascontiguousarray(rgba, np.float32).reshape(-1, 4)
        self._chk(self.lib.b200_set_material_colors(self.h, c.ctypes.data, c.shape[0]))

    def SetMaterials(self, materials):
        """CoreAPI_Base::SetMaterials: CoreMaterial bytes [n, 688] (core_materials) or a list of dicts for core_materials."""
        c = materials if isinstance(materials, np.ndarray) else core_materials(materials)
        c = np.ascontiguousarray(c, np.uint8).reshape(-1, 688)
        self._chk(self.lib.b200_set_materials(self.h, c.ctypes.data, c.shape[0]))

    def SetBlueNoise(self, words):
        """b200_set_blue_noise: uint32[5 * 65536] in the reference's layout, or None for xor32 numbers."""
        if words is None:
            self._chk(self.lib.b200_set_blue_noise(self.h, None, 0))
            return
        t = np.ascontiguousarray(words, np.uint32).reshape(-1)
        self._chk(self.lib.b200_set_blue_noise(self.h, t.ctypes.data, t.shape[0]))

    def DisneyProbe(self, mode: int, items: np.ndarray) -> np.ndarray:
        """b200_disney_probe: float32[n, 28] -> float32[n, 8] (see include/rendercore_b200.h)."""
        a = np.ascontiguousarray(items, np.float32).reshape(-1, 28)
        out = np.zeros((a.shape[0], 8), np.float32)
        self._chk(self.lib.b200_disney_probe(self.h, mode, a.shape[0], a.ctypes.data, out.ctypes.data))
        return out

    def SetMaterialTable(self, props):
        """float32[n, 16]: albedo rgb, roughness | transmission, eta, flags (1 SMOOTH + 2 HASALPHA), 0 | extinction rgb,
        diffuse texture index or -1 | uv scale xy, uv offset xy."""
        c = np.ascontiguousarray(props, np.float32).reshape(-1, 16)
        self._chk(self.lib.b200_set_material_table(self.h, c.ctypes.data, c.shape[0]))

    def SetTextures(self, textures):
        """CoreAPI_Base::SetTextures.  textures: list of (width, height, texels uint32[pixelCount] ARGB32 with the MIP levels
        behind each other, mip_levels[, storage]); storage 0 = ARGB32 (default), 2 = NRM32 (normal maps), 1 = ARGB128 (HDR: texels are
        float32[pixelCount, 4]).  Builds the CoreTexDesc array (40 bytes each, common_classes.h:246-269)."""
        descs = np.zeros((len(textures), 10), np.uint32)
        keep = []
        for i, tex in enumerate(textures):
            w, h, texels, mips = tex[:4]
            storage = tex[4] if len(tex) > 4 else 0
            t = np.ascontiguousarray(texels, np.float32).reshape(-1, 4) if storage == 1 else np.ascontiguousarray(texels, np.uint32)
            keep.append(t)
            descs[i, 0:2] = np.array([t.ctypes.data], np.uint64).view(np.uint32)
            descs[i, 2], descs[i, 3], descs[i, 5], descs[i, 7], descs[i, 8] = w, h, len(t), mips, storage
        self._chk(self.lib.b200_set_textures(self.h, descs.ctypes.data if len(textures) else None, len(textures)))

    def SetAccumulatorBuffer(self, dptr: int):
        self._chk(self.lib.b200_set_accumulator_buffer(self.h, dptr))

    def Finalize(self, total_samples: int):
        self._chk(self.lib.b200_finalize(self.h, total_samples))

    # -- multi-GPU gather over peer memory (b200_peer_handle / b200_open_peers / b200_gather_finalize) --------
    def PeerHandle(self) -> bytes:
        buf = C.create_string_buffer(128)
        self._chk(self.lib.b200_peer_handle(self.h, buf))
        return buf.raw

    def OpenPeers(self, table: bytes, world: int, rank: int):
        assert len(table) == 128 * world
        self._chk(self.lib.b200_open_peers(self.h, table, world, rank))

    def ClosePeers(self):
        self._chk(self.lib.b200_close_peers(self.h))

    def GatherFinalize(self, total_samples: int):
        self._chk(self.lib.b200_gather_finalize(self.h, total_samples))

    def TraceConnectDevice(self, dO4: int, dD4: int, n: int, dE4: int, dAcc: int):
        self._chk(self.lib.b200_trace_connect_device(self.h, dO4, dD4, n, dE4, dAcc))

    def Render(self, view: dict, converge: int = 1):
        vp = ViewPyramid()
        for k in ("pos", "p1", "p2", "p3"):
            getattr(vp, k)[:] = [float(x) for x in view[k]]
        vp.aperture, vp.spreadAngle, vp.imagePlane, vp.focalDistance, vp.distortion = (
            float(view.get("aperture", 0.0)), 0.01, 0.01, 1.0, float(view.get("distortion", 0.0)))
        self._chk(self.lib.b200_render(self.h, C.byref(vp), converge))

    def ReadAccumulator(self) -> np.ndarray:
        out = np.empty((self.height, self.width, 4), np.float32)
        self._chk(self.lib.b200_read_accumulator(self.h, out.ctypes.data))
        return out

    def ReadFilterBuffer(self, which: int) -> np.ndarray:
        """A per-pixel buffer of the last filtered frame (Setting filterCore): see b200_read_filter_buffer."""
        out = np.empty((self.height, self.width, 2 if which == 4 else 4), np.uint32 if which == 0 else np.float32)
        self._chk(self.lib.b200_read_filter_buffer(self.h, which, out.ctypes.data))
        return out

    def ReadFrame(self) -> np.ndarray:
        out = np.empty((self.height, self.width, 4), np.float32)
        self._chk(self.lib.b200_read_frame(self.h, out.ctypes.data))
        return out

    def ReadFrameAsync(self, dst_ptr: int):
        """The frame leaves for `dst_ptr` (page-locked host memory, width * height float4) beside the next Render(); WaitFrame() before it is read."""
        self._chk(self.lib.b200_read_frame_async(self.h, dst_ptr))

    def WaitFrame(self):
        self._chk(self.lib.b200_wait_frame(self.h))

    def DeviceAccumulator(self) -> int:
        return self.lib.b200_device_accumulator(self.h) or 0

    def GetCoreStats(self) -> dict:
        s = Stats()
        self._chk(self.lib.b200_get_stats(self.h, C.byref(s)))
        return s.as_dict()
