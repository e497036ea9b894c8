"""CPU oracles for the extend/shadow traversal path -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import
this package.  ``Oracle("ref")`` drives oracle/_ref/liboracle_ref.so (the reference's own unmodified
lib/RenderCore_Bart/bvh.cpp + common.h compiled in place); ``Oracle("port")`` drives
oracle/liboracle_port.so (the plain-C restatement, bart_port.c).  Both expose the same calls.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
PORT_SO = os.path.join(HERE, "liboracle_port.so")
REF_SO = os.path.join(HERE, "_ref", "liboracle_ref.so")

_f = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
_i = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")


def build() -> None:
    """Compile the checkers (port always; _ref only where /root/reference exists)."""
    subprocess.run(["make", "-s", "-C", HERE], check=True)


def have_ref() -> bool:
    return os.path.exists(REF_SO)


def _load(kind: str):
    path = REF_SO if kind == "ref" else PORT_SO
    if not os.path.exists(path):
        build()
    lib = C.CDLL(path)
    p = kind + "_"
    g = lambda n: getattr(lib, p + n)
    g("scene_create").restype = C.c_void_p
    g("scene_destroy").argtypes = [C.c_void_p]
    g("set_geometry").argtypes = [C.c_void_p, C.c_int, _f, C.c_int]
    g("set_instance").argtypes = [C.c_void_p, C.c_int, C.c_int, _f]
    g("build_toplevel").argtypes = [C.c_void_p]
    g("blas_build_seconds").argtypes = [C.c_void_p]
    g("blas_build_seconds").restype = C.c_double
    g("tlas_build_seconds").argtypes = [C.c_void_p]
    g("tlas_build_seconds").restype = C.c_double
    g("get_inv_transform").argtypes = [C.c_void_p, C.c_int, _f]
    if kind == "ref":
        g("trace").argtypes = [C.c_void_p, C.c_int, _f, _f, _f, _i, C.c_int]
    else:
        g("trace").argtypes = [C.c_void_p, C.c_int, _f, _f, _f, _i, _i, C.c_int]
        g("trace_brute").argtypes = [C.c_void_p, C.c_int, _f, _f, _f, _i, _i, C.c_int]
        g("trace_visits").argtypes = [C.c_void_p, C.c_int, _f, _f, _f, _i, _i, _i, C.c_int]
        g("trace_single").argtypes = [C.c_void_p, C.c_int, _f, _f, _i, _i, _f]
        g("normalize_n").argtypes = [C.c_int, _f]
        g("blas_node_count").argtypes = [C.c_void_p, C.c_int]
    g("intersect_triangle").argtypes = [_f, _f, _f, _f]
    g("intersect_aabb").argtypes = [_f, _f, _f, _f, _f]
    g("normalize").argtypes = [_f, _f]
    return lib


class Oracle:
    """One scene in one of the two CPU oracles.  Method names follow the CoreAPI calls they restate."""

    def __init__(self, kind: str = "port"):
        assert kind in ("ref", "port")
        self.kind = kind
        self.lib = _load(kind)
        self._p = kind + "_"
        self.h = C.c_void_p(self._fn("scene_create")())
        self.n_inst = 0

    def _fn(self, n):
        return getattr(self.lib, self._p + n)

    def close(self):
        if self.h:
            self._fn("scene_destroy")(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # RenderCore::SetGeometry (lib/RenderCore_Bart/rendercore.cpp:47)
    def SetGeometry(self, mesh_idx: int, verts: np.ndarray):
        verts = np.ascontiguousarray(verts, np.float32).reshape(-1, 9)
        rc = self._fn("set_geometry")(self.h, mesh_idx, verts, verts.shape[0])
        assert rc == 0

    # RenderCore::SetInstance (rendercore.cpp:82)
    def SetInstance(self, inst_idx: int, mesh_idx: int, mat16=None):
        m = np.eye(4, dtype=np.float32).reshape(16) if mat16 is None else np.ascontiguousarray(mat16, np.float32).reshape(16)
        rc = self._fn("set_instance")(self.h, inst_idx, mesh_idx, m)
        assert rc == 0
        self.n_inst = max(self.n_inst, inst_idx + 1)

    # TopLevelBVH::Rebuild, called per frame from RenderCore::Render (rendercore.cpp:199)
    def UpdateToplevel(self):
        self._fn("build_toplevel")(self.h)

    def inv_transform(self, inst_idx: int) -> np.ndarray:
        out = np.empty(16, np.float32)
        self._fn("get_inv_transform")(self.h, inst_idx, out)
        return out

    @property
    def blas_build_seconds(self):
        return self._fn("blas_build_seconds")(self.h)

    @property
    def tlas_build_seconds(self):
        return self._fn("tlas_build_seconds")(self.h)

    def trace(self, O, D, nthreads: int = 1, mode: str = "bvh"):
        """Closest hit per ray.  Returns (t float32[n], prim int32[n], inst int32[n]); miss = (1e34, -1, -1).
        ``inst`` is -1-filled for kind='ref' (the reference API cannot return it)."""
        O = np.ascontiguousarray(O, np.float32).reshape(-1, 3)
        D = np.ascontiguousarray(D, np.float32).reshape(-1, 3)
        n = O.shape[0]
        t = np.empty(n, np.float32)
        prim = np.empty(n, np.int32)
        inst = np.full(n, -1, np.int32)
        if self.kind == "ref":
            assert mode == "bvh"
            self._fn("trace")(self.h, n, O, D, t, prim, nthreads)
        else:
            self._fn("trace" if mode == "bvh" else "trace_brute")(self.h, n, O, D, t, prim, inst, nthreads)
        return t, prim, inst

    def trace_visits(self, O, D, nthreads: int = 1):
        assert self.kind == "port"
        O = np.ascontiguousarray(O, np.float32).reshape(-1, 3)
        D = np.ascontiguousarray(D, np.float32).reshape(-1, 3)
        n = O.shape[0]
        t = np.empty(n, np.float32); prim = np.empty(n, np.int32); inst = np.empty(n, np.int32); vis = np.empty(n, np.int32)
        self._fn("trace_visits")(self.h, n, O, D, t, prim, inst, vis, nthreads)
        return t, prim, inst, vis

    def t_of(self, O, D, inst, prim):
        """Reference-arithmetic t of given (inst, prim) pairs (1e34 where the triangle test rejects)."""
        assert self.kind == "port"
        O = np.ascontiguousarray(O, np.float32).reshape(-1, 3)
        D = np.ascontiguousarray(D, np.float32).reshape(-1, 3)
        out = np.empty(O.shape[0], np.float32)
        self._fn("trace_single")(self.h, O.shape[0], O, D, np.ascontiguousarray(inst, np.int32), np.ascontiguousarray(prim, np.int32), out)
        return out

    def occluded(self, O, D, tmax, nthreads: int = 1):
        """Any-hit as the reference defines it: closest t < tmax (raytracer.cpp:135-139,155-159)."""
        t, _, _ = self.trace(O, D, nthreads)
        return t < np.asarray(tmax, np.float32)

    def intersect_triangle(self, O, D, v9, t0=3.0e38):
        t = np.array([t0], np.float32)
        r = self._fn("intersect_triangle")(np.asarray(O, np.float32), np.asarray(D, np.float32), np.asarray(v9, np.float32).reshape(9), t)
        return bool(r), float(t[0])

    def intersect_aabb(self, O, D, bmin, bmax, t0=3.0e38):
        t = np.array([t0], np.float32)
        r = self._fn("intersect_aabb")(np.asarray(O, np.float32), np.asarray(D, np.float32), np.asarray(bmin, np.float32), np.asarray(bmax, np.float32), t)
        return bool(r), float(t[0])

    def normalize(self, v):
        out = np.empty(3, np.float32)
        self._fn("normalize")(np.asarray(v, np.float32), out)
        return out


BSDF_REF_SO = os.path.join(HERE, "_ref", "liboracle_bsdf_ref.so")


def have_bsdf_ref() -> bool:
    return os.path.exists(BSDF_REF_SO)


class DisneyRef:
    """The reference's own Disney BSDF (lib/sharedBSDFs/disney.h + ggxmdf.h, unmodified, compiled for the host by
    oracle/Makefile; see ref_bsdf_driver.cpp), over arrays.  Materials are given the way the reference's ShadingData holds
    them: color[n,3], transmittance[n,3], tint4[n,4] (tint rgb, luminance), params uint32[n,4] (three words of 0.8
    fixed-point parameters, eta as float bits)."""

    def __init__(self):
        if not os.path.exists(BSDF_REF_SO):
            build()
        if not os.path.exists(BSDF_REF_SO):
            raise RuntimeError("oracle/_ref/liboracle_bsdf_ref.so is missing and /root/reference is not here to build it")
        lib = C.CDLL(BSDF_REF_SO)
        u = np.ctypeslib.ndpointer(np.uint32, flags="C_CONTIGUOUS")
        lib.ref_disney_sample.argtypes = [C.c_int, _f, _f, _f, u, _f, _f, _f, _f, _f, _f, _f, _f, _f, _i, _f]
        lib.ref_disney_evaluate.argtypes = [C.c_int, _f, _f, _f, u, _f, _f, _f, _f, _f, _f]
        lib.ref_consistent_normal.argtypes = [C.c_int, _f, _f, _f, _f]
        self.lib = lib

    @staticmethod
    def _c(a, dt=np.float32):
        return np.ascontiguousarray(a, dt)

    def sample(self, color, transmittance, tint4, params, iN, N, iT, wow, distance, r0, r1):
        """SampleBSDF (disney.h:227-312) -> (wiw[n,3], pdf[n], specular[n] bool, value[n,3])"""
        n = len(r0)
        wiw, pdf, spec, value = np.zeros((n, 3), np.float32), np.zeros(n, np.float32), np.zeros(n, np.int32), np.zeros((n, 3), np.float32)
        c = self._c
        self.lib.ref_disney_sample(n, c(color), c(transmittance), c(tint4), c(params, np.uint32), c(iN), c(N), c(iT), c(wow), c(distance), c(r0), c(r1),
                                   wiw, pdf, spec, value)
        return wiw, pdf, spec != 0, value

    def evaluate(self, color, transmittance, tint4, params, iN, iT, wow, wiw):
        """EvaluateBSDF (disney.h:314-348) -> (value[n,3], pdf[n])"""
        n = len(wiw)
        pdf, value = np.zeros(n, np.float32), np.zeros((n, 3), np.float32)
        c = self._c
        self.lib.ref_disney_evaluate(n, c(color), c(transmittance), c(tint4), c(params, np.uint32), c(iN), c(iT), c(wow), c(wiw), pdf, value)
        return value, pdf

    def consistent_normal(self, D, iN, alpha):
        """ConsistentNormal (tools_shared.h:296-310)"""
        out = np.zeros((len(alpha), 3), np.float32)
        self.lib.ref_consistent_normal(len(alpha), self._c(D), self._c(iN), self._c(alpha), out)
        return out


SHADE_REF_SO = os.path.join(HERE, "_ref", "liboracle_shade_ref.so")


def have_shade_ref() -> bool:
    return os.path.exists(SHADE_REF_SO)


class ShadeRef:
    """The reference's own Lambert BSDF and light-sampling helpers (lib/sharedBSDFs/lambert.h, lib/CUDA/shared_kernel_code/
    lights_shared.h, unmodified, compiled for the host by oracle/Makefile; see ref_shade_driver.cpp), over arrays.  Materials the way
    the reference's ShadingData holds them: color[n,3], transmittance[n,3], params uint32[n,4] (roughness in bits 24..31 of word 0,
    transmission in bits 16..23 of word 2, eta as float bits in word 3).  Lights: the record arrays of lighthouse2_b200/lights.py."""

    def __init__(self):
        if not os.path.exists(SHADE_REF_SO):
            build()
        if not os.path.exists(SHADE_REF_SO):
            raise RuntimeError("oracle/_ref/liboracle_shade_ref.so is missing and /root/reference is not here to build it")
        lib = C.CDLL(SHADE_REF_SO)
        u = np.ctypeslib.ndpointer(np.uint32, flags="C_CONTIGUOUS")
        lib.ref_lambert_sample.argtypes = [C.c_int, _f, _f, u, _f, _f, _f, _f, _f, _f, _f, _f, _i, _f]
        lib.ref_lambert_evaluate.argtypes = [C.c_int, _f, _f, u, _f, _f, _f, _f, _f]
        lib.ref_set_lights.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_int]
        lib.ref_set_lights.restype = C.c_int
        lib.ref_random_barycentrics.argtypes = [C.c_int, _f, _f]
        lib.ref_random_point_on_light.argtypes = [C.c_int, _f, _f, _f, _f, _f, _f, _f, _f]
        lib.ref_light_pick_prob.argtypes = [C.c_int, _i, _f, _f, _f, _f]
        lib.ref_calculate_light_pdf.argtypes = [C.c_int, _f, _f, _f, _f, _f]
        lib.ref_safe_origin.argtypes = [C.c_int, _f, _f, _f, C.c_float, _f]
        self.lib = lib
        self._lights = None

    @staticmethod
    def _c(a, dt=np.float32):
        return np.ascontiguousarray(a, dt)

    @staticmethod
    def params(roughness, transmission, eta):
        """uint32[n,4] parameter words as the cores pack them (TOUINT4, rendercore_optix7/rendercore.cpp:511-512)."""
        r = (np.asarray(roughness, np.float32) * np.float32(255.0)).astype(np.uint32)
        t = (np.asarray(transmission, np.float32) * np.float32(255.0)).astype(np.uint32)
        out = np.zeros((len(r), 4), np.uint32)
        out[:, 0] = r << np.uint32(24)
        out[:, 2] = t << np.uint32(16)
        out[:, 3] = np.asarray(eta, np.float32).view(np.uint32)
        return out

    def sample(self, color, transmittance, params, iN, N, wow, distance, r3, r4):
        """SampleBSDF (lambert.h:71-123) -> (wiw[n,3], pdf[n], specular[n] bool, value[n,3])"""
        n = len(r3)
        wiw, pdf, spec, value = np.zeros((n, 3), np.float32), np.zeros(n, np.float32), np.zeros(n, np.int32), np.zeros((n, 3), np.float32)
        c = self._c
        self.lib.ref_lambert_sample(n, c(color), c(transmittance), c(params, np.uint32), c(iN), c(N), c(wow), c(distance), c(r3), c(r4), wiw, pdf, spec, value)
        return wiw, pdf, spec != 0, value

    def evaluate(self, color, transmittance, params, iN, wow, wiw):
        """EvaluateBSDF (lambert.h:58-69) -> (value[n,3], pdf[n])"""
        n = len(wiw)
        pdf, value = np.zeros(n, np.float32), np.zeros((n, 3), np.float32)
        c = self._c
        self.lib.ref_lambert_evaluate(n, c(color), c(transmittance), c(params, np.uint32), c(iN), c(wow), c(wiw), pdf, value)
        return value, pdf

    def set_lights(self, area=None, point=None, spot=None, directional=None):
        arrs = []
        for a, width in ((area, 24), (point, 8), (spot, 12), (directional, 8)):
            arrs.append(np.zeros((0, width), np.float32) if a is None else np.ascontiguousarray(a, np.float32).reshape(-1, width))
        self._lights = arrs  # (the library copies; kept for symmetry)
        args = []
        for a in arrs:
            args += [a.ctypes.data if len(a) else None, len(a)]
        if self.lib.ref_set_lights(*args):
            raise ValueError("the reference keeps light potentials in float[MAXISLIGHTS = 8] (lights_shared.h:25,124,177): more than 8 lights overrun it")

    def random_barycentrics(self, r0):
        out = np.zeros((len(r0), 3), np.float32)
        self.lib.ref_random_barycentrics(len(r0), self._c(r0), out)
        return out

    def random_point_on_light(self, r0, r1, I, N):
        """RandomPointOnLight (lights_shared.h:171-262) -> P, pickProb, lightPdf, lightColor (NaN where the reference leaves it unset)"""
        n = len(r0)
        P, pick, pdf, col = np.zeros((n, 3), np.float32), np.zeros(n, np.float32), np.zeros(n, np.float32), np.zeros((n, 3), np.float32)
        c = self._c
        self.lib.ref_random_point_on_light(n, c(r0), c(r1), c(I), c(N), P, pick, pdf, col)
        return P, pick, pdf, col

    def light_pick_prob(self, idx, O, N, I):
        out = np.zeros(len(idx), np.float32)
        self.lib.ref_light_pick_prob(len(idx), self._c(idx, np.int32), self._c(O), self._c(N), self._c(I), out)
        return out

    def calculate_light_pdf(self, D, t, area, light_normal):
        out = np.zeros(len(t), np.float32)
        self.lib.ref_calculate_light_pdf(len(t), self._c(D), self._c(t), self._c(area), self._c(light_normal), out)
        return out

    def safe_origin(self, O, R, N, eps):
        out = np.zeros((len(O), 3), np.float32)
        self.lib.ref_safe_origin(len(O), self._c(O), self._c(R), self._c(N), float(eps), out)
        return out



FILTER_REF_SO = os.path.join(HERE, "_ref", "liboracle_filter_ref.so")


def have_filter_ref() -> bool:
    return os.path.exists(FILTER_REF_SO)


class FilterRef:
    """The reference's own SVGF-style filter and TAA kernels (lib/CUDA/shared_kernel_code/finalize_shared.h:102-483), compiled in
    place by nvcc (oracle/ref_filter_driver.cu).  CUDA code: usable on a GPU box only.  Every argument that names a buffer is a DEVICE
    pointer (int); the calls synchronise."""

    def __init__(self):
        if not have_filter_ref():
            if os.path.isdir("/root/reference"):
                build()
            if not have_filter_ref():
                raise RuntimeError("oracle/_ref/liboracle_filter_ref.so is missing and /root/reference is not here to build it")
        lib = C.CDLL(FILTER_REF_SO)
        vp, ip, fl, u = C.c_void_p, C.c_int, C.c_float, C.c_uint
        lib.ref_prepare_filter.argtypes = [vp] * 9 + [_f] + [fl] * 4 + [ip, ip, u, fl, fl, ip]
        lib.ref_apply_filter.argtypes = [vp] * 9 + [u, u, ip, u]
        lib.ref_taa_pass.argtypes = [vp] * 5 + [u, u]
        lib.ref_unsharpen_taa.argtypes = [vp, vp, u, u]
        lib.ref_finalize_no_taa.argtypes = [vp, vp, u, u]
        for fn in (lib.ref_prepare_filter, lib.ref_apply_filter, lib.ref_taa_pass, lib.ref_unsharpen_taa, lib.ref_finalize_no_taa):
            fn.restype = ip
        self.lib = lib

    @staticmethod
    def _ok(code, what):
        if code:
            raise RuntimeError("reference filter: %s failed" % what)

    def prepare(self, accumulator, features, world_pos, prev_world_pos, shading, motion, moments, prev_moments, delta_depth, prev_view, j, prev_j, w, h, spp,
                clamp_direct, clamp_indirect, stationary):
        view = np.ascontiguousarray(prev_view, np.float32).reshape(12)
        self._ok(self.lib.ref_prepare_filter(accumulator, features, world_pos, prev_world_pos, shading, motion, moments, prev_moments, delta_depth, view,
                                             j[0], j[1], prev_j[0], prev_j[1], w, h, spp, clamp_direct, clamp_indirect, int(stationary)), "prepareFilter")

    def apply(self, features, prev_world_pos, world_pos, delta_depth, motion, moments, A, B, Cout, w, h, phase, last):
        self._ok(self.lib.ref_apply_filter(features, prev_world_pos, world_pos, delta_depth, motion, moments, A, B, Cout, w, h, phase, int(last)), "applyFilter")

    def taa(self, pixels, prev_pixels, world_pos, prev_world_pos, motion, w, h):
        self._ok(self.lib.ref_taa_pass(pixels, prev_pixels, world_pos, prev_world_pos, motion, w, h), "TAApass")

    def unsharpen(self, pixels, out, w, h):
        self._ok(self.lib.ref_unsharpen_taa(pixels, out, w, h), "unsharpenTAA")

    def finalize_no_taa(self, pixels, out, w, h):
        self._ok(self.lib.ref_finalize_no_taa(pixels, out, w, h), "finalizeNoTAA")
