"""Light records in the layouts CoreAPI_Base::SetLights takes (lib/RenderSystem/common_classes.h:275-345), built the way
the reference's RenderSystem builds them (lib/RenderSystem/host_light.cpp:25-110).  Host side, numpy.

    CoreLightTri          24 floats: centre, energy | N, area | radiance, - | vertex0, triIdx | vertex1, instIdx | vertex2, -
    CorePointLight         8 floats: position, energy | radiance, -
    CoreSpotLight         12 floats: position, cosInner | radiance, cosOuter | direction, -
    CoreDirectionalLight   8 floats: direction, energy | radiance, -
"""
from __future__ import annotations

import numpy as np

f32 = np.float32


def _len(v):
    return np.sqrt(np.sum(v * v, dtype=f32), dtype=f32)


def area_light(v0, v1, v2, radiance, tri_idx: int, inst_idx: int, normal=None) -> np.ndarray:
    """HostAreaLight ctor + ConvertToCoreLightTri (host_light.cpp:25-63): vertices in WORLD space, Heron area,
    energy = sum(radiance * area).  `normal` defaults to normalize((v1-v0) x (v2-v0))."""
    v0, v1, v2 = (np.asarray(v, f32) for v in (v0, v1, v2))
    rad = np.asarray(radiance, f32)
    if normal is None:
        n = np.cross(v1 - v0, v2 - v0).astype(f32)
        normal = n * (f32(1) / _len(n))
    a, b, c = _len(v1 - v0), _len(v2 - v1), _len(v0 - v2)
    s = (a + b + c) * f32(0.5)
    area = np.sqrt(s * (s - a) * (s - b) * (s - c), dtype=f32)
    rec = np.zeros(24, f32)
    rec[0:3] = f32(0.333333) * (v0 + v1 + v2)
    rec[3] = np.sum(rad * area, dtype=f32)
    rec[4:7], rec[7] = normal, area
    rec[8:11] = rad
    rec[12:15], rec[16:19], rec[20:23] = v0, v1, v2
    rec[15:16].view(np.int32)[0] = tri_idx
    rec[19:20].view(np.int32)[0] = inst_idx
    return rec


def point_light(position, radiance) -> np.ndarray:
    rec = np.zeros(8, f32)
    rec[0:3], rec[4:7] = np.asarray(position, f32), np.asarray(radiance, f32)
    rec[3] = rec[4] + rec[5] + rec[6]
    return rec


def spot_light(position, direction, cos_inner, cos_outer, radiance) -> np.ndarray:
    rec = np.zeros(12, f32)
    d = np.asarray(direction, f32)
    rec[0:3], rec[3] = np.asarray(position, f32), f32(cos_inner)
    rec[4:7], rec[7] = np.asarray(radiance, f32), f32(cos_outer)
    rec[8:11] = d * (f32(1) / _len(d))
    return rec


def directional_light(direction, radiance) -> np.ndarray:
    rec = np.zeros(8, f32)
    d = np.asarray(direction, f32)
    rec[0:3], rec[4:7] = d * (f32(1) / _len(d)), np.asarray(radiance, f32)
    rec[3] = rec[4] + rec[5] + rec[6]
    return rec


def stack(records, width: int) -> np.ndarray:
    return np.stack(records).astype(f32) if len(records) else np.zeros((0, width), f32)
