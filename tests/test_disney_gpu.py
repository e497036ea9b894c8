"""Disney BSDF of the shade stage (Setting bsdf = 1; lighthouse2_b200/csrc/disney.cuh) against the reference's own
lib/sharedBSDFs/disney.h + ggxmdf.h, compiled unmodified for the host (oracle/ref_bsdf_driver.cpp ->
oracle/_ref/liboracle_bsdf_ref.so): SampleBSDF and EvaluateBSDF on the same random inputs, lobe by lobe and mixed."""
import numpy as np
import pytest

import oracle
import wavefront_ref as wr
from lighthouse2_b200 import core

pytestmark = pytest.mark.gpu
f32 = np.float32

MATERIALS = [
    dict(color=(0.8, 0.6, 0.3), roughness=0.6),                                            # diffuse + specular
    dict(color=(0.9, 0.9, 0.9), roughness=0.15, metallic=1.0),                             # metal
    dict(color=(0.7, 0.2, 0.2), roughness=0.4, anisotropic=0.8, specular=1.0, specularTint=0.7),
    dict(color=(0.2, 0.5, 0.8), roughness=0.5, sheen=1.0, sheenTint=0.5, subsurface=0.6),  # cloth-like
    dict(color=(0.1, 0.3, 0.1), roughness=0.3, clearcoat=1.0, clearcoatGloss=0.8),         # car paint
    dict(color=(0.9, 0.95, 1.0), roughness=0.2, transmission=0.7, eta=1.0 / 1.5, absorption=(0.8, 0.9, 1.0)),
    dict(color=(0.5, 0.5, 0.5), roughness=0.9, metallic=0.4, subsurface=1.0, clearcoat=0.5, clearcoatGloss=0.1, sheen=0.3),
]


def _inputs(n, seed):
    rng = np.random.default_rng(seed)
    mat = rng.integers(0, len(MATERIALS), n)
    words = core.disney_words(MATERIALS)[mat]
    color = np.array([m["color"] for m in MATERIALS], f32)[mat]
    trans = (f32(1) - np.array([m.get("absorption", (0, 0, 0)) for m in MATERIALS], f32))[mat]
    unit = lambda v: (v / np.linalg.norm(v, axis=1, keepdims=True)).astype(f32)
    N = unit(rng.normal(size=(n, 3)))
    iN = unit(N + 0.2 * rng.normal(size=(n, 3)))           # a shading normal near the geometric one
    iT = unit(rng.normal(size=(n, 3)))
    wow = unit(rng.normal(size=(n, 3)) + 1.5 * N * np.where(rng.random(n) < 0.85, 1, -1)[:, None])  # mostly front side
    wiw = unit(rng.normal(size=(n, 3)) + N)
    dist = (rng.random(n) * 3).astype(f32)
    r0, r1 = rng.random(n).astype(f32), rng.random(n).astype(f32)
    return dict(mat=mat, words=words, color=color, trans=trans, N=N, iN=iN, iT=iT, wow=wow, wiw=wiw, dist=dist, r0=r0, r1=r1)


def _pack(x, tail):
    n = len(x["mat"])
    a = np.zeros((n, 28), f32)
    a[:, 0:3], a[:, 3:6] = x["color"], x["trans"]
    a[:, 6:9] = x["words"][:, :3].view(f32)
    a[:, 9] = x["words"][:, 3].view(f32)
    a[:, 10:13], a[:, 13:16], a[:, 16:19], a[:, 19:22], a[:, 22:25] = x["iN"], x["N"], x["iT"], x["wow"], tail
    return a


def _close(a, b, rtol=2e-3, atol=1e-5):
    return np.abs(a - b) <= atol + rtol * np.abs(b)


@pytest.fixture(scope="module")
def rc():
    c = core.RenderCoreB200(0)
    yield c
    c.Shutdown()


@pytest.fixture(scope="module")
def ref():
    if not oracle.have_bsdf_ref():
        oracle.build()
    if not oracle.have_bsdf_ref():
        pytest.skip("oracle/_ref/liboracle_bsdf_ref.so not built (needs /root/reference at build time)")
    return oracle.DisneyRef()


def test_evaluate_matches_reference(rc, ref):
    x = _inputs(200_000, 3)
    tint4 = wr.disney_tint(x["color"])
    value, pdf = ref.evaluate(x["color"], x["trans"], tint4, x["words"], x["iN"], x["iT"], x["wow"], x["wiw"])
    out = rc.DisneyProbe(1, _pack(x, x["wiw"]))
    ok = _close(out[:, 3], pdf) & _close(out[:, 4:7], value).all(axis=1)
    print("evaluate: mismatching", int((~ok).sum()), "of", len(ok), "; pdf > 0:", int((pdf > 0).sum()))
    assert (pdf > 0).mean() > 0.5
    assert (~ok).mean() < 1e-4, (out[~ok][:5], value[~ok][:5], pdf[~ok][:5])


def test_sample_matches_reference(rc, ref):
    x = _inputs(200_000, 4)
    tint4 = wr.disney_tint(x["color"])
    wiw, pdf, spec, value = ref.sample(x["color"], x["trans"], tint4, x["words"], x["iN"], x["N"], x["iT"], x["wow"], x["dist"], x["r0"], x["r1"])
    out = rc.DisneyProbe(0, _pack(x, np.stack([x["dist"], x["r0"], x["r1"]], axis=1)))
    # where the reference ends the path (pdf 0, or total internal reflection: zero value, no direction) the core reports pdf 0
    tir = spec & (np.abs(wiw).sum(axis=1) == 0)
    dead = (pdf <= 0) | tir
    assert np.all(out[dead, 3] <= 1e-6) or (out[dead, 3] > 1e-6).mean() < 1e-4
    live = ~dead
    ok = _close(out[live, 3], pdf[live]) & _close(out[live, 4:7], value[live]).all(axis=1) & (np.abs(out[live, 0:3] - wiw[live]).max(axis=1) < 2e-3)
    ok &= (out[live, 7] > 0.5) == spec[live]
    print("sample: live", int(live.sum()), "mismatching", int((~ok).sum()), "specular", int(spec.sum()), "tir", int(tir.sum()))
    assert live.mean() > 0.5 and spec.sum() > 1000
    assert (~ok).mean() < 2e-3, (out[live][~ok][:5], wiw[live][~ok][:5], pdf[live][~ok][:5], value[live][~ok][:5])
