"""Pins the checker of the shade stage.  The image tests compare the CUDA frames with tests/wavefront_ref.py, a numpy
restatement of the reference's wavefront loop; this file holds every shading function of that restatement against the REFERENCE'S OWN
code -- lib/sharedBSDFs/lambert.h and lib/CUDA/shared_kernel_code/lights_shared.h compiled unmodified for the host
(oracle/ref_shade_driver.cpp -> oracle/_ref/liboracle_shade_ref.so) -- on random inputs.  CPU only.

Bars: functions made of + - * / sqrt only are compared BIT FOR BIT (both sides round every operation separately: numpy float32 here,
-ffp-contract=off there); where a transcendental differs by construction (the restatement's diffuse directions use the separately
rounded sin / cos it shares with the device, the reference __sincosf -> libm here) the bar is 2e-6 absolute on unit vectors.
The reference's quirks that the restatement does not repeat are named where they are excluded."""
import numpy as np
import pytest

import wavefront_ref as wr
from lighthouse2_b200 import lights as L

f32 = np.float32


@pytest.fixture(scope="module")
def ref():
    import oracle
    if not oracle.have_shade_ref():
        oracle.build()
    if not oracle.have_shade_ref():
        pytest.skip("oracle/_ref/liboracle_shade_ref.so needs /root/reference at build time")
    return oracle.ShadeRef()


def _unit(v):
    return (v / np.linalg.norm(v, axis=1, keepdims=True)).astype(f32)


def _bits(a):
    return np.ascontiguousarray(a, f32).view(np.uint32)


def _lights():
    area = L.stack([L.area_light((-1, 3, -1), (1, 3, -1), (0, 3, 1), (12, 11, 9), 5, 0), L.area_light((2, 2.5, 0), (2, 2.5, 1), (3, 3.5, 0.5), (6, 7, 8), 9, 2)], 24)
    point = L.stack([L.point_light((0.5, 3.0, -1.0), (40, 36, 30)), L.point_light((-2, 1.5, 2), (9, 9, 12))], 8)
    spot = L.stack([L.spot_light((1.5, 4, 1), (-0.3, -1, -0.2), 0.95, 0.8, (30, 30, 25))], 12)
    directional = L.stack([L.directional_light((0.2, -1, 0.3), (1.5, 1.4, 1.2))], 8)
    return area, point, spot, directional


def _surface_points(n, seed):
    rng = np.random.default_rng(seed)
    I = rng.uniform(-3, 3, (n, 3)).astype(f32)
    I[:, 1] = rng.uniform(-1, 1.5, n).astype(f32)
    N = _unit(rng.normal(size=(n, 3)) + np.array([0, 1.5, 0]))
    return rng, I, N


def test_random_barycentrics_bit_exact(ref):
    r0 = np.random.default_rng(1).random(200_000).astype(f32)
    r0[:4] = (0.0, 0.25, 0.5, 0.99999994)
    assert np.array_equal(_bits(wr.random_barycentrics(r0)), _bits(ref.random_barycentrics(r0)))


def test_safe_origin_bit_exact(ref):
    rng, I, N = _surface_points(100_000, 2)
    R = _unit(rng.normal(size=I.shape))
    assert np.array_equal(_bits(wr.safe_origin(I, R, N, f32(1e-4))), _bits(ref.safe_origin(I, R, N, 1e-4)))


def test_random_point_on_light_matches_the_reference(ref):
    """RandomPointOnLight (lights_shared.h:171-262) with all four light kinds in play: the picked point, the pick probability, the solid-angle
    pdf and the radiance.  Bit for bit, except the radiance of a picked POINT light: the reference assigns it to a local that shadows the
    output (:196), the restatement (like the device code) returns the light's radiance, which is what is meant."""
    area, point, spot, directional = _lights()
    ref.set_lights(area, point, spot, directional)
    lights = wr.Lights(area, point, spot, directional)
    n = 200_000
    rng, I, N = _surface_points(n, 3)
    r0, r1 = rng.random(n).astype(f32), rng.random(n).astype(f32)
    P, pick, pdf, col = wr.random_point_on_light(lights, r0, r1, I, N)
    Pr, pickr, pdfr, colr = ref.random_point_on_light(r0, r1, I, N)
    lit = pickr > 0  # (no light has potential: the reference leaves pickProb unset and returns a dummy point; both sides report pdf 0)
    assert lit.mean() > 0.5
    assert np.array_equal(_bits(pdf), _bits(pdfr))
    assert np.array_equal(_bits(pick[lit]), _bits(pickr[lit])) and (pick[~lit] == 0).all()
    assert np.array_equal(_bits(P[lit]), _bits(Pr[lit]))
    is_point = lit & np.isnan(colr).any(axis=1)  # the reference left lightColor unset: a point light was picked
    by_pos = lit & ((np.abs(Pr[:, None, :] - point[None, :, 0:3]).max(axis=2) == 0).any(axis=1))
    assert np.array_equal(is_point, by_pos) and is_point.sum() > 1000
    assert np.array_equal(_bits(col[lit & ~is_point]), _bits(colr[lit & ~is_point]))
    which = (np.abs(Pr[is_point][:, None, :] - point[None, :, 0:3]).max(axis=2) == 0).argmax(axis=1)
    assert np.array_equal(col[is_point], point[which, 4:7])
    # every kind of light gets picked
    kinds = [(np.abs(P[lit][:, None, :] - p[None, :, 0:3]).max(axis=2) == 0).any(axis=1).sum() for p in (point, spot)]
    assert min(kinds) > 1000 and (P[lit][:, 1] > 2.4).sum() > 1000


def test_light_pick_prob_and_light_pdf_bit_exact(ref):
    """The implicit-connection side of MIS (pathtracer.h:132-142): LightPickProb for an area light seen at a point on it, and CalculateLightPDF."""
    area, point, spot, directional = _lights()
    ref.set_lights(area, point, spot, directional)
    lights = wr.Lights(area, point, spot, directional)
    n = 100_000
    rng, O, N = _surface_points(n, 4)
    idx = rng.integers(0, len(area), n).astype(np.int32)
    b = rng.dirichlet((1, 1, 1), n).astype(f32)
    I = (b[:, 0:1] * area[idx, 12:15] + b[:, 1:2] * area[idx, 16:19] + b[:, 2:3] * area[idx, 20:23]).astype(f32)
    assert np.array_equal(_bits(wr.light_pick_prob(lights, idx, O, N, I)), _bits(ref.light_pick_prob(idx, O, N, I)))
    # CalculateLightPDF( D, t, area, N ) = t^2 / (-dot( D, N ) area): the restatement computes it inline as t * t / (ddn * area) with ddn = -dot( D, N )
    D = _unit(I - O)
    t = np.linalg.norm(I - O, axis=1).astype(f32)
    ddn = -wr.dot(D, area[idx, 4:7])
    mine = ((t * t) / (ddn * area[idx, 7])).astype(f32)
    assert np.array_equal(_bits(mine), _bits(ref.calculate_light_pdf(D, t, area[idx, 7], area[idx, 4:7])))
    with pytest.raises(ValueError):
        ref.set_lights(point=np.tile(point[:1], (9, 1)))  # the reference's float[8] cannot hold nine lights


def _materials(n, rng):
    color = rng.uniform(0.05, 1.0, (n, 3)).astype(f32)
    absorption = rng.uniform(0.0, 0.8, (n, 3)).astype(f32)
    # 0.8 fixed point, as the cores hold roughness and transmission (tools_shared.h:44,51); a third of the materials each: rough opaque, glossy, dielectric
    kind = rng.integers(0, 3, n)
    rough_q = np.where(kind == 0, 255, rng.integers(0, 256, n)).astype(np.uint32)
    trans_q = np.where(kind == 2, rng.integers(128, 256, n), 0).astype(np.uint32)
    eta = np.where(kind == 2, rng.uniform(1.1, 1.8, n), 1.0).astype(f32)
    rough = np.maximum(f32(0.001), rough_q.astype(f32) * f32(1.0 / 255.0)).astype(f32)
    transmission = (trans_q.astype(f32) * f32(1.0 / 255.0)).astype(f32)
    params = np.zeros((n, 4), np.uint32)
    params[:, 0], params[:, 2], params[:, 3] = rough_q << np.uint32(24), trans_q << np.uint32(16), eta.view(np.uint32)
    return color, absorption, rough, transmission, eta, params


def test_lambert_sample_matches_the_reference(ref):
    """SampleBSDF of lambert.h:71-123 -- specular reflection, Fresnel reflection / refraction with Beer absorption, cosine-weighted diffuse
    bounce -- for rays arriving from either side of the surface, shading normal != geometric normal."""
    n = 300_000
    rng = np.random.default_rng(6)
    color, absorption, rough, transmission, eta, params = _materials(n, rng)
    N = _unit(rng.normal(size=(n, 3)))
    sN = _unit(N + 0.25 * rng.normal(size=(n, 3)))
    D = _unit(rng.normal(size=(n, 3)))
    t = rng.uniform(0.1, 5.0, n).astype(f32)
    r3, r4 = rng.random(n).astype(f32), rng.random(n).astype(f32)
    face = np.where(wr.dot(D, N) > 0, f32(-1), f32(1)).astype(f32)  # pathtracer.h:156: the side the ray came from
    fN, gF = (sN * face[:, None]).astype(f32), (N * face[:, None]).astype(f32)
    R, pdf, spec, bsdf, _ = wr.lambert_sample(color, absorption, rough, transmission, eta, fN, gF, face, D, t, r3, r4)
    wow = (D * f32(-1)).astype(f32)
    Rr, pdfr, specr, bsdfr = ref.sample(color, absorption, params, sN, N, wow, t, r3, r4)
    # total internal reflection: the reference returns value 0 with pdf 1 and no direction (lambert.h:100), the restatement ends the path
    tir = specr & (np.abs(Rr).sum(axis=1) == 0)
    assert 100 < tir.sum() < 0.1 * n and (pdf[tir] == 0).all() and (bsdfr[tir] == 0).all()
    ok = ~tir
    assert np.array_equal(spec[ok], specr[ok])
    s = ok & spec
    # specular branches: no transcendental but exp() in the Beer term (numpy exp vs expf: a few ulp)
    assert np.array_equal(_bits(R[s]), _bits(Rr[s])) and np.array_equal(_bits(pdf[s]), _bits(pdfr[s]))
    opaque = s & (transmission == 0)
    assert opaque.sum() > 10_000 and np.array_equal(_bits(bsdf[opaque]), _bits(bsdfr[opaque]))
    assert np.allclose(bsdf[s], bsdfr[s], rtol=2e-6, atol=0)
    # diffuse branch: directions agree to the accuracy of the two sin / cos implementations; a direction that lands within that of the geometric
    # horizon may be rejected (pdf = 0, lambert.h:121) on one side only
    d = ok & ~spec
    assert d.sum() > 50_000
    assert np.abs(R[d] - Rr[d]).max() < 2e-6
    assert np.array_equal(_bits(bsdf[d]), _bits(bsdfr[d]))
    horizon = np.abs(wr.dot(gF[d], Rr[d])) < 4e-6
    assert np.allclose(pdf[d][~horizon], pdfr[d][~horizon], rtol=0, atol=2e-6) and horizon.mean() < 1e-3
    print("lambert SampleBSDF: %d specular, %d diffuse, %d TIR; max direction difference %.2e" % (s.sum(), d.sum(), tir.sum(), np.abs(R[d] - Rr[d]).max()))


def test_lambert_evaluate_matches_what_next_event_estimation_uses(ref):
    """EvaluateBSDF (lambert.h:58-69) as the shade stage uses it for next-event estimation (pathtracer.h:184-189): value = color / pi,
    pdf = |cos| / pi, nothing for specular vertices.  The restatement (and the device code) computes the same two numbers inline."""
    n = 100_000
    rng = np.random.default_rng(8)
    color, absorption, rough, transmission, eta, params = _materials(n, rng)
    N = _unit(rng.normal(size=(n, 3)))
    wow, wiw = _unit(rng.normal(size=(n, 3))), _unit(rng.normal(size=(n, 3)))
    value, pdf = ref.evaluate(color, absorption, params, N, wow, wiw)
    specular = (transmission > f32(0.999)) | (rough <= f32(0.001))
    assert (pdf[specular] == 0).all() and (value[specular] == 0).all() and specular.sum() > 50
    assert np.array_equal(_bits(pdf[~specular]), _bits((np.abs(wr.dot(wiw, N)) * wr.INVPI).astype(f32)[~specular]))
    assert np.array_equal(_bits(value[~specular]), _bits((color * wr.INVPI).astype(f32)[~specular]))
