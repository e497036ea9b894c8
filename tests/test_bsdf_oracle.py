"""CPU-side checks of the BSDF checker itself (oracle/_ref/liboracle_bsdf_ref.so: the reference's disney.h compiled for the host)."""
import numpy as np
import pytest


def test_reference_disney_bsdf_is_self_consistent():
    """The reference's Disney BSDF compiled for the host (oracle/ref_bsdf_driver.cpp): a sampled direction evaluates to
    the same value and pdf (SampleBSDF sums the lobes EvaluateBSDF sums, disney.h:227-348) for materials without sheen
    (EvaluateBSDF lets the sheen value replace the diffuse one, :329-330) and away from the dielectric branch."""
    import oracle
    if not oracle.have_bsdf_ref():
        oracle.build()
    if not oracle.have_bsdf_ref():
        pytest.skip("oracle/_ref/liboracle_bsdf_ref.so needs /root/reference at build time")
    from lighthouse2_b200 import core
    import wavefront_ref as wr
    d = oracle.DisneyRef()
    n = 20000
    rng = np.random.default_rng(2)
    mats = [dict(color=(0.8, 0.5, 0.3), roughness=0.5, clearcoat=0.8, clearcoatGloss=0.5, metallic=0.3, anisotropic=0.5)]
    words = np.repeat(core.disney_words(mats), n, axis=0)
    color = np.repeat(np.array([mats[0]["color"]], np.float32), n, axis=0)
    zero = np.zeros((n, 3), np.float32)
    unit = lambda v: (v / np.linalg.norm(v, axis=1, keepdims=True)).astype(np.float32)
    N = unit(rng.normal(size=(n, 3)))
    T = unit(np.cross(N, rng.normal(size=(n, 3))))
    wow = unit(rng.normal(size=(n, 3)) + 2 * N)
    wow = np.where((np.einsum("ij,ij->i", wow, N) < 0.05)[:, None], N, wow).astype(np.float32)
    tint4 = wr.disney_tint(color)
    wiw, pdf, spec, value = d.sample(color, zero, tint4, words, N, N, T, wow, np.ones(n, np.float32), rng.random(n).astype(np.float32), rng.random(n).astype(np.float32))
    assert not spec.any()
    live = (pdf > 1e-4) & (np.einsum("ij,ij->i", wiw, N) > 1e-3)
    assert live.mean() > 0.9
    v2, p2 = d.evaluate(color, zero, tint4, words, N, T, wow, wiw)
    assert np.allclose(p2[live], pdf[live], rtol=1e-3, atol=1e-6)
    assert np.allclose(v2[live], value[live], rtol=1e-3, atol=1e-6)
    # consistent normals: alpha = 0 keeps the normal
    cn = d.consistent_normal(wow, N, np.zeros(n, np.float32))
    assert np.abs(cn - N).max() < 1e-3
