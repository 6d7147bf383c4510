"""CPU tests that PIN THE ORACLE: known-answer values taken from the reference's own unit tests (file:line cited at
each test), the golden vectors produced by the reference's header-only arithmetic (tests/golden/ref_vectors.npz, see
make_ref_vectors.py) and -- where oracle/_ref/libmts_ref.so exists -- a live comparison on a second input set."""
import math

import numpy as np
import pytest

import helpers as H
from helpers import Oracle, _p
from mitsuba2_b200 import _abi as A
from mitsuba2_b200 import mesh as M
from mitsuba2_b200.scenedata import SceneData, look_at

# comparison tolerances oracle <-> reference arithmetic (relative, absolute).  Everything not listed must be bit-equal.
TOL = {
    "fresnel": (1e-6, 2e-7),              # fma contraction differences
    "erfinv": (1e-6, 3e-7),               # different polynomial for the same function
    "microfacet_sample": (2e-4, 1e-4),    # visible-normal sampling amplifies erfinv/sincos ulps (microfacet.h:330-413)
    # hierarchical warp: a few ulps in a level's interval_to_linear are amplified near patch borders
    "h2d_sample_a": (1e-4, 1e-4), "h2d_sample_b": (1e-4, 1e-4), "h2d_sample_c": (1e-4, 1e-4),
    "h2d_eval_a": (3e-6, 1e-6), "h2d_eval_b": (3e-6, 1e-6), "h2d_eval_c": (3e-6, 1e-6),
    # spectral variants: libm atanh/cosh/rsqrt vs enoki's vectorised approximations
    "sample_wavelength": (3e-6, 1e-5), "srgb_model_eval": (0, 3e-6), "regular_eval": (0, 2e-7), "irregular_eval": (0, 2e-7),
    "spectrum_to_xyz": (1e-6, 3e-7),
}


def _compare(got, want):
    for k, w in want.items():
        g = got[k]
        assert g.shape == w.shape, k
        if k in TOL:
            np.testing.assert_allclose(g, w, rtol=TOL[k][0], atol=TOL[k][1], err_msg=k)
        else:
            assert np.array_equal(g, w, equal_nan=(g.dtype.kind == "f")), k


def test_oracle_matches_reference_vectors():
    want = dict(np.load(H.os.path.join(H.ROOT, "tests", "golden", "ref_vectors.npz")))
    _compare(H.math_cases(H.oracle_math()), want)


def test_oracle_matches_reference_live():
    ref = H.ref_math()
    if ref is None:
        pytest.skip("oracle/_ref/libmts_ref.so not built (reference tree absent)")
    _compare(H.math_cases(H.oracle_math(), seed=1234), H.math_cases(ref, seed=1234))


# ---- PCG32 / TEA ----------------------------------------------------------------------------------------------
def test_pcg32_published_vector():
    """pcg32 demo stream (initstate 42, initseq 54): 0xa15c02b7 0x7b47f409 0xba1d3330 0x83d2f293 0xbfa4784b 0xcbed606e;
    next_float32 keeps the top 23 bits (enoki random.h:96-101)."""
    want = [0xa15c02b7, 0x7b47f409, 0xba1d3330, 0x83d2f293, 0xbfa4784b, 0xcbed606e]
    st, sq = np.array([42], np.uint64), np.array([54], np.uint64)
    out = np.zeros((1, 6), np.float32)
    Oracle.lib().orc_pcg32_floats(1, _p(st), _p(sq), 6, _p(out))
    assert [int(round(float(f) * 2**23)) for f in out[0]] == [w >> 9 for w in want]


def test_tea_floats():
    """src/libcore/tests/test_random.py:6-28"""
    l = Oracle.lib()
    f32 = {(1, 1): 0.5424730777740479, (1, 2): 0.5079904794692993, (1, 3): 0.4171961545944214, (1, 4): 0.008385419845581055,
           (1, 5): 0.8085528612136841, (2, 1): 0.6939879655838013, (3, 1): 0.6978365182876587, (4, 1): 0.4897364377975464}
    f64 = {(1, 1): 0.5424730799533735, (1, 2): 0.5079905082233922, (1, 3): 0.4171962610608142, (1, 4): 0.008385529523330604,
           (1, 5): 0.80855288317879, (2, 1): 0.6939880404156831, (3, 1): 0.6978365636630994, (4, 1): 0.48973647949223253}
    for (a, b), want in f32.items():
        u = l.orc_tea32(a, b, 4)
        got = np.array([(u >> 9) | 0x3f800000], np.uint32).view(np.float32)[0] - np.float32(1)
        assert float(got) == want
    for (a, b), want in f64.items():
        u = l.orc_tea64(a, b, 4)
        got = np.array([(u >> 12) | 0x3ff0000000000000], np.uint64).view(np.float64)[0] - 1.0
        assert got == want


def test_sampler_seed_streams_distinct():
    st, inc = np.zeros(4096, np.uint64), np.zeros(4096, np.uint64)
    Oracle.lib().orc_sampler_seed(4096, 7, _p(st), _p(inc))
    assert len(set(zip(st.tolist(), inc.tolist()))) == 4096


# ---- Fresnel (src/librender/tests/test_fresnel.py:7-88) ---------------------------------------------------------
def _fresnel(c, eta):
    o = np.zeros(4, np.float32)
    Oracle.lib().orc_fresnel(float(c), float(eta), _p(o))
    return o


def test_fresnel_kat():
    ct_crit = -math.sqrt(1 - 1 / 1.5**2)
    np.testing.assert_allclose(_fresnel(1, 1.5), (0.04, -1, 1.5, 1 / 1.5), rtol=1e-5)
    np.testing.assert_allclose(_fresnel(-1, 1.5), (0.04, 1, 1 / 1.5, 1.5), rtol=1e-5)
    np.testing.assert_allclose(_fresnel(1, 1 / 1.5), (0.04, -1, 1 / 1.5, 1.5), rtol=1e-5)
    np.testing.assert_allclose(_fresnel(-1, 1 / 1.5), (0.04, 1, 1.5, 1 / 1.5), rtol=1e-5)
    np.testing.assert_allclose(_fresnel(0, 1.5), (1, ct_crit, 1.5, 1 / 1.5), rtol=1e-5)
    np.testing.assert_allclose(_fresnel(0, 1 / 1.5), (1, 0, 1 / 1.5, 1.5), rtol=1e-5, atol=1e-7)
    F, ct, _, scale = _fresnel(math.cos(math.radians(45)), 1.5)
    assert abs(ct + math.cos(math.radians(28.1255057020557))) < 1e-6
    assert abs(F - 0.5 * (0.09201336304552442**2 + 0.3033370452904235**2)) < 1e-6
    assert abs((scale * math.sin(math.radians(45)))**2 + ct**2 - 1) < 1e-6
    F, ct, _, _ = _fresnel(math.cos(math.radians(45)), 1 / 1.5)
    assert abs(F - 1) < 1e-6 and abs(ct) < 1e-6
    F, ct, _, _ = _fresnel(math.cos(math.radians(10)), 1 / 1.5)
    assert abs(ct + math.cos(math.radians(15.098086605159006))) < 1e-6
    assert abs(F - 0.5 * (0.19046797197779405**2 + 0.20949431963852014**2)) < 1e-6


def test_fresnel_conductor_equals_dielectric_for_real_ior():
    for eta in (1.5, 1 / 1.5):
        for th in np.linspace(0, math.pi / 2, 20):
            c = math.cos(th)
            assert abs(_fresnel(c, eta)[0] - Oracle.lib().orc_fresnel_conductor(float(c), float(eta), 0.0)) < 2e-6


def test_snell():
    th = np.linspace(0, math.pi / 2, 20)
    ct = np.array([_fresnel(math.cos(t), 1.5)[1] for t in th])
    np.testing.assert_allclose(np.sin(th) - 1.5 * np.sin(np.arccos(-ct)), 0, atol=1e-5)


# ---- microfacet (src/librender/tests/test_microfacet.py:18-190; values from Mitsuba 0.6) -------------------------
def _dirs(theta, phi):
    return np.stack([np.cos(phi) * np.sin(theta), np.sin(phi) * np.sin(theta), np.cos(theta)], -1).astype(np.float32)


def test_microfacet_beckmann_tables():
    l = Oracle.lib()
    B = A.MF_BECKMANN
    wi = np.array([0, 0, 1], np.float32)
    v = _dirs(np.linspace(0, math.pi, 20, dtype=np.float32), np.full(20, math.pi / 2, np.float32))
    ev = [l.orc_microfacet_eval(B, 0.1, 0.3, _p(x)) for x in v]
    np.testing.assert_allclose(ev[:7], [1.06103287e+01, 8.22650051e+00, 3.57923722e+00, 6.84863329e-01, 3.26460004e-02,
                                        1.01964230e-04, 5.87322635e-10], rtol=2e-4)
    assert all(e == 0 for e in ev[7:])
    pd = [l.orc_microfacet_pdf(B, 0.1, 0.3, 0, _p(wi), _p(x)) for x in v]
    np.testing.assert_allclose(pd[:7], [1.06103287e+01, 8.11430168e+00, 3.38530421e+00, 6.02319300e-01, 2.57622823e-02,
                                        6.90584930e-05, 3.21235011e-10], rtol=2e-4)
    ev_i = [l.orc_microfacet_eval(B, 0.1, 0.1, _p(x)) for x in v]
    np.testing.assert_allclose(ev_i[:4], [3.18309879e+01, 2.07673073e+00, 3.02855828e-04, 1.01591990e-11], rtol=3e-4)
    v = _dirs(np.full(20, 0.1, np.float32), np.linspace(0, 2 * math.pi, 20, dtype=np.float32))
    ref = [3.95569706, 4.34706259, 5.54415846, 7.4061389, 9.17129803, 9.62056446, 8.37803268, 6.42071199, 4.84459257, 4.05276537,
           4.05276537, 4.84459257, 6.42071199, 8.37803268, 9.62056446, 9.17129803, 7.4061389, 5.54415846, 4.34706259, 3.95569706]
    np.testing.assert_allclose([l.orc_microfacet_eval(B, 0.1, 0.3, _p(x)) for x in v], ref, rtol=2e-5)
    np.testing.assert_allclose([l.orc_microfacet_eval(B, 0.1, 0.1, _p(x)) for x in v], 11.86709118, rtol=2e-5)


def test_microfacet_smith_g1_tables():
    l = Oracle.lib()
    B = A.MF_BECKMANN
    wi = np.array([0, 0, 1], np.float32)
    v = _dirs(np.linspace(math.pi / 3, math.pi / 2, 20, dtype=np.float32), np.full(20, math.pi / 2, np.float32))
    ref = [1.0, 1.0, 1.0, 1.0000523, 9.9941480e-01, 9.9757767e-01, 9.9420297e-01, 9.8884594e-01, 9.8091525e-01, 9.6961778e-01,
           9.5387781e-01, 9.3222123e-01, 9.0260512e-01, 8.6216795e-01, 8.0686140e-01, 7.3091686e-01, 6.2609726e-01, 4.8074335e-01,
           2.7883825e-01, 1.9197471e-06]
    np.testing.assert_allclose([l.orc_microfacet_smith_g1(B, 0.1, 0.3, _p(x), _p(wi)) for x in v], ref, rtol=2e-4, atol=2e-6)
    v = _dirs(np.full(20, math.pi / 2 * 0.98, np.float32), np.linspace(0, 2 * math.pi, 20, dtype=np.float32))
    ref = [0.67333597, 0.56164336, 0.42798978, 0.35298213, 0.31838724, 0.31201753, 0.33166203, 0.38421196, 0.48717275, 0.63746351,
           0.63746351, 0.48717275, 0.38421196, 0.33166203, 0.31201753, 0.31838724, 0.35298213, 0.42798978, 0.56164336, 0.67333597]
    np.testing.assert_allclose([l.orc_microfacet_smith_g1(B, 0.1, 0.3, _p(x), _p(wi)) for x in v], ref, rtol=5e-5)


def test_microfacet_sample_beckmann_table():
    """test_microfacet.py:142-190 (sample_visible = false), first rows of the 6x6 grid"""
    l = Oracle.lib()
    wi = np.array([0, 0, 1], np.float32)
    ref = {(0.2, 0.0): [4.71862517e-02, 1.23754589e-08, 9.98886108e-01], (0.4, 0.0): [7.12896436e-02, 1.86970155e-08, 9.97455657e-01],
           (0.2, 0.2): [1.44650340e-02, 1.33556545e-01, 9.90935624e-01], (0.6, 0.4): [-7.61397704e-02, 1.65956154e-01, 9.83189344e-01]}
    for (sx, sy), want in ref.items():
        o = np.zeros(4, np.float32)
        l.orc_microfacet_sample(A.MF_BECKMANN, 0.1, 0.3, 0, _p(wi), np.float32(sx), np.float32(sy), _p(o))
        np.testing.assert_allclose(o[:3], want, rtol=1e-4, atol=1e-6)


# ---- reconstruction filters (src/rfilters/tests/test_rfilter.py:8-58) --------------------------------------------
def test_rfilter_spot_values():
    l = Oracle.lib()
    assert l.orc_rfilter_eval(A.RFILTER_BOX, 0.5, 0.0, 0.49) == 1 and l.orc_rfilter_eval(A.RFILTER_BOX, 0.5, 0.0, 0.51) == 0
    assert abs(l.orc_rfilter_eval(A.RFILTER_GAUSSIAN, 2.0, 0.5, 0.2) - 0.9227) < 8e-3
    assert l.orc_rfilter_eval(A.RFILTER_GAUSSIAN, 2.0, 0.5, 2.1) == 0
    assert abs(l.orc_rfilter_eval(A.RFILTER_TENT, 1.0, 0.0, 0.1) - 0.903) < 5e-2
    assert l.orc_rfilter_eval(A.RFILTER_TENT, 1.0, 0.0, 1.1) == 0
    # test_rfilter.py:24-45 (lanczos 3 lobes, mitchell B = C = 1/3, catmullrom)
    e2 = l.orc_rfilter_eval2
    assert abs(e2(A.RFILTER_LANCZOS, 3.0, 0.0, 0.0, 1.4) - (-0.14668)) < 1e-2 and e2(A.RFILTER_LANCZOS, 3.0, 0.0, 0.0, 3.1) == 0
    assert e2(A.RFILTER_LANCZOS, 3.0, 0.0, 0.0, 0.0) == 1
    assert abs(e2(A.RFILTER_MITCHELL, 2.0, 1 / 3, 1 / 3, 0.0) - 0.8888) < 1e-3 and e2(A.RFILTER_MITCHELL, 2.0, 1 / 3, 1 / 3, 2.1) == 0
    assert abs(e2(A.RFILTER_CATMULLROM, 2.0, 0.0, 0.0, 0.0) - 0.9765) < 5e-2 and e2(A.RFILTER_CATMULLROM, 2.0, 0.0, 0.0, 2.1) == 0
    assert e2(A.RFILTER_CATMULLROM, 2.0, 0.0, 0.0, 1.5) < 0                      # negative lobe


# ---- BSDF spot checks through the oracle scene hooks -----------------------------------------------------------------
def _bsdf_scene(add):
    sd = SceneData()
    q = M.quad([-1, 0, -1], [-1, 0, 1], [1, 0, 1], [1, 0, -1])
    q.bsdf = add(sd)
    sd.add_mesh(q)
    sd.add_perspective(look_at([0, 3, 0], [0, 0, 0], [0, 0, 1]), 45.0, 0.01, 100.0, 8, 8, "box")
    return sd


def test_dielectric_spot_check():
    """src/bsdfs/tests/test_dielectric.py:150-172"""
    sd = _bsdf_scene(lambda s: s.add_dielectric(1.5, 1.0, 0.3, 0.6))
    o = Oracle(sd)
    ang = math.radians(80)
    wi = np.array([[math.sin(ang), 0, math.cos(ang)]], np.float32)
    uv = np.zeros((1, 2), np.float32)
    r = o.bsdf_sample(0, wi, uv, np.array([0], np.float32), np.zeros((1, 2), np.float32))[0]   # wo[3], pdf, eta, weight[3], type
    assert abs(r[3] - 0.387704354691473) < 1e-6
    np.testing.assert_allclose(r[:3], [-math.sin(ang), 0, math.cos(ang)], atol=1e-6)
    r = o.bsdf_sample(0, wi, uv, np.array([1], np.float32), np.zeros((1, 2), np.float32))[0]
    a2 = math.radians(41.03641052520335)
    assert abs(r[3] - (1 - 0.387704354691473)) < 1e-6
    np.testing.assert_allclose(r[:3], [-math.sin(a2), 0, -math.cos(a2)], atol=1e-6)
    r2 = o.bsdf_sample(0, r[None, :3].copy(), uv, np.array([1], np.float32), np.zeros((1, 2), np.float32))[0]
    np.testing.assert_allclose(r2[:3], wi[0], atol=1e-6)


def test_diffuse_eval_pdf():
    """src/bsdfs/tests/test_diffuse.py:16-38"""
    o = Oracle(_bsdf_scene(lambda s: s.add_diffuse(0.5)))
    th = np.arange(20) / 19.0 * (math.pi / 2)
    wo = np.stack([np.sin(th), np.zeros(20), np.cos(th)], -1).astype(np.float32)
    wi = np.tile(np.array([[0, 0, 1]], np.float32), (20, 1))
    ev, pdf = o.bsdf_eval_pdf(0, wi, np.zeros((20, 2), np.float32), wo)
    np.testing.assert_allclose(pdf, wo[:, 2] / math.pi, atol=1e-6)
    np.testing.assert_allclose(ev[:, 0], 0.5 * wo[:, 2] / math.pi, atol=1e-6)
