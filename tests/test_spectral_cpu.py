"""CPU tests of the spectral mode (scalar_spectral semantics): the sRGB->spectrum model on the host against the reference's own
outputs, the spectral stage bodies (host simulator) against the oracle, and the loader's conversions."""
import os

import numpy as np
import pytest

import helpers as H
from helpers import HostSim, Oracle
from mitsuba2_b200 import _abi as A
from mitsuba2_b200 import rgb2spec
from mitsuba2_b200 import xml as X
from mitsuba2_b200.scenedata import SceneData


def test_rgb2spec_matches_reference_vectors():
    """rgb2spec.py (table optimisation + fetch) vs ext/rgb2spec/{rgb2spec_opt.cpp, rgb2spec.c} run at resolution 16
    (tests/golden/make_rgb2spec_vectors.py)."""
    g = np.load(os.path.join(H.ROOT, "tests", "golden", "rgb2spec_ref16.npz"))
    coeff = rgb2spec.fetch(g["rgb"], model=rgb2spec.table(16))
    spectra = rgb2spec.model_eval(coeff, g["wavelengths"])
    assert np.abs(spectra - g["spectra"]).max() < 1e-4          # measured 7e-6
    np.testing.assert_allclose(coeff, g["coeff"], rtol=2e-3, atol=1e-5)


def test_rgb2spec_table_equals_reference_table_live():
    path = os.path.join(H.ROOT, "oracle", "_ref", "srgb_16.coeff")
    if not os.path.exists(path):
        pytest.skip("oracle/_ref/srgb_16.coeff not built (reference tree absent)")
    raw = open(path, "rb").read()
    assert raw[:4] == b"SPEC" and np.frombuffer(raw[4:8], np.uint32)[0] == 16
    scale = np.frombuffer(raw[8:8 + 64], np.float32)
    data = np.frombuffer(raw[8 + 64:], np.float32).reshape(3, 16, 16, 16, 3)
    s2, d2 = rgb2spec.table(16)
    assert np.array_equal(scale, s2)
    lam = np.linspace(360, 830, 32)
    a, b = rgb2spec.model_eval(data.reshape(-1, 3), lam), rgb2spec.model_eval(d2.reshape(-1, 3), lam)
    # the zero-brightness plane (k = 0) is degenerate: the optimiser has nothing to fit there
    diff = np.abs(a - b).reshape(3, 16, 16 * 16, -1)[:, 1:]
    assert diff.max() < 1e-4


def test_rgb2spec_special_cases_and_roundtrip():
    c = rgb2spec.fetch(np.array([[0, 0, 0], [1, 1, 1]], np.float32), model=rgb2spec.table(16))
    assert np.isneginf(c[0, 2]) and np.isposinf(c[1, 2])               # srgb.cpp:34-37
    assert np.allclose(rgb2spec.model_eval(c, [400.0, 700.0]), [[0, 0], [1, 1]])
    rgb = np.random.default_rng(0).uniform(0.1, 0.9, (64, 3)).astype(np.float32)
    lam = 360 + rgb2spec._LAM01 * 470
    back = rgb2spec.model_eval(rgb2spec.fetch(rgb, model=rgb2spec.table(16)), lam) @ rgb2spec._RGB_TBL.T
    assert np.abs(back - rgb).max() < 0.03                             # resolution-16 interpolation error


def test_spectral_cbox_hostsim_matches_oracle():
    sd = H.spectral(H.cbox_scene(res=16, spp=4))
    a, b = Oracle(sd).render_xyzaw(seed=3), HostSim(sd).render_xyzaw(seed=3)
    assert np.abs(a - b).max() < 2e-3 * max(1.0, np.abs(a).max())
    assert a[..., :3].max() > 0.05


def test_spectral_material_scene_hostsim_close_to_oracle():
    sd = H.spectral(H.material_scene(res=(24, 16), spp=4, named_conductor=True))
    a, b = Oracle(sd).render_xyzaw(seed=1), HostSim(sd).render_xyzaw(seed=1)
    bad = np.abs(a - b).max(axis=2) > 1e-3 * (1 + np.abs(a).max(axis=2))
    assert bad.mean() < 0.02


def test_spectral_n1_plugins_hostsim_matches_oracle():
    """twosided + constant + point emitters in spectral mode (srgb_d65 radiance / intensity spectra)"""
    sd = H.spectral(H.n1_scene(named_conductor=True))
    a, b = Oracle(sd).render_xyzaw(seed=2), HostSim(sd).render_xyzaw(seed=2)
    bad = np.abs(a - b).max(axis=2) > 1e-3 * (1 + np.abs(a).max(axis=2))
    assert bad.mean() < 0.02 and a[..., 1].max() > 0


def test_spectral_render_is_colorimetrically_close_to_rgb():
    """A spectral render of an RGB-authored scene reproduces the rgb render's XYZ film up to metamerism (the upsampled
    spectra integrate back to the authored colours; interreflections differ slightly)."""
    rgb = Oracle(H.cbox_scene(res=12, spp=96)).render_xyzaw(seed=5)
    spec = Oracle(H.spectral(H.cbox_scene(res=12, spp=96))).render_xyzaw(seed=5)
    m_rgb, m_spec = rgb[..., :3].reshape(-1, 3).mean(axis=0), spec[..., :3].reshape(-1, 3).mean(axis=0)
    assert np.all(np.abs(m_spec - m_rgb) < 0.08 * m_rgb), (m_rgb, m_spec)
    assert np.array_equal(rgb[..., 3:], spec[..., 3:])                 # alpha and weights do not depend on the colour mode


def test_spectrum_textures_and_named_conductor():
    sd = SceneData()
    t = sd.add_texture_spectrum([400, 500, 600, 700], [0.1, 0.5, 0.9, 0.3])
    sd.add_diffuse(t)
    with pytest.raises(ValueError):
        sd.to_desc()                                                   # rgb mode takes colours only
    H.spectral(sd)
    desc, keep = sd.to_desc()
    assert desc.mode == A.MODE_SPECTRAL and desc.textures[t].type == A.TEX_IRREGULAR and desc.textures[t].width == 4
    sd2 = H.spectral(SceneData())
    b = sd2.add_roughconductor(material="Cu")
    desc, keep = sd2.to_desc()
    eta = desc.textures[sd2.bsdfs[b]["eta_tex"]]
    assert eta.type == A.TEX_IRREGULAR and eta.width == 55 and abs(eta.data[0] - 302.400421) < 1e-3
    sd3 = H.spectral(SceneData())
    sd3.add_diffuse((1.2, 0.5, 0.5))
    with pytest.raises(ValueError, match="Invalid RGB reflectance"):
        sd3.to_desc()                                                  # srgb.cpp:31-32


def test_emitter_conversion_follows_srgb_d65():
    sd = H.spectral(H.cbox_scene(res=8, spp=1))
    desc, keep = sd.to_desc()
    e = next(i for i, em in enumerate(sd.emitters) if em["type"] == A.EMITTER_AREA)
    tex = desc.textures[desc.emitters[e].radiance_tex]
    rad = np.asarray(sd.emitters[e]["radiance"], np.float32)
    assert tex.type == A.TEX_CONSTANT and abs(tex.d65_scale - 2 * rad.max()) < 1e-5      # srgb_d65.cpp:44-52
    want = rgb2spec.fetch(rad / (2 * rad.max()), model=sd.rgb2spec_model)
    assert np.allclose(list(tex.value), want)


def test_xml_spectrum_tag_and_variant():
    xml = """<scene version="2.0.0">
      <integrator type="path"><integer name="max_depth" value="3"/></integrator>
      <sensor type="perspective"><float name="fov" value="40"/>
        <film type="hdrfilm"><integer name="width" value="8"/><integer name="height" value="8"/></film></sensor>
      <bsdf type="diffuse" id="m"><spectrum name="reflectance" value="400:0.2, 550:0.8, 700:0.4"/></bsdf>
      <shape type="ply"><string name="filename" value="%s"/><ref id="m"/></shape>
      <emitter type="constant"><rgb name="radiance" value="1.0"/></emitter>
    </scene>"""
    import tempfile
    from mitsuba2_b200 import mesh as M
    with tempfile.TemporaryDirectory() as d:
        M.write_ply(os.path.join(d, "q.ply"), M.quad([-1, -1, 3], [1, -1, 3], [1, 1, 3], [-1, 1, 3]))
        sd = X.load_string(xml % os.path.join(d, "q.ply"), base_dir=d)
    H.spectral(sd)
    desc, keep = sd.to_desc()
    assert any(desc.textures[i].type == A.TEX_IRREGULAR for i in range(desc.n_textures))
    img = Oracle(sd).render_xyzaw(seed=0, spp=2)
    assert np.isfinite(img).all() and img[..., 1].max() > 0
