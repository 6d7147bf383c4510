"""GPU parity tests of the spectral mode (`cuda_b200_spectral`, scalar_spectral semantics) through the C ABI vs the oracle."""
import numpy as np
import pytest

import helpers as H

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def mi():
    import mitsuba2_b200 as mi
    mi.set_variant("cuda_b200_spectral")
    yield mi
    mi.set_variant("cuda_b200_rgb")


@pytest.mark.parametrize("scene_fn,spp,tol", [
    (lambda: H.spectral(H.cbox_scene(64, 16)), 16, 2e-3),
    (lambda: H.spectral(H.cbox_scene(48, 8, "gaussian")), 8, 2e-3),
    (lambda: H.spectral(H.material_scene((96, 64), 16, named_conductor=True)), 16, 2e-2),
    (lambda: H.spectral(H.n1_scene(named_conductor=True)), 8, 2e-2),
])
def test_spectral_render_matches_oracle_per_sample_streams(mi, scene_fn, spp, tol):
    """Same PCG32 streams (and therefore the same 4 wavelengths per path) on both sides."""
    sd = scene_fn()
    scene, orc = mi.Scene(sd), H.Oracle(sd)
    scene.film_clear()
    st = scene.render_accumulate(spp=spp, seed=7)
    f_g = scene.film_xyzaw()
    cnt = np.zeros(4, np.uint64)
    f_c = orc.render_xyzaw(spp=spp, seed=7, counts=cnt)
    assert st["samples"] == cnt[0]
    assert abs(int(st["closest_rays"]) - int(cnt[1])) <= 1e-3 * cnt[1] + 2
    assert np.allclose(f_g[..., 4], f_c[..., 4], rtol=1e-4, atol=1e-4)
    assert np.allclose(f_g[..., 3], f_c[..., 3], rtol=1e-4, atol=1e-3)
    num, den = np.abs(f_g[..., :3] - f_c[..., :3]).sum(), np.abs(f_c[..., :3]).sum()
    assert den > 0 and num / den < tol, num / den
    d = np.abs(f_g[..., :3] - f_c[..., :3]).max(axis=2) / (np.abs(f_c[..., :3]).max(axis=2) + 1e-2)
    assert (d > 1e-2).mean() < 0.02


def test_spectral_sample_sharding_is_exact_union(mi):
    sd = H.spectral(H.cbox_scene(32, 8))
    scene = mi.Scene(sd)
    scene.film_clear(); scene.render_accumulate(spp=8, seed=3, spp_begin=0, spp_total=8); full = scene.film_xyzaw()
    scene.film_clear()
    scene.render_accumulate(spp=4, seed=3, spp_begin=0, spp_total=8)
    scene.render_accumulate(spp=4, seed=3, spp_begin=4, spp_total=8)
    assert np.allclose(full, scene.film_xyzaw(), rtol=1e-5, atol=1e-5)


def test_spectral_z_test_against_oracle_with_independent_seeds(mi):
    sd = H.spectral(H.material_scene(res=(32, 24), spp=16, named_conductor=True))
    sc = mi.Scene(sd)
    batches = np.stack([sc.render(spp=16, seed=100 + k) for k in range(32)])
    o = H.Oracle(sd)
    ref_batches = np.stack([o.develop(o.render_xyzaw(spp=16, seed=1000 + k)) for k in range(32)])
    mean, ref = batches.mean(axis=0), ref_batches.mean(axis=0)
    var_per_sample = np.maximum(batches.var(axis=0, ddof=1), ref_batches.var(axis=0, ddof=1)) * 16
    frac = H.z_test_pass_fraction(mean, ref, var_per_sample, 32 * 16, 32 * 16)
    assert frac >= 0.9975, frac


def test_spectral_close_to_rgb_and_errors(mi):
    """The spectral film of an RGB-authored scene is colorimetrically close to the rgb film; unsupported combinations fail loudly."""
    sd_rgb, sd_spec = H.cbox_scene(32, 64), H.spectral(H.cbox_scene(32, 64))
    def film_mean(sd):
        sc = mi.Scene(sd)
        sc.film_clear(); sc.render_accumulate(spp=64, seed=1)
        return sc.film_xyzaw()[..., :3].reshape(-1, 3).mean(axis=0)
    a, b = film_mean(sd_rgb), film_mean(sd_spec)
    assert np.all(np.abs(a - b) < 0.08 * a), (a, b)
    sd, tex = H.textured_cbox()
    H.spectral(sd)
    s2 = mi.Scene(sd)
    with pytest.raises(Exception, match="RGB only|rgb only|MTSB_MODE_RGB"):
        s2.render_grad(np.ones((16, 16, 3), np.float32), tex_id=tex, spp=4)
