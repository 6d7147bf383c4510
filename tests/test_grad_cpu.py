"""Gradient pass (texture adjoint, csrc/path.h:texture_adjoint) checked on the CPU: the product's replay adjoint, run by
the host simulator, against central finite differences of the ORACLE render at identical seeds.  The reference has no
implementation to compare with (parity unpinned); finite differences of the pinned forward model are
the ground truth."""
import numpy as np

import helpers as H
from helpers import HostSim, Oracle


def _develop(o, film):
    return o.develop(film)


def test_texture_gradient_matches_finite_differences_of_oracle():
    sd, tex = H.textured_cbox(res=12, spp=24, tex_res=5)
    rng = np.random.default_rng(0)
    wgt = rng.normal(size=(12, 12, 3)).astype(np.float32)          # L = sum(wgt * image)  =>  dL/dimage = wgt
    g = HostSim(sd).render_grad(wgt, tex, seed=5)
    assert g.shape == sd.textures[tex]["data"].shape and np.isfinite(g).all() and np.abs(g).max() > 0
    o = Oracle(sd)
    base = sd.textures[tex]["data"].copy()

    def loss(data):
        o.update_texture(tex, data)
        return float(np.sum(np.float64(wgt) * o.develop(o.render_xyzaw(seed=5))))

    eps = 0.02
    picks = [(y, x, c) for y, x, c in zip(rng.integers(0, 5, 24), rng.integers(0, 5, 24), rng.integers(0, 3, 24))]
    fd, an = [], []
    for y, x, c in picks:
        d = base.copy(); d[y, x, c] += eps
        lp = loss(d)
        d[y, x, c] -= 2 * eps
        fd.append((lp - loss(d)) / (2 * eps))
        an.append(g[y, x, c])
    fd, an = np.array(fd), np.array(an)
    scale = np.abs(fd).max()
    assert scale > 0
    np.testing.assert_allclose(an, fd, rtol=2e-2, atol=2e-3 * scale)
    cos = float(np.dot(an, fd) / (np.linalg.norm(an) * np.linalg.norm(fd)))
    assert cos > 0.999


def test_texture_gradient_is_linear_in_dl_dimage_and_additive_over_sample_ranges():
    sd, tex = H.textured_cbox(res=8, spp=6, tex_res=4)
    h = HostSim(sd)
    rng = np.random.default_rng(1)
    w1, w2 = rng.normal(size=(2, 8, 8, 3)).astype(np.float32)
    g1, g2, g12 = h.render_grad(w1, tex, seed=2), h.render_grad(w2, tex, seed=2), h.render_grad(w1 + 2 * w2, tex, seed=2)
    np.testing.assert_allclose(g12, g1 + 2 * g2, rtol=1e-4, atol=1e-5)
    parts = h.render_grad(w1, tex, seed=2, spp=2, spp_begin=0, spp_total=6) + h.render_grad(w1, tex, seed=2, spp=4, spp_begin=2, spp_total=6)
    np.testing.assert_allclose(parts, g1, rtol=1e-4, atol=1e-5)
