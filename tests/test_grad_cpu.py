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


# ---- vertex-position adjoint (csrc/adjoint.h), run by the host simulator ----------------------------------------------
import pytest


def _ramp(res):
    yy, xx = np.mgrid[0:res, 0:res]
    return np.stack([xx / res, yy / res, 0.5 + 0 * xx], -1).astype(np.float32)


@pytest.mark.parametrize("bsdf", ["diffuse", "beckmann", "ggx"])
def test_vertex_gradient_interior_point_light_matches_fd(bsdf):
    """A plane filling the view under a point light: no visibility change, no random light sampling -> the adjoint and the
    finite difference of the oracle are the same estimator sample by sample.  Exercises the dual-number BSDF derivative,
    the geometry-term adjoint and the slide of x1 along the camera ray.  Tolerance 2e-3 relative (fp32 FD)."""
    sd = H.vgrad_scene("plane", bsdf, res=16, spp=32)
    wgt = np.random.default_rng(1).uniform(0.5, 1.5, (16, 16, 3)).astype(np.float32)
    g = HostSim(sd).render_grad_vertices(wgt, 0, seed=3, boundary=False).reshape(-1, 3)
    P0 = np.asarray(sd.meshes[0].positions, np.float32).reshape(-1, 3)
    F = H.displacement_fields(P0)
    for name in ("tz", "tilt"):
        fd = H.oracle_fd_vertices(sd, 0, wgt, F[name].astype(np.float32), 2e-3, 32, 3)
        ad = float((g * F[name]).sum())
        assert abs(ad - fd) <= 2e-3 * abs(fd) + 1e-3, (bsdf, name, ad, fd)
    for name in ("tx", "ty"):                                    # in-plane motion changes nothing
        assert abs(float((g * F[name]).sum())) < 1e-3


def test_vertex_gradient_smooth_normals_matches_fd():
    """Smooth-shaded mesh, direct light only: at a fixed pixel the hit point slides over the triangle, so the interpolated
    normal changes through the barycentrics, and the vertex normals themselves are re-derived from the moved positions by
    `parameters_changed` (mesh.cpp:794-804, angle-weighted face normals) -- both are differentiated exactly (the second by
    vnormal_adjoint_body after the replay).  FD = oracle with recomputed normals.  Global fields 1 %, single vertices 1 %
    of the largest derivative."""
    sd = H.vgrad_scene("wavy", "diffuse", res=16, spp=8)
    wgt = np.random.default_rng(1).uniform(0.5, 1.5, (16, 16, 3)).astype(np.float32)
    g = HostSim(sd).render_grad_vertices(wgt, 0, seed=3, boundary=True).reshape(-1, 3)     # no silhouette in view: boundary = 0
    P0 = np.asarray(sd.meshes[0].positions, np.float32).reshape(-1, 3)
    F = H.displacement_fields(P0)
    fds = {n: H.oracle_fd_vertices(sd, 0, wgt, F[n].astype(np.float32), 1e-3, 8, 3) for n in F}
    ref = max(abs(v) for v in fds.values())
    for n, fd in fds.items():
        assert abs(float((g * F[n]).sum()) - fd) < 0.01 * ref, (n, fd)
    single = []
    for vi, ax in ((6 * 13 + 6, 2), (6 * 13 + 7, 2), (5 * 13 + 6, 0), (7 * 13 + 5, 2)):      # vertices near the view centre
        D = np.zeros_like(P0); D[vi, ax] = 1.0
        single.append((H.oracle_fd_vertices(sd, 0, wgt, D, 5e-3, 8, 3), float(g[vi, ax])))
    ref1 = max(abs(a) for a, _ in single)
    for fd, ad in single:
        assert abs(ad - fd) < 0.01 * ref1, single


def test_vertex_gradient_area_light_and_indirect_matches_fd_statistically():
    """Area light + a wall (MIS, emitter hits, one indirect bounce): the detached adjoint and the same-seed finite
    difference are different estimators of the same derivative; they must agree within 4 standard errors over seeds."""
    sd = H.vgrad_scene("corner", "diffuse", res=12, spp=32)
    wgt = np.random.default_rng(1).uniform(0.5, 1.5, (12, 12, 3)).astype(np.float32)
    P0 = np.asarray(sd.meshes[0].positions, np.float32).reshape(-1, 3)
    D = H.displacement_fields(P0)["tz"].astype(np.float32)
    ad, fd = [], []
    for seed in range(1, 7):
        g = HostSim(sd).render_grad_vertices(wgt, 0, seed=seed, boundary=False).reshape(-1, 3)
        ad.append(float((g * D).sum())); fd.append(H.oracle_fd_vertices(sd, 0, wgt, D, 2e-3, 32, seed))
    ad, fd = np.array(ad), np.array(fd)
    se = np.sqrt(ad.var(ddof=1) / len(ad) + fd.var(ddof=1) / len(fd))
    assert abs(ad.mean() - fd.mean()) < 4 * se + 0.01 * abs(fd.mean()), (ad.mean(), fd.mean(), se)


def test_vertex_gradient_boundary_term_matches_fd():
    """A quad in front of a constant background: the image changes only through the silhouette (edge-sampling term).
    FD of a discontinuous integrand needs a wide step and many samples; tolerance 6 % of the largest derivative."""
    sd = H.vgrad_scene("quadenv", "diffuse", res=16, spp=128)
    wgt = _ramp(16)
    g = HostSim(sd).render_grad_vertices(wgt, 0, seed=3, boundary=True, edge_fraction=1.0).reshape(-1, 3)
    gi = HostSim(sd).render_grad_vertices(wgt, 0, seed=3, boundary=False).reshape(-1, 3)
    P0 = np.asarray(sd.meshes[0].positions, np.float32).reshape(-1, 3)
    F = H.displacement_fields(P0)
    fds = {n: H.oracle_fd_vertices(sd, 0, wgt, F[n].astype(np.float32), 0.02, 1024, 3) for n in ("tx", "ty", "scale")}
    ref = max(abs(v) for v in fds.values())
    for n, fd in fds.items():
        ad = float((g * F[n]).sum())
        assert abs(ad - fd) < 0.06 * ref, (n, ad, fd)
        assert abs(float((gi * F[n]).sum())) < 0.01 * ref          # the interior term alone sees nothing


def test_vertex_gradient_shadow_boundary_matches_fd():
    """A quad floating above a lit plane: most of the derivative is the quad's moving shadow / sky occlusion on the plane,
    i.e. the boundary term seen from the first path vertex (boundary=2).  Without it the scale derivative even has the
    wrong sign.  FD of the oracle, wide step; tolerance 12 % of the largest derivative (Monte Carlo on both sides)."""
    sd = H.vgrad_scene("quadshadow", "diffuse", res=16, spp=128)
    wgt = _ramp(16)
    h = HostSim(sd)
    # both sides are Monte Carlo estimates (a single seed of either scatters by ~8 %): average the adjoint over 4 seeds and the
    # differences over 2 seeds at 2048 spp.  Measured over 8 / 4 seeds: tx -6.83 +- 0.08 vs -7.14 +- 0.11, tz -1.55 +- 0.05 vs
    # -1.37 +- 0.09, scale -8.01 +- 0.10 vs -8.03 +- 0.19.
    g2 = np.mean([h.render_grad_vertices(wgt, 0, seed=s_, boundary=2, edge_fraction=1.0).reshape(-1, 3) for s_ in (3, 4, 5, 6)], axis=0)
    g1 = h.render_grad_vertices(wgt, 0, seed=3, boundary=1, edge_fraction=1.0).reshape(-1, 3)
    P0 = np.asarray(sd.meshes[0].positions, np.float32).reshape(-1, 3)
    F = H.displacement_fields(P0)
    fds = {n: float(np.mean([H.oracle_fd_vertices(sd, 0, wgt, F[n].astype(np.float32), 0.02, 2048, s_) for s_ in (3, 4)])) for n in ("tx", "tz", "scale")}
    ref = max(abs(v) for v in fds.values())
    for n, fd in fds.items():
        assert abs(float((g2 * F[n]).sum()) - fd) < 0.10 * ref, (n, float((g2 * F[n]).sum()), fd)
    assert abs(float((g1 * F["scale"]).sum()) - fds["scale"]) > 0.5 * ref      # the term matters here


def test_vertex_gradient_sample_ranges_add_up():
    """Splitting the samples over two calls (ranks) gives the same gradient as one call: interior term exactly (float
    association), boundary term because edge-sample indices are derived from the sample range."""
    sd = H.vgrad_scene("quadshadow", "diffuse", res=8, spp=6)
    wgt = _ramp(8)
    h = HostSim(sd)
    full = h.render_grad_vertices(wgt, 0, seed=2, spp=6, boundary=2, edge_fraction=0.5)
    parts = h.render_grad_vertices(wgt, 0, seed=2, spp=2, spp_begin=0, spp_total=6, boundary=2, edge_fraction=0.5) + \
        h.render_grad_vertices(wgt, 0, seed=2, spp=4, spp_begin=2, spp_total=6, boundary=2, edge_fraction=0.5)
    assert np.allclose(full, parts, rtol=1e-4, atol=1e-5 * np.abs(full).max())


def test_vertex_gradient_texture_sliding_matches_fd():
    """A bitmap-textured plane: moving it in its own plane slides the texture under the fixed pixels (the barycentrics of
    x1 change -> the interpolated uv changes -> the texel mix changes).  Direct light only; FD of the oracle, 2 %."""
    sd = H.vgrad_scene("texplane", "diffuse", res=16, spp=8)
    wgt = np.random.default_rng(1).uniform(0.5, 1.5, (16, 16, 3)).astype(np.float32)
    g = HostSim(sd).render_grad_vertices(wgt, 0, seed=3, boundary=False).reshape(-1, 3)
    P0 = np.asarray(sd.meshes[0].positions, np.float32).reshape(-1, 3)
    F = H.displacement_fields(P0)
    fds = {n: H.oracle_fd_vertices(sd, 0, wgt, F[n].astype(np.float32), 1e-3, 8, 3) for n in ("tx", "ty", "tz", "scale")}
    ref = max(abs(v) for v in fds.values())
    assert abs(fds["tx"]) > 0.02 * ref and abs(fds["ty"]) > 0.02 * ref              # the sliding term is not negligible here
    for n, fd in fds.items():
        assert abs(float((g * F[n]).sum()) - fd) < 0.02 * ref, (n, float((g * F[n]).sum()), fd)


def test_vertex_gradient_smooth_sphere_silhouette_and_terminator():
    """A smooth-shaded sphere lit from the camera's side on black.  Its image has two moving discontinuities: the polygon
    silhouette and, a fraction of a pixel inside it, the shading-normal terminator (a one-sided BSDF is black where the
    interpolated normal looks away from the camera).  Interior term + silhouette pairs + terminator samples (adjoint.h,
    kTermTag) against Richardson-extrapolated finite differences of the oracle: scale within 1.5 %, translations within
    1.2 % of the scale derivative (they are differences of two large opposite rims).  Without the terminator samples the
    scale derivative is 4 % short (HS_NO_TERM=1 in the host simulator reproduces that)."""
    sd = H.vgrad_scene("smoothsphere", "diffuse", res=32, spp=64)
    yy, xx = np.mgrid[0:32, 0:32]
    wgt = np.stack([0.5 + xx / 32, 0.5 + yy / 32, 1.0 + 0 * xx], -1).astype(np.float32)
    h = HostSim(sd)
    P0 = np.asarray(sd.meshes[0].positions, np.float32).reshape(-1, 3)
    F = H.displacement_fields(P0)
    g = np.mean([h.render_grad_vertices(wgt, 0, seed=s_, boundary=1, edge_fraction=1.0).reshape(-1, 3) for s_ in range(3, 7)], axis=0)
    fd = {}
    for n in ("tx", "ty", "scale"):
        f1 = H.oracle_fd_vertices(sd, 0, wgt, F[n].astype(np.float32), 0.005, 1024, 3)
        f2 = H.oracle_fd_vertices(sd, 0, wgt, F[n].astype(np.float32), 0.01, 1024, 3)
        fd[n] = f1 + (f1 - f2) / 3
    an = {n: float((g * F[n]).sum()) for n in fd}
    assert abs(an["scale"] - fd["scale"]) < 0.015 * abs(fd["scale"]), (an, fd)
    assert abs(an["tx"] - fd["tx"]) < 0.012 * abs(fd["scale"]) and abs(an["ty"] - fd["ty"]) < 0.012 * abs(fd["scale"]), (an, fd)
    g0 = h.render_grad_vertices(wgt, 0, seed=3, boundary=0, edge_fraction=1.0).reshape(-1, 3)
    assert abs(float((g0 * F["scale"]).sum()) - fd["scale"]) > 0.1 * abs(fd["scale"])      # the boundary terms matter here


@pytest.mark.parametrize("rfilter", ["gaussian", "mitchell"])
def test_gradients_under_wide_reconstruction_filters(rfilter):
    """Filters wider than a pixel (the film's default is gaussian, film.cpp:46-50): a sample's adjoint weight is the gather of
    dL/dI over its footprint divided by the film's weight channel (path.h film_gather).  The film positions do not depend on the
    parameters, so the texture and the interior vertex term equal finite differences of the oracle at identical seeds up to
    rounding; the film-boundary term is held to 1024-spp differences."""
    sd, tex = H.textured_cbox(res=16, spp=32, tex_res=6)
    H.set_rfilter(sd, rfilter)
    wgt = np.random.default_rng(0).normal(size=(16, 16, 3)).astype(np.float32)
    g = HostSim(sd).render_grad(wgt, tex, seed=5).reshape(sd.textures[tex]["data"].shape)
    scale = np.abs(g).max()
    o = H.Oracle(sd)
    base = sd.textures[tex]["data"].copy()

    def loss(data):
        o.update_texture(tex, data)
        return float(np.sum(np.float64(wgt) * o.develop(o.render_xyzaw(seed=5))))

    for y, x, c in ((1, 2, 0), (3, 3, 1), (4, 0, 2), (2, 5, 1)):
        d = base.copy(); d[y, x, c] += 0.02
        lp = loss(d)
        d[y, x, c] -= 0.04
        assert abs(g[y, x, c] - (lp - loss(d)) / 0.04) <= 1e-2 * scale
    sd = H.set_rfilter(H.vgrad_scene("plane", "beckmann", res=16, spp=32), rfilter)
    wgt = np.random.default_rng(1).uniform(0.5, 1.5, (16, 16, 3)).astype(np.float32)
    gv = HostSim(sd).render_grad_vertices(wgt, 0, seed=3, boundary=False).reshape(-1, 3)
    F = H.displacement_fields(np.asarray(sd.meshes[0].positions, np.float32).reshape(-1, 3))
    for name in ("tz", "tilt"):
        a, fd = float((gv * F[name]).sum()), H.oracle_fd_vertices(sd, 0, wgt, F[name].astype(np.float32), 2e-3, 32, 3)
        assert abs(a - fd) <= 3e-3 * abs(fd) + 1e-3, (name, a, fd)
    sd = H.set_rfilter(H.vgrad_scene("quadenv", "diffuse", res=16, spp=128), rfilter)
    yy, xx = np.mgrid[0:16, 0:16]
    wgt = np.stack([xx / 16, yy / 16, 0.5 + 0 * xx], -1).astype(np.float32)
    h = HostSim(sd)
    gv = np.mean([h.render_grad_vertices(wgt, 0, seed=s_, boundary=1, edge_fraction=1.0).reshape(-1, 3) for s_ in (3, 4, 5, 6)], axis=0)
    F = H.displacement_fields(np.asarray(sd.meshes[0].positions, np.float32).reshape(-1, 3))
    fds = {n: H.oracle_fd_vertices(sd, 0, wgt, F[n].astype(np.float32), 0.02, 1024, 3) for n in ("tx", "scale")}
    ref = max(abs(v) for v in fds.values())
    for n, fd in fds.items():
        assert abs(float((gv * F[n]).sum()) - fd) < 0.04 * ref, (n, float((gv * F[n]).sum()), fd)


def test_vertex_gradient_boundary_term_of_the_second_path_vertex():
    """`boundary=3`: the edge-sampling term of adjoint.h evaluated at the deeper path vertices too.  A wall is lit by a floor that
    carries the soft shadow of the moving quad; only image rows that see the wall count, so the quad and its shadow act through one
    diffuse bounce: the quad occludes the floor as seen from the wall (term of x1) and the light as seen from the floor (term of
    x2).  With the x1 term alone the vertical translation and the scale derivative are far off (7.2 / 7.4 against finite
    differences of 5.3 / 1.9); with the x2 term they agree within the Monte Carlo noise of the two estimates."""
    res = 16
    sd = H.vgrad_scene("indirectshadow", "diffuse", res=res, spp=256)
    wgt = np.zeros((res, res, 3), np.float32)
    wgt[:res // 2 - 1] = 1.0
    h = HostSim(sd)
    F = H.displacement_fields(np.asarray(sd.meshes[0].positions, np.float32).reshape(-1, 3))
    fd = {n: float(np.mean([H.oracle_fd_vertices(sd, 0, wgt, F[n].astype(np.float32), 0.03, 4096, s_) for s_ in (3, 4)])) for n in ("ty", "tz", "scale")}
    g2 = np.mean([h.render_grad_vertices(wgt, 0, seed=s_, boundary=2, edge_fraction=1.0).reshape(-1, 3) for s_ in range(3, 7)], axis=0)
    g3 = np.mean([h.render_grad_vertices(wgt, 0, seed=s_, boundary=3, edge_fraction=1.0).reshape(-1, 3) for s_ in range(3, 11)], axis=0)
    a2 = {n: float((g2 * F[n]).sum()) for n in fd}
    a3 = {n: float((g3 * F[n]).sum()) for n in fd}
    ref = max(abs(v) for v in fd.values())
    assert abs(a2["scale"] - fd["scale"]) > 0.5 * ref and abs(a2["ty"] - fd["ty"]) > 0.2 * ref, (a2, fd)     # the x2 term matters here
    for n in fd:                                                                                        # standard errors: 0.4 - 0.5
        assert abs(a3[n] - fd[n]) < 0.25 * ref, (n, a3, fd)


@pytest.mark.parametrize("kind,bsdf", [("planearea", "diffuse"), ("corner", "beckmann")])
def test_vertex_gradient_of_an_emitter_mesh(kind, bsdf):
    """The differentiated mesh carries the area light (outside the view: the image changes only through the illumination).
    Emitter samples are material points of the moving mesh -- triangle and barycentrics as the sampler fixed them, the
    attached-surface factor |N . d| / r^3 of vertex_triple with the unnormalised normal carries the area Jacobian, the
    sampling probabilities and MIS weights are detached -- and a hit by BSDF sampling is an ordinary path vertex.  Against
    finite differences of the oracle: translations, scale (changes the light's area) and tilt."""
    sd = H.vgrad_scene(kind, bsdf, res=16, spp=256)
    wgt = np.random.default_rng(1).uniform(0.5, 1.5, (16, 16, 3)).astype(np.float32)
    h = HostSim(sd)
    g = np.mean([h.render_grad_vertices(wgt, 1, seed=s_).reshape(-1, 3) for s_ in (3, 4, 5, 6)], axis=0)
    F = H.displacement_fields(np.asarray(sd.meshes[1].positions, np.float32).reshape(-1, 3))
    fds = {n: float(np.mean([H.oracle_fd_vertices(sd, 1, wgt, F[n].astype(np.float32), 0.02, 2048, s_) for s_ in (3, 4)])) for n in ("tx", "ty", "tz", "scale")}
    ref = max(abs(v) for v in fds.values())
    for n, fd in fds.items():
        assert abs(float((g * F[n]).sum()) - fd) < 0.015 * abs(fd) + 0.004 * ref, (n, float((g * F[n]).sum()), fd)


def test_x1_boundary_term_against_analytic_form_factors():
    """Ground truth without finite-difference noise for the boundary term seen from the first path vertex: on the `quadshadow`
    scene a plane pixel that does not see the quad has radiance albedo * (1 - F), F = the form factor of the quad seen from the
    plane point (closed form for a polygon: sum over edges of angle * n . unit(v_i x v_{i+1}) / 2 pi), so d(sum of those pixels) /
    d(quad motion) follows from differentiating F numerically in double.  The adjoint (8 seeds) must agree within 3 % for the
    motion along the view axis, the in-plane motions and the scale; same-seed differences of the renderer scatter by 12 % per
    pair at 2048 spp on this loss (16 pairs: -7.27 +- 0.22 against -7.081), which is why they are not the yardstick here."""
    res = 24
    sd = H.vgrad_scene("quadshadow", "diffuse", res=res, spp=128)
    h, o = HostSim(sd), Oracle(sd)
    yy, xx = np.mgrid[0:res, 0:res]
    om, dm = o.sensor_sample_ray(np.stack([(xx.ravel() + 0.5) / res, (yy.ravel() + 0.5) / res], 1).astype(np.float32))
    prim = h.intersect_closest(om, dm).view(np.uint32)[:, 3].reshape(res, res)
    cover = prim < len(sd.meshes[0].faces)
    pad, dil = np.pad(cover, 2), np.zeros_like(cover)
    for dy in range(5):
        for dx in range(5):
            dil |= pad[dy:dy + res, dx:dx + res]
    mask = ~dil                                                   # plane pixels at least two pixels away from the quad
    wgt = np.repeat(mask[..., None], 3, 2).astype(np.float32)
    P = np.asarray(sd.meshes[0].positions, np.float64).reshape(-1, 3)
    lo, hi, centre = P.min(0), P.max(0), P.mean(0)
    base = [np.array([lo[0], lo[1], lo[2]]), np.array([hi[0], lo[1], lo[2]]), np.array([hi[0], hi[1], lo[2]]), np.array([lo[0], hi[1], lo[2]])]

    def form_factor(x, corners):
        v = [c - x for c in corners]
        v = [q / np.linalg.norm(q) for q in v]
        tot = 0.0
        for i in range(4):
            a, b = v[i], v[(i + 1) % 4]
            cr = np.cross(a, b)
            tot += np.arccos(np.clip(np.dot(a, b), -1, 1)) * np.dot([0.0, 0.0, -1.0], cr / np.linalg.norm(cr))
        return abs(tot) / (2 * np.pi)

    S = 3
    sub = (np.arange(S) + 0.5) / S
    pts = []
    for py in range(res):
        for px in range(res):
            if mask[py, px]:
                oo, dd = o.sensor_sample_ray(np.array([[(px + a) / res, (py + b) / res] for b in sub for a in sub], np.float32))
                for k in range(S * S):
                    O, D = oo[k, :3].astype(np.float64), dd[k, :3].astype(np.float64)
                    pts.append(O + (2.0 - O[2]) / D[2] * D)      # the plane z = 2
    albedo_sum = 0.3 + 0.5 + 0.7                                  # the plane's reflectance, summed over the channels; sky radiance 1

    def total(shift, s):
        corners = [q + shift + s * (q - centre) for q in base]
        return sum(albedo_sum * (1.0 - form_factor(x, corners)) for x in pts) / (S * S)

    e = 1e-3
    exact = {"tz": (total(np.array([0, 0, e]), 0) - total(np.array([0, 0, -e]), 0)) / (2 * e),
             "tx": (total(np.array([e, 0, 0]), 0) - total(np.array([-e, 0, 0]), 0)) / (2 * e),
             "scale": (total(np.zeros(3), e) - total(np.zeros(3), -e)) / (2 * e)}
    F = H.displacement_fields(P.astype(np.float32))
    g = np.mean([h.render_grad_vertices(wgt, 0, seed=s_, boundary=2, edge_fraction=1.0).reshape(-1, 3) for s_ in range(3, 11)], axis=0)
    ref = max(abs(v) for v in exact.values())
    for n, ex in exact.items():
        a = float((g * F[n]).sum())
        assert abs(a - ex) < 0.03 * abs(ex) + 0.004 * ref, (n, a, ex)


def test_deeper_boundary_term_against_a_quadrature_of_the_transport():
    """The boundary term of the second path vertex against a deterministic answer.  `indirectshadow` under a uniform sky with
    max_depth 3: the wall's pixels receive sky light (static), the floor's reflection -- the floor is lit by the sky minus the
    closed-form form factors of the quad and of the wall (plus the part both count), and is hidden from the wall where the quad is
    in the way -- and the reflection of the quad's own lit top.  Integrating that over the floor and the quad on a grid and
    differencing it in the quad's motion gives ty / tz / scale = -4.34 / 2.46 / -4.21 (the grid moves them by about 0.1; the floor
    radiance of the model was checked against renders at four points: 0.1762 / 0.1776, 0.613 / 0.6156, ...).  The adjoint with the x1
    term only misses a third of the scale derivative (-2.95); with the deeper term it gives -4.15 / 2.30 / -4.10."""
    res = 16
    sd = H.vgrad_scene("indirectshadow", "diffuse_sky", res=res, spp=256)
    o, h = Oracle(sd), HostSim(sd)
    rows = res // 2 - 1
    wgt = np.zeros((res, res, 3), np.float32)
    wgt[:rows] = 1.0
    Pq = np.asarray(sd.meshes[0].positions, np.float64).reshape(-1, 3)
    qlo, qhi, qc = Pq.min(0), Pq.max(0), Pq.mean(0)
    alb_wall, alb_quad, alb_floor = np.array([0.3, 0.5, 0.7]), np.array([0.6, 0.5, 0.4]), 0.8

    def ff_poly(X, n, corners):          # form factor of a polygon seen from the points X with normal n
        V = [c[None] - X for c in corners]
        V = [v / np.linalg.norm(v, axis=1, keepdims=True) for v in V]
        tot = np.zeros(len(X))
        for i in range(len(corners)):
            a, b = V[i], V[(i + 1) % len(corners)]
            cr = np.cross(a, b)
            tot += np.arccos(np.clip((a * b).sum(1), -1, 1)) * (cr @ n) / np.maximum(np.linalg.norm(cr, axis=1), 1e-30)
        return np.abs(tot) / (2 * np.pi)

    def ff_rect(X, x0, x1_, zlo, z1, y):          # rectangle [x0, x1] x [zlo, z1] in the plane y, zlo per point
        mk = lambda x, z: np.stack([np.full(len(X), x), np.full(len(X), y), np.broadcast_to(z, (len(X),))], 1)  # noqa: E731
        V = [q - X for q in (mk(x0, zlo), mk(x1_, zlo), mk(x1_, z1), mk(x0, z1))]
        V = [v / np.linalg.norm(v, axis=1, keepdims=True) for v in V]
        tot = np.zeros(len(X))
        for i in range(4):
            a, b = V[i], V[(i + 1) % 4]
            cr = np.cross(a, b)
            ln = np.linalg.norm(cr, axis=1)
            tot += np.where(ln > 1e-14, np.arccos(np.clip((a * b).sum(1), -1, 1)) * cr[:, 1] / np.maximum(ln, 1e-30), 0.0)
        return np.abs(tot) / (2 * np.pi)

    oo, dd = o.sensor_sample_ray(np.array([[(px + a) / res, (py + b) / res] for py in range(rows) for px in range(res)
                                           for b in (0.25, 0.75) for a in (0.25, 0.75)], np.float32))
    O, D = oo[:, :3].astype(np.float64), dd[:, :3].astype(np.float64)
    X1 = O + ((1.0 - O[:, 2]) / D[:, 2])[:, None] * D                                   # the wall z = 1, 2 x 2 points per pixel
    NX, NZ, NQ = 360, 210, 32
    GX, GZ = np.meshgrid((np.arange(NX) + 0.5) / NX * 12 - 6, (np.arange(NZ) + 0.5) / NZ * 7 - 6)
    X2 = np.stack([GX.ravel(), np.full(GX.size, -1.0), GZ.ravel()], 1)                  # the floor in front of the wall
    dA, up = (12 / NX) * (7 / NZ), np.array([0, 1.0, 0])
    FW = ff_poly(X2, up, [np.array([-4, -1, 1.0]), np.array([4, -1, 1.0]), np.array([4, 4, 1.0]), np.array([-4, 4, 1.0])])

    def total(shift, s):
        c = [q + shift + s * (q - qc) for q in (np.array([qlo[0], qlo[1], qlo[2]]), np.array([qhi[0], qlo[1], qlo[2]]),
                                                  np.array([qhi[0], qlo[1], qhi[2]]), np.array([qlo[0], qlo[1], qhi[2]]))]
        x0, x1_, z0, z1, yq = c[0][0], c[1][0], c[0][2], c[2][2], c[0][1]
        # seen from a floor point the quad hides part of the wall: the part of the quad with z >= zl (beyond the floor point and below
        # the line that projects onto the wall's top edge) is counted by both form factors
        zl = np.minimum(np.maximum(np.maximum(z0, X2[:, 2] + (yq + 1) / 5.0 * (1 - X2[:, 2])), X2[:, 2]), z1)
        both = np.where(zl < z1 - 1e-12, ff_rect(X2, x0, x1_, zl, z1, yq), 0.0)
        Lf = alb_floor * (1 - ff_poly(X2, up, c) - FW + both)
        u = (np.arange(NQ) + 0.5) / NQ
        XQ = np.array([c[0] + a * (c[1] - c[0]) + b * (c[3] - c[0]) for b in u for a in u])
        dAq = np.linalg.norm(np.cross(c[1] - c[0], c[3] - c[0])) / (NQ * NQ)
        LQ = 1 - ff_poly(XQ, up, [np.array([-4, yq, 1.0]), np.array([4, yq, 1.0]), np.array([4, 4, 1.0]), np.array([-4, 4, 1.0])])
        floor_part = quad_part = 0.0
        for x in X1:
            d = X2 - x[None]
            r2 = (d * d).sum(1)
            K = np.clip(-d[:, 2], 0, None) * np.clip(-d[:, 1], 0, None) / (np.pi * r2 * r2)
            t = (yq - x[1]) / d[:, 1]
            qx, qz = x[0] + t * d[:, 0], x[2] + t * d[:, 2]
            floor_part += (K * ~((qx > x0) & (qx < x1_) & (qz > z0) & (qz < z1)) * Lf).sum() * dA
            d = XQ - x[None]
            r2 = (d * d).sum(1)
            quad_part += (np.clip(-d[:, 2], 0, None) * np.clip(-d[:, 1], 0, None) / (np.pi * r2 * r2) * LQ).sum() * dAq
        return (alb_wall.sum() * floor_part + (alb_wall * alb_quad).sum() * quad_part) / 4

    e = 0.04
    exact = {"ty": (total(np.array([0, e, 0]), 0) - total(np.array([0, -e, 0]), 0)) / (2 * e),
             "tz": (total(np.array([0, 0, e]), 0) - total(np.array([0, 0, -e]), 0)) / (2 * e),
             "scale": (total(np.zeros(3), e) - total(np.zeros(3), -e)) / (2 * e)}
    F = H.displacement_fields(Pq.astype(np.float32))
    g2 = np.mean([h.render_grad_vertices(wgt, 0, seed=s_, boundary=2, edge_fraction=1.0).reshape(-1, 3) for s_ in range(3, 7)], axis=0)
    g3 = np.mean([h.render_grad_vertices(wgt, 0, seed=s_, boundary=3, edge_fraction=1.0).reshape(-1, 3) for s_ in range(3, 11)], axis=0)
    ref = max(abs(v) for v in exact.values())
    print("quadrature", exact, "adjoint x1 term only", {n: float((g2 * F[n]).sum()) for n in exact}, "with the deeper term", {n: float((g3 * F[n]).sum()) for n in exact})
    assert abs(float((g2 * F["scale"]).sum()) - exact["scale"]) > 0.25 * ref
    for n, ex in exact.items():
        assert abs(float((g3 * F[n]).sum()) - ex) < 0.07 * ref, (n, float((g3 * F[n]).sum()), ex)
