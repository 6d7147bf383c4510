"""The C oracle (oracle/mts_oracle.c) against the UNMODIFIED reference built from /root/reference by
oracle/build_reference.sh (`dist/mitsuba`, scalar_rgb) and queried through oracle/ref_query.cpp: kd-tree hit ids,
BSDF / emitter / sensor functions on identical inputs, and converged images (the reference's own z-test,
src/librender/tests/test_renders.py:60-126).  Skipped where the reference build is absent.  CPU only."""
import numpy as np
import pytest

import helpers as H

pytestmark = pytest.mark.skipif(not H.RefBuild.available(query=True), reason="reference build (oracle/_ref/mitsuba) not present")


def _prims(h):
    return np.ascontiguousarray(h[:, 3]).view(np.uint32)


def _mixed_rays(sd, orc, n, seed=0):
    rng = np.random.default_rng(seed)
    om_c, dm_c = orc.sensor_sample_ray(rng.random((n // 2, 2)))
    lo = np.min([m.positions.min(axis=0) for m in sd.meshes], axis=0)
    hi = np.max([m.positions.max(axis=0) for m in sd.meshes], axis=0)
    om_i, dm_i = H.random_rays(n // 2, seed + 1, radius=1.0)
    om_i[:, :3] = (lo + hi) / 2 + om_i[:, :3] * (hi - lo) * 0.6
    return np.concatenate([om_c, om_i]), np.concatenate([dm_c, dm_i])


@pytest.mark.parametrize("scene_fn,n", [(lambda: H.cbox_scene(32, 4), 20000), (lambda: H.sphere_scene(128, 256), 100000)])
def test_hit_ids_oracle_vs_reference_kdtree(scene_fn, n):
    """`Scene::ray_intersect` of the reference (SAH kd-tree + Moeller-Trumbore) vs the oracle (median-split BVH, same
    triangle test): identical primitive ids except equal-t ties, which the two accelerators visit in different orders."""
    sd = scene_fn()
    orc, ref = H.Oracle(sd), H.RefBuild(sd)
    om, dm = _mixed_rays(sd, orc, n)
    h_o, h_r = orc.intersect_closest(om, dm), ref.intersect(om, dm)
    mism = _prims(h_o) != _prims(h_r)
    with np.errstate(invalid="ignore"):
        tie = np.abs(h_o[:, 0] - h_r[:, 0]) <= 1e-5 * np.maximum(1, np.abs(h_r[:, 0]))
    assert (mism & ~tie).sum() == 0, f"{(mism & ~tie).sum()} hard mismatches of {n}"
    assert mism.mean() < 5e-3
    ok = ~mism & np.isfinite(h_r[:, 0])
    assert np.array_equal(np.isfinite(h_o[:, 0]), np.isfinite(h_r[:, 0]) | mism & np.isfinite(h_o[:, 0]))
    assert np.allclose(h_o[ok, 0], h_r[ok, 0], rtol=2e-5, atol=1e-6)      # the reference is built with FMA contraction (-ffp-contract=fast)
    # any-hit
    dm2 = dm.copy()
    dm2[:, 3] = np.random.default_rng(5).random(n) * 3.0 * max(1.0, float(np.abs(om[:, :3]).max()) / 3.0)
    a, b = orc.intersect_any(om, dm2).astype(bool), ref.ray_test(om, dm2)
    assert np.mean(a != b) < 1e-4


def test_sensor_rays_match_reference():
    sd = H.cbox_scene(64, 4)
    pos = np.random.default_rng(1).random((5000, 2))
    o_o, d_o = H.Oracle(sd).sensor_sample_ray(pos)
    o_r, d_r = H.RefBuild(sd).sensor_sample_ray(pos)
    assert np.allclose(o_o, o_r, rtol=1e-5, atol=1e-5) and np.allclose(d_o, d_r, rtol=1e-5, atol=1e-5)


def _bsdf_inputs(n, seed, both_sides=False):
    rng = np.random.default_rng(seed)
    wi, wo = H._hemi(rng, n), H._hemi(rng, n)
    if both_sides:
        wi[n // 2:, 2] *= -1
        wo[::3, 2] *= -1
    return wi, rng.random((n, 2)).astype(np.float32), rng.random(n).astype(np.float32), rng.random((n, 2)).astype(np.float32), wo


@pytest.mark.parametrize("shape,name", [(0, "diffuse+bitmap"), (1, "roughconductor"), (2, "dielectric")])
def test_bsdf_functions_match_reference(shape, name):
    """`BSDF::sample/eval/pdf` of the reference plugins vs the oracle's restatement, same inputs."""
    sd = H.material_scene()
    orc, ref = H.Oracle(sd), H.RefBuild(sd)
    wi, uv, s1, s2, wo = _bsdf_inputs(4000, 7 + shape, both_sides=(name == "dielectric"))
    r = ref.bsdf(shape, wi, uv, s1, s2, wo)
    bid = sd.meshes[shape].bsdf
    smp = orc.bsdf_sample(bid, wi, uv, s1, s2)
    ev, pdf = orc.bsdf_eval_pdf(bid, wi, uv, wo)
    valid = r[:, 3] > 0
    assert np.array_equal(valid, smp[:, 3] > 0) or np.mean(valid != (smp[:, 3] > 0)) < 2e-3
    both = valid & (smp[:, 3] > 0)
    # the rough conductor's visible-normal sampling runs an erfinv + Newton solve: a few 1e-4 relative there
    tol = 2e-3 if name == "roughconductor" else 2e-5
    assert np.allclose(smp[both, :3], r[both, :3], atol=tol), np.abs(smp[both, :3] - r[both, :3]).max()
    assert np.allclose(smp[both, 3], r[both, 3], rtol=max(tol, 1e-4) * 5, atol=1e-6)
    assert np.allclose(smp[both, 4], r[both, 4], rtol=1e-6)
    assert np.allclose(smp[both, 5:8], r[both, 5:8], rtol=max(tol, 1e-4) * 5, atol=1e-5)
    assert np.array_equal(smp[both, 8], r[both, 8])
    assert np.allclose(ev, r[:, 9:12], rtol=1e-4, atol=1e-6)
    assert np.allclose(pdf, r[:, 12], rtol=1e-4, atol=1e-6)


@pytest.mark.parametrize("kind", ["area", "envmap", "both"])
def test_emitter_sampling_matches_reference(kind):
    """`Scene::sample_emitter_direction` (uniform emitter choice with sample reuse, area: `Mesh::sample_position`,
    envmap: `Hierarchical2D` warp) -- the reference has no envmap unit test, this pins it."""
    sd = H.material_scene(envmap=kind != "area", area=kind != "envmap")
    orc, ref = H.Oracle(sd), H.RefBuild(sd)
    rng = np.random.default_rng(3)
    p = (rng.random((4000, 3)).astype(np.float32) - 0.5) * np.array([3, 1, 3], np.float32)
    s2 = rng.random((4000, 2)).astype(np.float32)
    a, b = orc.sample_emitter_direction(p, s2), ref.sample_emitter_direction(p, s2)
    ok = (b[:, 4] > 0) & (a[:, 4] > 0)
    assert np.mean((b[:, 4] > 0) != (a[:, 4] > 0)) < 2e-3
    assert np.allclose(a[ok, :3], b[ok, :3], atol=2e-4)
    close = np.isclose(a[ok, 4], b[ok, 4], rtol=2e-3, atol=1e-6) & np.all(np.isclose(a[ok, 5:8], b[ok, 5:8], rtol=2e-3, atol=1e-5), axis=1)
    assert close.mean() > 0.998, close.mean()          # a texel boundary of the envmap warp can flip for a sample in a thousand


def _ztest(img, ref_img, var, n_img, n_ref):
    return H.z_test_pass_fraction(img.astype(np.float64), ref_img.astype(np.float64), var, n_img, n_ref)


def test_c1_cbox_image_oracle_vs_reference_cli():
    """C1 of BASELINE.json: Cornell box 256 x 256, 64 spp, max_depth 6 rendered by the reference CLI (`scalar_rgb`) vs
    the oracle with its own sample streams: per-pixel z-test (variance from 8 independent oracle renders)."""
    sd = H.cbox_scene(res=256, spp=64)
    ref_img, secs = H.RefBuild(sd).render(spp=256)
    orc = H.Oracle(sd)
    runs = np.stack([orc.render(spp=64, seed=100 + k) for k in range(8)]).astype(np.float64)
    mean, var = runs.mean(axis=0), runs.var(axis=0, ddof=1) * 64          # per-sample variance
    frac = _ztest(mean, ref_img, var, 8 * 64, 256)
    assert frac >= 0.9975, frac
    assert abs(mean.mean() - ref_img.mean()) < 0.01 * ref_img.mean()


def test_material_scene_image_oracle_vs_reference_cli():
    """rough conductor + dielectric + textured diffuse + area light + envmap: every plugin of C3, small."""
    sd = H.material_scene(res=(96, 64), spp=64)
    ref_img, _ = H.RefBuild(sd).render(spp=1024)
    orc = H.Oracle(sd)
    runs = np.stack([orc.render(spp=128, seed=200 + k) for k in range(8)]).astype(np.float64)
    mean, var = runs.mean(axis=0), runs.var(axis=0, ddof=1) * 128
    frac = _ztest(mean, ref_img, var, 8 * 128, 1024)
    assert frac >= 0.9975, frac
    assert abs(mean.mean() - ref_img.mean()) < 0.02 * ref_img.mean()


@pytest.mark.parametrize("per_face", [False, True])
def test_mesh_attribute_textures_oracle_vs_reference_cli(per_face):
    """`mesh_attribute` (src/textures/mesh_attribute.cpp): per-vertex / per-face colours read from the PLY by the reference's
    own loader (find_other_fields, ply.cpp:594-690) vs the oracle's restatement of Mesh::eval_attribute."""
    sd = H.attribute_scene(res=(64, 48), spp=64, per_face=per_face)
    ref_img, _ = H.RefBuild(sd).render(spp=512)
    orc = H.Oracle(sd)
    runs = np.stack([orc.render(spp=64, seed=300 + k) for k in range(16)]).astype(np.float64)
    frac = _ztest(runs.mean(axis=0), ref_img, runs.var(axis=0, ddof=1) * 64, 16 * 64, 512)
    assert frac >= 0.9975, frac
    assert abs(runs.mean() - ref_img.mean()) < 0.01 * ref_img.mean()


# ---- analytic shapes (src/shapes/{sphere,disk,cylinder,rectangle}.cpp) ---------------------------------------------
def test_analytic_shapes_surface_interaction_matches_reference():
    """`Scene::ray_intersect` on a scene of spheres / a cylinder / disks / rectangles + a triangle floor: primitive ids
    (mesh faces first, then one primitive per shape), t, uv, p, n, shading normal and wi (which needs dp_du) of the
    oracle's restatement of `ray_intersect` + `fill_surface_interaction` vs the reference's own plugins; then `ray_test`."""
    sd = H.shapes_scene(emitters="disk sphere cylinder rectangle")
    orc, ref = H.Oracle(sd), H.RefBuild(sd)
    rng = np.random.default_rng(0)
    om_c, dm_c = orc.sensor_sample_ray(rng.random((20000, 2)))
    om_i, dm_i = H.random_rays(20000, 3, radius=4.0)
    om, dm = np.concatenate([om_c, om_i]), np.concatenate([dm_c, dm_i])
    a, b = orc.surface_interaction(om, dm), ref.intersect(om, dm)
    assert np.array_equal(_prims(a), _prims(b))
    assert set(np.unique(_prims(b))) >= set(range(11))                 # every primitive was hit
    ok = np.isfinite(b[:, 0])
    assert np.allclose(a[ok, 0], b[ok, 0], rtol=2e-5, atol=1e-5)
    assert np.abs(a[ok, 1:3] - b[ok, 1:3]).max() < 5e-5                # uv
    assert np.abs(a[ok, 4:7] - b[ok, 4:7]).max() < 5e-5                # p
    assert np.abs(a[ok, 7:16] - b[ok, 7:16]).max() < 1e-4              # n, sh_frame.n, wi
    dm2 = dm.copy()
    dm2[:, 3] = rng.random(len(dm2)) * 6.0
    assert np.mean(orc.intersect_any(om, dm2).astype(bool) != ref.ray_test(om, dm2)) < 1e-4


@pytest.mark.parametrize("kind", ["disk", "sphere", "cylinder", "rectangle"])
def test_analytic_shape_emitter_sampling_matches_reference(kind):
    """Area light on an analytic shape: `Sphere::sample_direction` (cone sampling, sphere.cpp:182-255) and
    `Shape::sample_direction` over the plugins' `sample_position` -- direction, distance, pdf, value and normal."""
    sd = H.shapes_scene(emitters=kind)
    orc, ref = H.Oracle(sd), H.RefBuild(sd)
    rng = np.random.default_rng(3)
    p = (rng.random((4000, 3)).astype(np.float32) - 0.5) * np.array([6, 3, 6], np.float32)
    s2 = rng.random((4000, 2)).astype(np.float32)
    a, b = orc.sample_emitter_direction(p, s2), ref.sample_emitter_direction(p, s2)
    assert np.array_equal(a[:, 4] > 0, b[:, 4] > 0)
    ok = b[:, 4] > 0
    assert np.abs(a[ok, :3] - b[ok, :3]).max() < 1e-5
    assert np.allclose(a[ok, 3], b[ok, 3], rtol=1e-4)
    assert np.allclose(a[ok, 4], b[ok, 4], rtol=2e-3) and np.allclose(a[ok, 5:8], b[ok, 5:8], rtol=2e-3, atol=1e-6)
    assert np.abs(a[ok, 8:11] - b[ok, 8:11]).max() < 2e-4


@pytest.mark.parametrize("emitters,tri_floor", [("disk sphere cylinder rectangle", True), ("sphere", False)])
def test_analytic_shapes_image_oracle_vs_reference_cli(emitters, tri_floor):
    """Converged image of the shapes scene by the reference CLI vs the oracle (z-test, test_renders.py:60-126); the second
    case has no triangle at all (every primitive analytic)."""
    sd = H.shapes_scene(res=(72, 48), spp=64, emitters=emitters, tri_floor=tri_floor)
    ref_img, _ = H.RefBuild(sd).render(spp=1024)
    orc = H.Oracle(sd)
    # 24 runs: an 8-run variance estimate alone fails ~0.3 % of the pixels (heavy-tailed glossy / small-light pixels)
    runs = np.stack([orc.render(spp=128, seed=400 + k) for k in range(24)]).astype(np.float64)
    frac = _ztest(runs.mean(axis=0), ref_img, runs.var(axis=0, ddof=1) * 128, 24 * 128, 1024)
    # the small sphere light seen through the glossy sphere gives heavy-tailed pixels whose variance 24 runs underestimate:
    # measured rms z = 1.02 with direct light only (each of the four emitter shapes), 1.12 at depth 3 -- no bias, fat tails
    assert frac >= (0.9975 if tri_floor else 0.995), frac
    assert abs(runs.mean() - ref_img.mean()) < 0.01 * ref_img.mean()


@pytest.mark.parametrize("kind", ["disk", "sphere", "cylinder", "rectangle"])
def test_analytic_shape_emitters_direct_light_image_vs_reference_cli(kind):
    """Direct illumination only (max_depth 2, light-tailed): next-event estimation on each emitter shape + MIS with
    `pdf_direction` against the reference CLI; every primitive analytic."""
    sd = H.shapes_scene(res=(72, 48), spp=64, emitters=kind, tri_floor=False)
    sd.max_depth = 2
    ref_img, _ = H.RefBuild(sd).render(spp=2048)
    orc = H.Oracle(sd)
    runs = np.stack([orc.render(spp=128, seed=500 + k) for k in range(16)]).astype(np.float64)
    frac = _ztest(runs.mean(axis=0), ref_img, runs.var(axis=0, ddof=1) * 128, 16 * 128, 2048)
    assert frac >= 0.9975, frac
    assert abs(runs.mean() - ref_img.mean()) < 0.005 * ref_img.mean()


# ---- gradients: finite differences of the REFERENCE renderer ----------------------------------------------------------
def _reference_fd(sd, mesh_id, wgt, D, eps, spp):
    """Central finite difference of sum(wgt * image) along the vertex displacement field D, both images rendered by the
    unmodified reference CLI (its sampler is seeded per block, so the two renders share their random numbers)."""
    import copy
    P0 = np.asarray(sd.meshes[mesh_id].positions, np.float32).reshape(-1, 3).copy()
    vals = []
    for sgn in (+1.0, -1.0):
        s2 = copy.copy(sd)
        s2.meshes = list(sd.meshes)
        m = copy.copy(sd.meshes[mesh_id])
        m.positions = (P0 + sgn * eps * D).astype(np.float32)
        s2.meshes[mesh_id] = m
        img, _ = H.RefBuild(s2).render(spp=spp)
        vals.append(float((img.astype(np.float64) * wgt).sum()))
    return (vals[0] - vals[1]) / (2 * eps)


@pytest.mark.parametrize("kind,fields,boundary", [("plane", ("tz", "tilt"), 0), ("quadenv", ("tx", "scale"), 1), ("quadshadow", ("tz", "scale"), 2)])
def test_vertex_gradient_vs_finite_differences_of_the_reference(kind, fields, boundary):
    """The vertex adjoint (the product's stage bodies on the host simulator) against finite differences of the REFERENCE
    renderer itself: interior term (plane under a point light), film boundary (quad on a background), boundary seen from x1
    (quad floating above a lit plane).  Tolerance 6 % / 6 % / 12 % of the largest derivative, as for the oracle's differences."""
    sd = H.vgrad_scene(kind, "diffuse", res=16, spp=128)
    yy, xx = np.mgrid[0:16, 0:16]
    wgt = np.stack([xx / 16, yy / 16, 0.5 + 0 * xx], -1).astype(np.float32)
    h = H.HostSim(sd)
    g = np.mean([h.render_grad_vertices(wgt, 0, seed=s_, boundary=boundary, edge_fraction=1.0).reshape(-1, 3) for s_ in (3, 4, 5, 6)], axis=0)
    P0 = np.asarray(sd.meshes[0].positions, np.float32).reshape(-1, 3)
    F = H.displacement_fields(P0)
    fds = {n: _reference_fd(sd, 0, wgt, F[n].astype(np.float32), 0.02, 4096) for n in fields}
    ref = max(abs(v) for v in fds.values())
    tol = 0.12 if boundary == 2 else 0.06
    for n, fd in fds.items():
        assert abs(float((g * F[n]).sum()) - fd) < tol * ref, (kind, n, float((g * F[n]).sum()), fd)


def test_texture_gradient_vs_finite_differences_of_the_reference():
    """The texture adjoint against finite differences of the reference renderer: textured Cornell box, a few texels."""
    import copy
    sd, tex = H.textured_cbox(res=16, spp=256, tex_res=4)
    yy, xx = np.mgrid[0:16, 0:16]
    wgt = np.stack([0.5 + xx / 16.0, 0.5 + yy / 16.0, 1.5 - xx / 16.0], axis=2).astype(np.float32)
    h = H.HostSim(sd)
    g = np.mean([h.render_grad(wgt, tex, seed=s_) for s_ in (5, 6, 7, 8)], axis=0).reshape(sd.textures[tex]["data"].shape)
    base = sd.textures[tex]["data"].copy()

    def loss(data):
        s2 = copy.copy(sd)
        s2.textures = [dict(t) for t in sd.textures]
        s2.textures[tex]["data"] = data.astype(np.float32)
        img, _ = H.RefBuild(s2).render(spp=4096)
        return float((img.astype(np.float64) * wgt).sum())

    fd, an = [], []
    for y, x, c in ((1, 1, 0), (2, 1, 1), (1, 2, 2), (2, 2, 0)):
        d = base.copy(); d[y, x, c] += 0.05; lp = loss(d)
        d[y, x, c] -= 0.10; lm = loss(d)
        fd.append((lp - lm) / 0.10); an.append(g[y, x, c])
    fd, an = np.array(fd), np.array(an)
    print("reference FD", fd, "adjoint", an)
    assert np.linalg.norm(fd - an) < 0.05 * np.linalg.norm(fd), (fd, an)
