"""End-to-end pinning against the REFERENCE'S OWN stored renders (resources/data/tests/scenes/bsdf_spheres/refs, 32k-spp
images + variance images written by src/librender/tests/test_renders.py): the oracle and the product's stage bodies (host
simulator) render the reference's test scene from its XML and must pass the reference's z-test (test_renders.py:60-126:
per-pixel two-sided z-test on XYZ against the stored variance, Sidak-corrected significance 0.01, >= 99.75 % of the pixels).

Runs where the reference tree is present (this container); the scene needs the reference's 3 MB texture and mesh files, which
are not copied into the repo.  The GPU path is tied to the oracle by the same-stream tests of tests/test_gpu_parity.py."""
import math
import os

import numpy as np
import pytest

import helpers as H

SCENES = "/root/reference/resources/data/tests/scenes/bsdf_spheres"
pytestmark = pytest.mark.skipif(not os.path.isdir(SCENES), reason="reference tree not present")

RGB2XYZ = np.array([[0.412453, 0.357580, 0.180423], [0.212671, 0.715160, 0.072169], [0.019334, 0.119193, 0.950227]], np.float64)


def _read_exr_xyz(path):
    os.environ.setdefault("OPENCV_IO_ENABLE_OPENEXR", "1")
    import cv2
    im = cv2.imread(path, cv2.IMREAD_UNCHANGED)
    assert im is not None, path
    return im[:, :, :3][:, :, ::-1].astype(np.float64) @ RGB2XYZ.T       # read_rgb_bmp_to_xyz (test_renders.py:46-48)


def _z_test_pass_fraction(xyz_mean, spp, ref, ref_var, significance=0.01):
    ref_var = np.maximum(ref_var, 1e-4)                                        # test_renders.py:60-78
    z = np.abs(xyz_mean - ref) * np.sqrt(spp / ref_var)
    p = 2.0 * (1.0 - 0.5 * (1.0 + np.vectorize(math.erf)(z / math.sqrt(2.0))))
    npix = ref.shape[0] * ref.shape[1]
    alpha = 1.0 - (1.0 - significance) ** (1.0 / npix)
    return np.count_nonzero(p > alpha) / 3 / npix


@pytest.mark.parametrize("renderer,mode,scene", [("oracle", "rgb", "test_textured_diffuse"), ("hostsim", "rgb", "test_textured_diffuse"),
                                                 ("oracle", "spectral", "test_textured_diffuse"), ("hostsim", "spectral", "test_textured_diffuse"),
                                                 ("oracle", "rgb", "test_roughdielectric"), ("hostsim", "rgb", "test_roughdielectric"),
                                                 ("oracle", "rgb", "test_roughplastic_linear"), ("hostsim", "rgb", "test_roughplastic_nonlinear")])
def test_single_sphere_scenes_pass_the_reference_z_test(renderer, mode, scene):
    """test_textured_diffuse.xml: OBJ sphere with a bitmap (JPEG) diffuse texture, area light + envmap + point light,
    `moment(path, max_depth 2)`, box filter, 48 x 48 -- every plugin on the hot path except the rough conductor.
    test_roughdielectric.xml: the same with a Beckmann (alpha 0.02) rough dielectric sphere."""
    from mitsuba2_b200 import xml
    spp = int(2e6) // (48 * 48)                                                # test_renders.py:102-105
    sd = xml.load_file(os.path.join(SCENES, scene + ".xml"), spp=spp)
    sd.mode = mode                      # `scalar_rgb` / `scalar_spectral` references (get_ref_fname, test_renders.py:31-38)
    if mode == "spectral":
        # needs the reference-resolution (64^3) sRGB -> spectrum table: a 16^3 one fails this z-test.  Building it takes
        # ~4 min on CPU cores (7 s on a GPU) and is cached next to rgb2spec.py, so only run where the cache exists or
        # MTSB_BUILD_SRGB_TABLE=1 asks for the build.
        import mitsuba2_b200
        cache = os.path.join(os.path.dirname(mitsuba2_b200.__file__), "_srgb_coeff_64.npz")
        if not os.path.exists(cache) and os.environ.get("MTSB_BUILD_SRGB_TABLE") != "1":
            pytest.skip("64^3 rgb2spec table not cached (set MTSB_BUILD_SRGB_TABLE=1 to build it)")
    ref = _read_exr_xyz(os.path.join(SCENES, "refs", f"{scene}_ref_{mode}.exr"))
    var = _read_exr_xyz(os.path.join(SCENES, "refs", f"{scene}_ref_{mode}_var.exr"))
    r = H.Oracle(sd) if renderer == "oracle" else H.HostSim(sd)
    film = r.render_xyzaw(spp=spp, seed=7).astype(np.float64)
    xyz = film[:, :, :3] / np.maximum(film[:, :, 4:5], 1e-12)
    frac = _z_test_pass_fraction(xyz, spp, ref, var)
    assert frac >= 0.9975, frac
    assert abs(xyz.mean() - ref.mean()) < 0.01 * ref.mean()


def _supported_pixel_mask(sd, sub=4):
    """Pixels none of whose sub-pixel camera rays hits a shape rendered with a placeholder BSDF."""
    s = sd.sensors[0]
    w, h = s["width"], s["height"]
    bad_meshes = [i for i, m in enumerate(sd.meshes) if m.bsdf in getattr(sd, "placeholder_bsdfs", set())]
    starts = np.cumsum([0] + [len(m.faces) for m in sd.meshes])
    orc = H.Oracle(sd)
    yy, xx, sy, sx = np.meshgrid(np.arange(h), np.arange(w), (np.arange(sub) + 0.5) / sub, (np.arange(sub) + 0.5) / sub, indexing="ij")
    pos = np.stack([(xx + sx).ravel() / w, (yy + sy).ravel() / h], 1)
    om, dm = orc.sensor_sample_ray(pos)
    hits = orc.intersect_closest(om, dm)
    prim = np.ascontiguousarray(hits[:, 3]).view(np.uint32).astype(np.int64)
    mesh = np.searchsorted(starts, prim, side="right") - 1
    bad = np.isfinite(hits[:, 0]) & np.isin(mesh, bad_meshes)
    bad = bad.reshape(h, w, sub * sub).any(axis=2)
    grown = bad.copy()                       # one pixel of margin: a sliver of a placeholder shape between the sub-pixel rays
    grown[1:, :] |= bad[:-1, :]; grown[:-1, :] |= bad[1:, :]; grown[:, 1:] |= bad[:, :-1]; grown[:, :-1] |= bad[:, 1:]
    return ~grown


@pytest.mark.parametrize("scene", ["test_arealight", "test_pointlight", "test_constant", "test_envmap"])
def test_sphere_grid_scenes_pass_the_reference_z_test_on_supported_pixels(scene):
    """The reference's sphere-grid scenes: 13 shapes with one BSDF each (diffuse, bitmap-textured diffuse, `diamond` dielectric,
    plastic, two rough dielectrics, two rough plastics, conductor, rough conductor, two leaf-masked rectangles) under one
    emitter, `max_depth` 2, compared with the reference's stored render by its own z-test on the whole image.  (The harness
    can still leave out pixels that show a plugin the loader does not know: `_placeholder_unsupported`.)"""
    from mitsuba2_b200 import xml
    spp = int(2e6) // (48 * 48)
    sd = xml.load_file(os.path.join(SCENES, scene + ".xml"), spp=spp, _placeholder_unsupported=True)
    ref = _read_exr_xyz(os.path.join(SCENES, "refs", scene + "_ref_rgb.exr"))
    var = _read_exr_xyz(os.path.join(SCENES, "refs", scene + "_ref_rgb_var.exr"))
    mask = _supported_pixel_mask(sd)
    assert mask.all() and not getattr(sd, "placeholder_bsdfs", set())      # every plugin of these scenes is supported now
    film = H.Oracle(sd).render_xyzaw(spp=spp, seed=11).astype(np.float64)
    xyz = film[:, :, :3] / np.maximum(film[:, :, 4:5], 1e-12)
    ref_var = np.maximum(var, 1e-4)
    z = np.abs(xyz - ref) * np.sqrt(spp / ref_var)
    p = 2.0 * (1.0 - 0.5 * (1.0 + np.vectorize(math.erf)(z / math.sqrt(2.0))))
    alpha = 1.0 - (1.0 - 0.01) ** (1.0 / (48 * 48))
    ok = (p > alpha).all(axis=2)
    frac = ok[mask].mean()
    assert frac >= 0.9975, (scene, frac, int(mask.sum()))
