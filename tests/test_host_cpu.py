"""CPU tests (no GPU): the C-ABI library loads and exports everything include/mtsb200.h declares, the product's
stage bodies (compiled for the host by tests/hostsim) agree with the oracle, the host-side mirror (loaders, scene
description) behaves like the reference's, and the multi-rank sharding logic reduces to the single-rank film (gloo)."""
import ctypes as C
import os
import re
import socket
import tempfile

import numpy as np
import pytest

import helpers as H
from helpers import HostSim, Oracle
from mitsuba2_b200 import _abi as A
from mitsuba2_b200 import distributed as D
from mitsuba2_b200 import mesh as M
from mitsuba2_b200 import xml as X
from mitsuba2_b200.scenedata import SceneData, look_at, perspective_sample_to_camera


# ---- C ABI ---------------------------------------------------------------------------------------------------------
def _declared_functions():
    txt = open(os.path.join(H.ROOT, "include", "mtsb200.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(mtsb_[a-z0-9_]+)\s*\(", txt)))


def test_abi_exports_every_declared_symbol():
    lib = A.lib()
    names = _declared_functions()
    assert len(names) >= 25
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing
    bound = {n for n, _, _ in A.SYMBOLS}
    assert set(names) == bound, set(names) ^ bound       # the ctypes mirror binds exactly the header
    ver = int(re.search(r"#define\s+MTSB_ABI_VERSION\s+(\d+)", open(os.path.join(H.ROOT, "include", "mtsb200.h")).read()).group(1))
    assert lib.mtsb_abi_version() == ver


def test_abi_fails_loudly_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    ctx = C.c_void_p()
    rc = A.lib().mtsb_ctx_create(0, C.byref(ctx))
    assert rc != 0
    assert A.lib().mtsb_last_error()
    with pytest.raises(RuntimeError):
        import mitsuba2_b200 as mi
        mi.Scene(H.sphere_scene())


# ---- product stage bodies on the host vs the oracle ---------------------------------------------------------------
@pytest.mark.parametrize("which", ["sphere", "cbox", "material"])
def test_hostsim_hit_ids_match_oracle(which):
    sd = {"sphere": H.sphere_scene, "cbox": H.cbox_scene, "material": H.material_scene}[which]()
    o, h = Oracle(sd), HostSim(sd)
    om, dm = H.random_rays(20000, 5, radius={"sphere": 1.5, "cbox": 0.9, "material": 3.0}[which])
    if which == "cbox":
        om[:, :3] = om[:, :3] * np.float32(250) + np.float32([278, 273, 280])
    a, b = o.intersect_closest(om, dm), h.intersect_closest(om, dm)
    pa, pb = a.view(np.uint32)[:, 3], b.view(np.uint32)[:, 3]
    assert np.mean(pa != pb) < 5e-3                       # Moeller-Trumbore vs watertight: only edge/tie rays differ
    same = (pa == pb) & (pa != 0xffffffff)
    np.testing.assert_allclose(a[same, 0], b[same, 0], rtol=3e-4, atol=1e-5)   # grazing hits are ill-conditioned in t
    assert np.mean(o.intersect_any(om, dm) != h.intersect_any(om, dm)) < 5e-3
    hn = o.intersect_closest(om[:2000], dm[:2000], naive=True)      # Scene::ray_intersect_naive (kdtree.h:2304-2329)
    assert np.mean(hn.view(np.uint32)[:, 3] != pa[:2000]) < 5e-3


def test_hostsim_bvh_shape():
    n, depth = HostSim(H.sphere_scene()).bvh_info()
    assert n >= 1 and 1 <= depth <= 64


@pytest.mark.parametrize("rfilter", ["box", "gaussian", "tent", "mitchell", "catmullrom", "lanczos"])
def test_hostsim_cbox_film_matches_oracle(rfilter):
    sd = H.cbox_scene(res=16, spp=4, rfilter=rfilter)
    a, b = Oracle(sd).render_xyzaw(seed=3), HostSim(sd).render_xyzaw(seed=3)
    assert np.abs(a - b).max() < 2e-3 * max(1.0, np.abs(a).max())


def test_hostsim_material_film_close_to_oracle():
    sd = H.material_scene(res=(24, 16), spp=4)
    a, b = Oracle(sd).render_xyzaw(seed=1), HostSim(sd).render_xyzaw(seed=1)
    bad = np.abs(a - b).max(axis=2) > 1e-3 * (1 + np.abs(a).max(axis=2))
    assert bad.mean() < 0.02                               # a path through a specular chain may flip at an edge


def test_hostsim_n4_plugins_match_oracle():
    """`conductor` and `plastic` (src/bsdfs/{conductor,plastic}.cpp): BSDF::sample of the product's stage bodies vs the
    oracle on random inputs, and the films of a scene that uses them on the same sample streams."""
    sd = H.n4_scene(res=(24, 16), spp=4)
    orc, hs = Oracle(sd), HostSim(sd)
    rng = np.random.default_rng(2)
    n = 2048
    wi = rng.normal(size=(n, 3)); wi[:, 2] = np.abs(wi[:, 2]) + 0.02; wi /= np.linalg.norm(wi, axis=1, keepdims=True)
    wi[: n // 8, 2] *= -1
    uv, s1, s2 = rng.random((n, 2)), rng.random(n), rng.random((n, 2))
    for bid in range(len(sd.bsdfs)):
        a, b = orc.bsdf_sample(bid, wi, uv, s1, s2), hs.bsdf_sample(bid, wi, uv, s1, s2)
        bad = ~np.isclose(a, b, rtol=2e-3, atol=1e-5).all(axis=1)        # the Newton solve of sample_visible_11 is ill-conditioned
        assert bad.mean() <= 2e-3, (bid, int(bad.sum()))               # at near-normal incidence (see test_gpu_parity)
    # plastic KAT: normal incidence on eta = 1.9 -> F = ((eta - 1) / (eta + 1))^2; specular lobe weight = F / P(specular)
    pl = next(i for i, b in enumerate(sd.bsdfs) if b["type"] == H.A.BSDF_PLASTIC)
    out = orc.bsdf_sample(pl, np.array([[0, 0, 1.0]]), np.zeros((1, 2)), np.array([0.0]), np.zeros((1, 2)))[0]
    F = ((1.9 / 1.000277 - 1) / (1.9 / 1.000277 + 1)) ** 2
    assert out[8] == 1 and np.allclose(out[:3], [0, 0, 1]) and abs(out[3] * out[5] - F) < 1e-5
    a, b = orc.render_xyzaw(seed=1), hs.render_xyzaw(seed=1)
    bad = np.abs(a - b).max(axis=2) > 1e-3 * (1 + np.abs(a).max(axis=2))
    assert bad.mean() < 0.02


def test_sample_range_union_equals_full_render():
    sd = H.cbox_scene(res=12, spp=6)
    h = HostSim(sd)
    full = h.render_xyzaw(seed=9)
    parts = sum(h.render_xyzaw(seed=9, spp=c, spp_begin=b, spp_total=6) for b, c in ((0, 2), (2, 3), (5, 1)))
    np.testing.assert_allclose(parts, full, rtol=1e-5, atol=1e-6)


# ---- host mirror -----------------------------------------------------------------------------------------------------
def test_obj_ply_roundtrip_and_normals():
    m = M.displaced_sphere(6, 12)
    with tempfile.TemporaryDirectory() as d:
        p = os.path.join(d, "m.ply")
        M.write_ply(p, m)
        m2 = M.load_ply(p)
        np.testing.assert_array_equal(m.faces, m2.faces)
        np.testing.assert_allclose(m.positions, m2.positions)
        obj = os.path.join(d, "q.obj")
        open(obj, "w").write("v 0 0 0\nv 1 0 0\nv 1 1 0\nv 0 1 0\nvt 0 0\nvt 1 0\nvt 1 1\nvt 0 1\nf 1/1 2/2 3/3 4/4\n")
        q = M.load_obj(obj)
        assert q.faces.shape == (2, 3) and q.positions.shape[0] == 4 and q.texcoords is not None
    n = M.compute_vertex_normals(m.positions, m.faces)
    r = m.positions / np.linalg.norm(m.positions, axis=1, keepdims=True)
    inner = np.abs(r[:, 1]) < 0.99                         # pole rows only touch degenerate triangles -> (1,0,0) fallback (mesh.cpp:262-270)
    assert np.all(np.sum(n * r, axis=1)[inner] > 0.8)      # outward


def test_perspective_matrix_matches_oracle():
    o = np.zeros(16, np.float32)
    Oracle.lib().orc_perspective_sample_to_camera(45.0, 0.01, 100.0, 16 / 9, H._p(o))
    np.testing.assert_allclose(perspective_sample_to_camera(45.0, 0.01, 100.0, 16 / 9).ravel(), o, rtol=1e-6, atol=1e-7)


def test_xml_subset_loader():
    xml = """<scene version="2.0.0">
      <default name="spp" value="4"/>
      <integrator type="path"><integer name="max_depth" value="5"/></integrator>
      <sensor type="perspective"><float name="fov" value="40"/>
        <transform name="to_world"><lookat origin="0,0,4" target="0,0,0" up="0,1,0"/></transform>
        <sampler type="independent"><integer name="sample_count" value="$spp"/></sampler>
        <film type="hdrfilm"><integer name="width" value="20"/><integer name="height" value="10"/><rfilter type="box"/></film>
      </sensor>
      <bsdf type="diffuse" id="red"><rgb name="reflectance" value="0.8,0.1,0.1"/></bsdf>
      <shape type="rectangle"><ref id="red"/></shape>
      <shape type="rectangle"><transform name="to_world"><rotate x="1" angle="90"/><translate y="2" z="1"/></transform>
        <emitter type="area"><rgb name="radiance" value="5"/></emitter></shape>
    </scene>"""
    sd = X.load_string(xml)
    assert sd.spp == 4 and sd.max_depth == 5
    assert sd.sensors[0]["width"] == 20 and sd.sensors[0]["height"] == 10
    assert len(sd.shapes) == 2 and len(sd.meshes) == 0 and len(sd.emitters) == 1       # rectangles are analytic primitives
    assert Oracle(sd).render_xyzaw(seed=0).shape == (10, 20, 5)
    sd_t = X.load_string(xml, _triangulate_rectangles=True)
    assert len(sd_t.meshes) == 2 and len(sd_t.shapes) == 0
    a, b = Oracle(sd).render_xyzaw(seed=0), Oracle(sd_t).render_xyzaw(seed=0)
    # same picture either way; the light samples differ (Mesh::sample_position picks a triangle first), so only coverage is
    # exact at 4 spp
    assert np.array_equal(a[..., 3:], b[..., 3:]) and abs(a[..., :3].mean() - b[..., :3].mean()) < 0.1 * b[..., :3].mean()
    with pytest.raises(Exception):
        X.load_string("<scene version='2.0.0'><shape type='nosuchshape'/></scene>")


def test_scenedata_npz_roundtrip(tmp_path):
    sd = H.material_scene(res=(8, 8), spp=1)
    p = str(tmp_path / "s.npz")
    sd.save_npz(p)
    sd2 = SceneData.load_npz(p)
    np.testing.assert_array_equal(Oracle(sd).render_xyzaw(seed=2), Oracle(sd2).render_xyzaw(seed=2))


# ---- multi-rank logic over gloo -----------------------------------------------------------------------------------------
def test_sample_share_partition():
    for total in (0, 1, 7, 16, 1024):
        for world in (1, 2, 3, 8):
            shares = [D.sample_share(total, r, world) for r in range(world)]
            assert shares[0][0] == 0 and sum(c for _, c in shares) == total
            assert all(shares[i][0] + shares[i][1] == shares[i + 1][0] for i in range(world - 1))
            assert max(c for _, c in shares) - min(c for _, c in shares) <= 1


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _gloo_worker(rank, world, port, out_dir):
    import torch
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        sd = H.cbox_scene(res=12, spp=5)
        h = HostSim(sd)

        def render(b, c, t, seed):
            return torch.from_numpy(h.render_xyzaw(seed=seed, spp=c, spp_begin=b, spp_total=t).ravel().copy())

        film = D.render_sharded(render, 5, seed=4)
        g = [torch.full((3,), float(rank + 1))]
        D.allreduce_gradients(g)
        np.save(os.path.join(out_dir, f"film{rank}.npy"), film.numpy())
        np.save(os.path.join(out_dir, f"grad{rank}.npy"), g[0].numpy())
    finally:
        dist.destroy_process_group()


def test_world_size_2_sharded_render_equals_single(tmp_path):
    import torch.multiprocessing as mp
    H.build_hostsim()
    mp.spawn(_gloo_worker, args=(2, _free_port(), str(tmp_path)), nprocs=2, join=True)
    full = HostSim(H.cbox_scene(res=12, spp=5)).render_xyzaw(seed=4).ravel()
    f0, f1 = np.load(tmp_path / "film0.npy"), np.load(tmp_path / "film1.npy")
    np.testing.assert_array_equal(f0, f1)
    np.testing.assert_allclose(f0, full, rtol=1e-5, atol=1e-6)
    np.testing.assert_array_equal(np.load(tmp_path / "grad0.npy"), np.full(3, 3.0, np.float32))


# ---- 8(f) N1: twosided / constant / point -------------------------------------------------------------------------
def test_n1_plugins_hostsim_matches_oracle():
    sd = H.n1_scene()
    a, b = Oracle(sd).render_xyzaw(seed=2), HostSim(sd).render_xyzaw(seed=2)
    assert a[..., :3].max() > 0
    bad = np.abs(a - b).max(axis=2) > 1e-3 * (1 + np.abs(a).max(axis=2))
    assert bad.mean() < 0.02


def test_twosided_mirrors_the_nested_brdf():
    """twosided.cpp:98-200: f(wi, wo) on the back side equals the nested f(-wi.z, -wo.z)."""
    sd = SceneData()
    q = M.quad([-1, 0, -1], [-1, 0, 1], [1, 0, 1], [1, 0, -1])
    q.bsdf = sd.make_two_sided(sd.add_diffuse(0.5))
    sd.add_mesh(q)
    sd.add_perspective(look_at([0, 3, 0], [0, 0, 0], [0, 0, 1]), 45.0, 0.01, 100.0, 4, 4, "box")
    o = Oracle(sd)
    wi = np.array([[0.3, 0.1, 0.9], [0.3, 0.1, -0.9]], np.float32)
    wo = np.array([[-0.2, 0.4, 0.8], [-0.2, 0.4, -0.8]], np.float32)
    ev, pdf = o.bsdf_eval_pdf(0, wi, np.zeros((2, 2), np.float32), wo)
    assert pdf[0] > 0 and pdf[0] == pdf[1] and np.array_equal(ev[0], ev[1])
    ev, pdf = o.bsdf_eval_pdf(0, wi, np.zeros((2, 2), np.float32), wo[::-1].copy())     # opposite sides: no transmission
    assert np.all(pdf == 0) and np.all(ev == 0)
    bs = o.bsdf_sample(0, wi, np.zeros((2, 2), np.float32), np.array([0.3, 0.3], np.float32), np.array([[0.2, 0.7]] * 2, np.float32))
    assert bs[0, 2] > 0 and bs[1, 2] == -bs[0, 2] and np.array_equal(bs[0, :2], bs[1, :2])
    with pytest.raises(ValueError):
        sd.make_two_sided(sd.add_dielectric())


def test_constant_and_point_emitter_sampling():
    sd = H.n1_scene()
    o = Oracle(sd)
    ref = np.zeros((64, 3), np.float32)
    s2 = np.random.default_rng(0).random((64, 2)).astype(np.float32)
    out = o.sample_emitter_direction(ref, s2)        # d[3], dist, pdf, value[3], n[3]
    point = s2[:, 0] < 0.5                           # two emitters: sample reuse picks index floor(2 * s.x)
    np.testing.assert_allclose(out[point, 4], 0.5)                                  # delta light: pdf 1 x emitter pdf
    np.testing.assert_allclose(out[~point, 4], 0.5 / (4 * np.pi), rtol=1e-6)        # constant.cpp:98-103
    d = np.array([-1.0, 1.5, -1.0]); dist = np.linalg.norm(d)
    np.testing.assert_allclose(out[point, :3], np.tile(d / dist, (point.sum(), 1)), atol=1e-6)
    np.testing.assert_allclose(out[point, 5:8], np.tile(np.array([8.0, 7.0, 6.0]) / dist**2 / 0.5, (point.sum(), 1)), rtol=1e-5)
    np.testing.assert_allclose(out[~point, 5:8], np.tile(np.array([0.3, 0.4, 0.6]) * 4 * np.pi / 0.5, ((~point).sum(), 1)), rtol=1e-5)
    h = HostSim(sd)
    np.testing.assert_allclose(h.sample_emitter_direction(ref, s2), out, rtol=1e-5, atol=1e-6)


# ---- the ABI from C and C++ ---------------------------------------------------------------------------------------------
def test_header_is_valid_c99_and_cpp_host_links(tmp_path):
    import subprocess
    c = tmp_path / "abi.c"
    c.write_text('#include "mtsb200.h"\nint main(void) { return mtsb_abi_version() == MTSB_ABI_VERSION ? 0 : 1; }\n')
    inc = os.path.join(H.ROOT, "include")
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-I" + inc, "-c", str(c), "-o", str(tmp_path / "abi.o")])
    A.lib()                                      # library must exist
    libdir = os.path.join(H.ROOT, "mitsuba2_b200")
    exe = str(tmp_path / "host_render")
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-I" + inc, os.path.join(H.ROOT, "examples", "host_render.cpp"),
                           "-L" + libdir, "-lmtsb200", "-Wl,-rpath," + libdir, "-o", exe])
    import torch
    if not torch.cuda.is_available():            # without a GPU the host must fail loudly, not fall back
        r = subprocess.run([exe, str(tmp_path / "o.pfm"), "2"], capture_output=True, text=True)
        assert r.returncode != 0 and "failed" in r.stderr


def test_crop_window_equals_the_window_of_the_full_film():
    """`hdrfilm` crop (src/librender/film.cpp:16-44; perspective.cpp:106-133 folds it into camera_to_sample): the camera
    rays of a crop-window sensor are those of the same pixels of the full film."""
    from mitsuba2_b200.scenedata import SceneData, look_at
    full, crop = SceneData(), SceneData()
    tw = look_at([0, 0, -3], [0, 0.2, 0], [0, 1, 0])
    full.add_perspective(tw, 50.0, 0.1, 100.0, 40, 30, "box")
    crop.add_perspective(tw, 50.0, 0.1, 100.0, 40, 30, "box", crop=(8, 5, 16, 12))
    assert crop.sensors[0]["width"] == 16 and crop.sensors[0]["height"] == 12
    for sd in (full, crop):
        sd.add_mesh(H.Mesh(*H.grid_mesh(1, 1, 1.0, [0, 0, 0], [1, 0, 0], [0, 1, 0]), bsdf=sd.add_diffuse(0.5)))
    px = np.random.default_rng(0).random((64, 2))
    of, df = Oracle(full).sensor_sample_ray((px * [16, 12] + [8, 5]) / [40, 30])
    oc, dc = Oracle(crop).sensor_sample_ray(px)
    assert np.allclose(of, oc, atol=1e-5) and np.allclose(df[:, :3], dc[:, :3], atol=1e-5)
    from mitsuba2_b200 import xml
    sd = xml.load_string("""<scene version="2.0.0"><sensor type="perspective"><film type="hdrfilm">
        <integer name="width" value="40"/><integer name="height" value="30"/><integer name="crop_offset_x" value="8"/>
        <integer name="crop_offset_y" value="5"/><integer name="crop_width" value="16"/><integer name="crop_height" value="12"/>
        <rfilter type="mitchell"><float name="B" value="0.2"/></rfilter></film></sensor>
        <shape type="rectangle"/></scene>""")
    s = sd.sensors[0]
    assert (s["width"], s["height"]) == (16, 12) and abs(s["stddev"] - 0.2) < 1e-6 and abs(s["filter_c"] - 1 / 3) < 1e-6


def _write_serialized(path, meshes, version=4, double=False):
    """Writer for the format of src/shapes/serialized.cpp:56-135 (test helper): zlib streams + offset dictionary."""
    import struct
    import zlib
    blob, offsets = b"", []
    for name, pos, nrm, tex, faces in meshes:
        flags = (0x2000 if double else 0x1000) | (0x0001 if nrm is not None else 0) | (0x0002 if tex is not None else 0)
        dt = "<f8" if double else "<f4"
        body = struct.pack("<I", flags) + (name.encode() + b"\0" if version == 4 else b"") + struct.pack("<QQ", len(pos), len(faces))
        body += np.asarray(pos, dt).tobytes()
        if nrm is not None:
            body += np.asarray(nrm, dt).tobytes()
        if tex is not None:
            body += np.asarray(tex, dt).tobytes()
        body += np.asarray(faces, "<u4").tobytes()
        offsets.append(len(blob))
        blob += struct.pack("<hh", 0x041C, version) + zlib.compress(body)
    tail = b"".join(struct.pack("<Q" if version == 4 else "<I", o) for o in offsets) + struct.pack("<I", len(offsets))
    open(path, "wb").write(blob + tail)


@pytest.mark.parametrize("version,double", [(4, False), (3, False), (4, True)])
def test_serialized_loader(tmp_path, version, double):
    """`serialized` shape plugin (src/shapes/serialized.cpp:177-330): versions 3 / 4, shape_index, precision flags, normals
    computed when the file has none; plus the reference's own fixtures and expected values (test_mesh.py:131-170) when
    the reference tree is present."""
    from mitsuba2_b200 import mesh as M
    rng = np.random.default_rng(version)
    a = ("first", rng.random((5, 3)), None, rng.random((5, 2)), [[0, 1, 2], [2, 3, 4]])
    nb = rng.normal(size=(4, 3)); nb /= np.linalg.norm(nb, axis=1, keepdims=True)
    b = ("second", rng.random((4, 3)), nb, None, [[0, 1, 2], [0, 2, 3]])
    path = str(tmp_path / "two.serialized")
    _write_serialized(path, [a, b], version, double)
    m0, m1 = M.load_serialized(path), M.load_serialized(path, shape_index=1)
    assert np.allclose(m0.positions, a[1], atol=1e-6) and np.allclose(m0.texcoords, a[3], atol=1e-6) and m0.faces.tolist() == a[4]
    assert np.allclose(m0.normals, M.compute_vertex_normals(m0.positions, m0.faces))
    assert np.allclose(m1.positions, b[1], atol=1e-6) and np.allclose(m1.normals, nb, atol=1e-6) and m1.texcoords is None
    if version == 4:
        assert m1.name == "second"
    assert M.load_serialized(path, face_normals=True).normals is None
    with pytest.raises(ValueError):
        M.load_serialized(path, shape_index=2)
    ref = "/root/reference/resources/data/tests/serialized/rectangle_normals_uv.serialized"
    if os.path.exists(ref):
        m = M.load_serialized(ref)
        assert np.allclose(m.positions[[0, 2, 3]], [[-2.85, 0, -7.6], [2.85, 0, 0.6], [2.85, 0, -7.6]], atol=1e-3)
        assert np.allclose(m.texcoords[[0, 2, 3]], [[0.950589, 1 - 0.988416], [0.025105, 1 - 0.689127], [0.950589, 1 - 0.689127]], atol=1e-3)
        assert np.allclose(m.normals, [[0, 1, 0]] * 4)


def test_sample_stream_index_beyond_32_bits():
    """pixels * spp_total > 2^32 (e.g. 1080p x 2304 spp over 8 ranks): the sample-stream index is 64-bit, its high half is
    folded into the sampler seed; streams below 2^32 are unchanged, the ones above differ from them and agree with the oracle."""
    sd = H.cbox_scene(res=4, spp=2)
    big = 1 << 29                                       # 16 pixels x 2^29 = 2^33 streams
    o, h = Oracle(sd), HostSim(sd)
    a = o.render_xyzaw(seed=5, spp=2, spp_begin=big - 2, spp_total=big)
    b = h.render_xyzaw(seed=5, spp=2, spp_begin=big - 2, spp_total=big)
    assert np.abs(a - b).max() < 2e-3 * max(1.0, np.abs(a).max())
    # pixel 8 of this job (index 2^32 + 2^29 - 2, seed + 1) must not reuse the streams of pixel 0 (index 2^29 - 2, seed)
    assert not np.allclose(a.reshape(-1, 5)[8, :3], a.reshape(-1, 5)[0, :3])


# ---- analytic shapes as BVH leaf primitives (N4) -------------------------------------------------------------------
def test_hostsim_analytic_shapes_match_oracle():
    """sphere / disk / cylinder / rectangle leaves in the product's BVH8 (csrc/shapes.h, bvh.h) vs the oracle's restatement
    (pinned against the reference plugins in test_oracle_vs_reference_build.py): hit ids exact, any-hit, film, NEE."""
    sd = H.shapes_scene(emitters="disk sphere cylinder rectangle")
    o, h = Oracle(sd), HostSim(sd)
    om, dm = H.random_rays(20000, 3, radius=4.0)
    a, b = o.intersect_closest(om, dm), h.intersect_closest(om, dm)
    pa, pb = a.view(np.uint32)[:, 3], b.view(np.uint32)[:, 3]
    assert set(np.unique(pa)) >= set(range(11))
    tri = (pa < 2) | (pb < 2)                                  # the floor's two triangles: watertight vs Moeller-Trumbore edge rays
    assert np.array_equal(pa[~tri], pb[~tri]) and np.mean(pa != pb) < 1e-3
    same = (pa == pb) & (pa != 0xffffffff)
    np.testing.assert_allclose(a[same, :3], b[same, :3], rtol=1e-5, atol=1e-5)       # t and the cached object-space (x, y)
    dm2 = dm.copy()
    dm2[:, 3] = np.random.default_rng(2).random(len(dm2)).astype(np.float32) * 6
    assert np.mean(o.intersect_any(om, dm2) != h.intersect_any(om, dm2)) < 1e-3
    ra, rb = o.render_xyzaw(spp=8, seed=1), h.render_xyzaw(spp=8, seed=1)
    assert np.abs(ra - rb).sum() / np.abs(ra).sum() < 1e-5
    rng = np.random.default_rng(1)
    p, s2 = rng.uniform(-2, 2, (4000, 3)).astype(np.float32), rng.random((4000, 2)).astype(np.float32)
    ea, eb = o.sample_emitter_direction(p, s2), h.sample_emitter_direction(p, s2)
    assert np.allclose(ea, eb, rtol=2e-3, atol=1e-5, equal_nan=True)


def test_all_analytic_scene_and_npz_roundtrip(tmp_path):
    """No triangle at all; SceneData with shapes survives save_npz / load_npz."""
    sd = H.shapes_scene(emitters="sphere", tri_floor=False)
    assert len(sd.meshes) == 0 and len(sd.shapes) == 7
    sd.save_npz(str(tmp_path / "s.npz"))
    from mitsuba2_b200.scenedata import SceneData
    sd2 = SceneData.load_npz(str(tmp_path / "s.npz"))
    a, b = HostSim(sd).render_xyzaw(spp=4, seed=2), HostSim(sd2).render_xyzaw(spp=4, seed=2)
    assert np.array_equal(a, b)
    o = Oracle(sd).render_xyzaw(spp=4, seed=2)
    assert np.abs(a - o).sum() / np.abs(o).sum() < 1e-5


def test_xml_loader_analytic_shapes():
    """<shape type="sphere|disk|cylinder|rectangle"> load as analytic primitives with the plugins' parameters folded into
    to_world (sphere.cpp:100-110, cylinder.cpp:84-97); `_triangulate_rectangles` keeps the two-triangle rectangle."""
    from mitsuba2_b200 import xml
    src = '''<scene version="2.0.0">
      <shape type="sphere"><point name="center" x="1" y="2" z="3"/><float name="radius" value="0.5"/>
        <emitter type="area"><rgb name="radiance" value="2"/></emitter></shape>
      <shape type="cylinder"><point name="p0" x="0" y="0" z="0"/><point name="p1" x="0" y="1" z="0"/><float name="radius" value="0.25"/></shape>
      <shape type="disk"><boolean name="flip_normals" value="true"/></shape>
      <shape type="rectangle"><transform name="to_world"><translate x="0" y="0" z="5"/></transform></shape>
      <sensor type="perspective"><film type="hdrfilm"><integer name="width" value="8"/><integer name="height" value="8"/></film></sensor>
    </scene>'''
    sd = xml.load_string(src)
    assert [s["type"] for s in sd.shapes] == [1, 3, 2, 4] and len(sd.meshes) == 0
    assert np.allclose(sd.shapes[0]["to_world"], [[0.5, 0, 0, 1], [0, 0.5, 0, 2], [0, 0, 0.5, 3], [0, 0, 0, 1]])
    assert sd.shapes[0]["emitter"] == 0 and sd.emitters[0]["shape_id"] == 0
    cyl = sd.shapes[1]["to_world"]
    assert np.allclose(cyl[:3, 2], [0, 1, 0]) and np.allclose(np.linalg.norm(cyl[:3, :2], axis=0), 0.25)
    assert sd.shapes[2]["flip_normals"] and np.allclose(sd.shapes[3]["to_world"][:3, 3], [0, 0, 5])
    om = np.array([[1, 2, 0, 0], [0, 0.5, -3, 0], [0.6, -0.5, -3, 0], [0.3, 0.3, 2, 0]], np.float32)
    dm = np.array([[0, 0, 1, np.inf]] * 4, np.float32)
    hits = HostSim(sd).intersect_closest(om, dm)
    assert np.array_equal(hits.view(np.uint32)[:, 3], [0, 1, 2, 3]) and np.allclose(hits[:, 0], [2.5, 2.75, 3.0, 3.0], atol=1e-5)
    sd_t = xml.load_string(src, _triangulate_rectangles=True)
    assert len(sd_t.meshes) == 1 and len(sd_t.shapes) == 3


def test_film_drops_invalid_samples():
    """ImageBlock::put rejects samples with a NaN / Inf / negative channel, weight included (imageblock.cpp:84-108)."""
    sd = H.cbox_scene(8, 1, "gaussian")
    pos = np.array([[3.5, 3.5]] * 4, np.float32)
    val = np.array([[1, 1, 1, 1, 1], [np.nan, 1, 1, 1, 1], [1, np.inf, 1, 1, 1], [1, 1, -0.5, 1, 1]], np.float32)
    f = Oracle(sd).film_put(pos, val)
    g = Oracle(sd).film_put(pos[:1], val[:1])
    assert np.isfinite(f).all() and np.array_equal(f, g)


def test_hostsim_bvh_refit_equals_rebuild():
    """bvh.h refit_tri_body / refit_node_body: after the vertices moved (per-vertex noise + a scale + a shift of the whole
    mesh, far more than an optimiser step), the refitted BVH answers every ray like a BVH built from scratch on the new
    positions -- same primitive, same t (the triangle test does not depend on the tree; ties aside) -- and like the oracle."""
    import copy
    sd = H.sphere_scene(24, 48)
    h = HostSim(sd)
    P0 = np.asarray(sd.meshes[0].positions, np.float32).reshape(-1, 3)
    rng = np.random.default_rng(3)
    P1 = (1.15 * P0 + 0.03 * rng.normal(size=P0.shape) + np.float32([0.2, -0.1, 0.05])).astype(np.float32)
    h.refit_vertices(0, P1)
    sd2 = copy.deepcopy(sd)
    sd2.meshes[0].positions = P1
    h2, o2 = HostSim(sd2), Oracle(sd2)
    om, dm = H.random_rays(20000, 7, radius=2.0)
    a, b, c = h.intersect_closest(om, dm), h2.intersect_closest(om, dm), o2.intersect_closest(om, dm)
    pa, pb, pc = (x.view(np.uint32)[:, 3] for x in (a, b, c))
    assert (pa != 0xffffffff).mean() > 0.3
    assert np.mean(pa != pb) < 1e-4                       # equal-t ties can resolve differently in another visit order
    same = pa == pb
    np.testing.assert_array_equal(a[same, 0], b[same, 0])
    assert np.mean(pa != pc) < 5e-3
    np.testing.assert_array_equal(h.intersect_any(om, dm), h2.intersect_any(om, dm))
