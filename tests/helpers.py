"""Test helpers: ctypes bindings of the CPU oracle (oracle/) and of the host simulator (tests/hostsim), plus
small scene builders.  The oracle is only ever loaded from tests/, __graft_entry__.smoke() and bench.py."""
import ctypes as C
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from mitsuba2_b200 import _abi as A                      # noqa: E402
from mitsuba2_b200.scenedata import SceneData, Mesh, look_at   # noqa: E402
from mitsuba2_b200 import mesh as M                       # noqa: E402

ORACLE_SO = os.path.join(ROOT, "oracle", "libmts_oracle.so")
HOSTSIM_SO = os.path.join(ROOT, "tests", "hostsim", "libhostsim.so")
vp = C.c_void_p


def _newer(src_list, target):
    return not os.path.exists(target) or any(os.path.getmtime(s) > os.path.getmtime(target) for s in src_list)


def build_oracle():
    src = [os.path.join(ROOT, "oracle", f) for f in ("mts_oracle.c", "mts_oracle.h")] + [os.path.join(ROOT, "include", "mtsb200.h")]
    if _newer(src, ORACLE_SO):
        subprocess.check_call(["gcc", "-O2", "-std=c99", "-fopenmp", "-fPIC", "-shared", "-o", ORACLE_SO, src[0], "-lm"])
    return ORACLE_SO


def build_hostsim():
    csrc = os.path.join(ROOT, "mitsuba2_b200", "csrc")
    src = [os.path.join(ROOT, "tests", "hostsim", "hostsim.cu")] + [os.path.join(csrc, f) for f in os.listdir(csrc) if f.endswith(".h")]
    src.append(os.path.join(ROOT, "include", "mtsb200.h"))
    if _newer(src, HOSTSIM_SO):
        subprocess.check_call(["nvcc", "-O2", "-std=c++17", "-Wno-deprecated-gpu-targets", "-diag-suppress", "20013", "-Xcompiler", "-fPIC",
                               "-shared", "-o", HOSTSIM_SO, src[0]])
    return HOSTSIM_SO


def _p(a):
    return vp(a.ctypes.data)


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


class Oracle:
    """CPU oracle scene (oracle/mts_oracle.c)."""
    _lib = None

    @classmethod
    def lib(cls):
        if cls._lib is None:
            l = C.CDLL(build_oracle())
            l.orc_scene_create.restype = vp
            l.orc_scene_create.argtypes = [C.POINTER(A.SceneDesc)]
            l.orc_tea32.restype = C.c_uint32
            l.orc_tea32.argtypes = [C.c_uint32, C.c_uint32, C.c_int]
            l.orc_tea64.restype = C.c_uint64
            l.orc_tea64.argtypes = [C.c_uint32, C.c_uint32, C.c_int]
            for name in ("orc_fresnel_conductor", "orc_microfacet_eval", "orc_microfacet_pdf", "orc_microfacet_smith_g1", "orc_erfinv", "orc_rfilter_eval"):
                getattr(l, name).restype = C.c_float
            l.orc_fresnel_conductor.argtypes = [C.c_float] * 3
            l.orc_fresnel.argtypes = [C.c_float, C.c_float, vp]
            l.orc_erfinv.argtypes = [C.c_float]
            l.orc_rfilter_eval.argtypes = [C.c_int, C.c_float, C.c_float, C.c_float]
            l.orc_rfilter_eval2.restype = C.c_float
            l.orc_rfilter_eval2.argtypes = [C.c_int, C.c_float, C.c_float, C.c_float, C.c_float]
            l.orc_microfacet_eval.argtypes = [C.c_int, C.c_float, C.c_float, vp]
            l.orc_microfacet_pdf.argtypes = [C.c_int, C.c_float, C.c_float, C.c_int, vp, vp]
            l.orc_microfacet_smith_g1.argtypes = [C.c_int, C.c_float, C.c_float, vp, vp]
            l.orc_microfacet_sample.argtypes = [C.c_int, C.c_float, C.c_float, C.c_int, vp, C.c_float, C.c_float, vp]
            l.orc_square_to_uniform_disk_concentric.argtypes = [C.c_float, C.c_float, vp]
            l.orc_square_to_cosine_hemisphere.argtypes = [C.c_float, C.c_float, vp]
            l.orc_square_to_uniform_triangle.argtypes = [C.c_float, C.c_float, vp]
            l.orc_coordinate_system.argtypes = [vp, vp, vp]
            l.orc_perspective_sample_to_camera.argtypes = [C.c_float] * 4 + [vp]
            l.orc_pcg32_floats.argtypes = [C.c_uint64, vp, vp, C.c_uint32, vp]
            l.orc_sampler_seed.argtypes = [C.c_uint64, C.c_uint64, vp, vp]
            l.orc_scene_destroy.argtypes = [vp]
            l.orc_scene_update_vertices.argtypes = [vp, C.c_uint32, vp]
            l.orc_scene_update_texture.argtypes = [vp, C.c_uint32, vp]
            l.orc_intersect_closest.argtypes = [vp, C.c_uint64, vp, vp, vp, C.c_int]
            l.orc_intersect_any.argtypes = [vp, C.c_uint64, vp, vp, vp, C.c_int]
            l.orc_intersect_naive.argtypes = [vp, C.c_uint64, vp, vp, vp, C.c_int]
            l.orc_surface_interaction.argtypes = [vp, C.c_uint64, vp, vp, vp, C.c_int]
            l.orc_sensor_sample_ray.argtypes = [vp, C.c_uint32, C.c_uint64, vp, vp, vp]
            l.orc_bsdf_sample.argtypes = [vp, C.c_uint32, C.c_uint64, vp, vp, vp, vp, vp]
            l.orc_bsdf_eval_pdf.argtypes = [vp, C.c_uint32, C.c_uint64, vp, vp, vp, vp, vp]
            l.orc_sample_emitter_direction.argtypes = [vp, C.c_uint64, vp, vp, vp]
            l.orc_film_put.argtypes = [vp, C.c_uint32, C.c_uint64, vp, vp, vp]
            l.orc_film_develop.argtypes = [vp, C.c_uint32, vp, vp]
            l.orc_render.argtypes = [vp, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint64, vp, C.c_int, vp]
            l.orc_trace_sample.argtypes = [vp, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint64, vp]
            cls._lib = l
        return cls._lib

    def __init__(self, data: SceneData, threads=None):
        self.data = data
        self.threads = threads or os.cpu_count() or 1
        desc, keep = data.to_desc()
        self.h = vp(self.lib().orc_scene_create(C.byref(desc)))
        del keep

    def __del__(self):
        try:
            if self.h:
                self.lib().orc_scene_destroy(self.h)
                self.h = None
        except Exception:
            pass

    def intersect_closest(self, om, dm, naive=False):
        om, dm = _f32(om), _f32(dm)
        hits = np.empty((om.shape[0], 4), np.float32)
        fn = self.lib().orc_intersect_naive if naive else self.lib().orc_intersect_closest
        fn(self.h, om.shape[0], _p(om), _p(dm), _p(hits), self.threads)
        return hits

    def surface_interaction(self, om, dm):
        """-> (n,16): t, uv, prim bits, p, n, sh_n, wi (the layout of RefBuild.intersect)"""
        om, dm = _f32(om), _f32(dm)
        out = np.empty((om.shape[0], 16), np.float32)
        self.lib().orc_surface_interaction(self.h, om.shape[0], _p(om), _p(dm), _p(out), self.threads)
        return out

    def intersect_any(self, om, dm):
        om, dm = _f32(om), _f32(dm)
        occ = np.empty(om.shape[0], np.uint32)
        self.lib().orc_intersect_any(self.h, om.shape[0], _p(om), _p(dm), _p(occ), self.threads)
        return occ

    def render_xyzaw(self, sensor=0, spp=None, seed=0, spp_begin=0, spp_total=None, counts=None):
        spp = self.data.spp if spp is None else spp
        spp_total = spp_begin + spp if spp_total is None else spp_total
        s = self.data.sensors[sensor]
        film = np.zeros((s["height"], s["width"], 5), np.float32)
        cnt = np.zeros(4, np.uint64)
        self.lib().orc_render(self.h, sensor, spp_begin, spp, spp_total, seed, _p(film), self.threads, _p(cnt))
        if counts is not None:
            counts[:] = cnt
        return film

    def develop(self, xyzaw, sensor=0):
        xyzaw = _f32(xyzaw)
        rgb = np.empty(xyzaw.shape[:2] + (3,), np.float32)
        self.lib().orc_film_develop(self.h, sensor, _p(xyzaw), _p(rgb))
        return rgb

    def render(self, sensor=0, spp=None, seed=0):
        return self.develop(self.render_xyzaw(sensor, spp, seed), sensor)

    def film_put(self, pos, values, sensor=0):
        pos, values = _f32(pos), _f32(values)
        s = self.data.sensors[sensor]
        film = np.zeros((s["height"], s["width"], 5), np.float32)
        self.lib().orc_film_put(self.h, sensor, pos.shape[0], _p(pos), _p(values), _p(film))
        return film

    def sensor_sample_ray(self, pos, sensor=0):
        pos = _f32(pos)
        o, d = np.empty((pos.shape[0], 4), np.float32), np.empty((pos.shape[0], 4), np.float32)
        self.lib().orc_sensor_sample_ray(self.h, sensor, pos.shape[0], _p(pos), _p(o), _p(d))
        return o, d

    def bsdf_sample(self, bsdf_id, wi, uv, s1, s2):
        wi, uv, s1, s2 = map(_f32, (wi, uv, s1, s2))
        out = np.empty((wi.shape[0], 9), np.float32)
        self.lib().orc_bsdf_sample(self.h, bsdf_id, wi.shape[0], _p(wi), _p(uv), _p(s1), _p(s2), _p(out))
        return out

    def bsdf_eval_pdf(self, bsdf_id, wi, uv, wo):
        wi, uv, wo = map(_f32, (wi, uv, wo))
        ev, pdf = np.empty((wi.shape[0], 3), np.float32), np.empty(wi.shape[0], np.float32)
        self.lib().orc_bsdf_eval_pdf(self.h, bsdf_id, wi.shape[0], _p(wi), _p(uv), _p(wo), _p(ev), _p(pdf))
        return ev, pdf

    def sample_emitter_direction(self, ref_p, s2):
        ref_p, s2 = _f32(ref_p), _f32(s2)
        out = np.empty((ref_p.shape[0], 11), np.float32)
        self.lib().orc_sample_emitter_direction(self.h, ref_p.shape[0], _p(ref_p), _p(s2), _p(out))
        return out

    def update_vertices(self, mesh_id, positions):
        positions = _f32(positions)
        self.lib().orc_scene_update_vertices(self.h, mesh_id, _p(positions))

    def update_texture(self, tex_id, data):
        data = _f32(data)
        self.lib().orc_scene_update_texture(self.h, tex_id, _p(data))


class HostSim:
    """The product's stage bodies executed sequentially on the CPU (tests/hostsim/hostsim.cu)."""
    _lib = None

    @classmethod
    def lib(cls):
        if cls._lib is None:
            l = C.CDLL(build_hostsim())
            l.hs_scene_create.restype = vp
            l.hs_scene_create.argtypes = [C.POINTER(A.SceneDesc)]
            l.hs_scene_destroy.argtypes = [vp]
            l.hs_bvh_info.argtypes = [vp, vp, vp]
            l.hs_refit_vertices.argtypes = [vp, C.c_uint32, vp]
            l.hs_intersect.argtypes = [vp, C.c_uint64, vp, vp, vp, vp, C.c_int]
            l.hs_render.argtypes = [vp, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint64, vp, vp]
            l.hs_render_grad.argtypes = [vp, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint64, vp, C.c_int32, vp, C.c_int32, vp,
                                         C.c_int, C.c_float]
            l.hs_bsdf_sample.argtypes = [vp, C.c_uint32, C.c_uint64, vp, vp, vp, vp, vp]
            l.hs_sample_emitter_direction.argtypes = [vp, C.c_uint64, vp, vp, vp]
            cls._lib = l
        return cls._lib

    def __init__(self, data: SceneData):
        self.data = data
        desc, keep = data.to_desc()
        self.h = vp(self.lib().hs_scene_create(C.byref(desc)))

    def __del__(self):
        try:
            if self.h:
                self.lib().hs_scene_destroy(self.h)
                self.h = None
        except Exception:
            pass

    def bvh_info(self):
        a, b = C.c_uint32(), C.c_uint32()
        self.lib().hs_bvh_info(self.h, C.byref(a), C.byref(b))
        return a.value, b.value

    def refit_vertices(self, mesh_id, positions):
        """bvh.h refit_*_body under the topology of the build (positions only: normals and emitter tables are not refreshed)."""
        positions = np.ascontiguousarray(positions, np.float32)
        self.lib().hs_refit_vertices(self.h, mesh_id, _p(positions))

    def intersect_closest(self, om, dm):
        om, dm = _f32(om), _f32(dm)
        hits = np.empty((om.shape[0], 4), np.float32)
        self.lib().hs_intersect(self.h, om.shape[0], _p(om), _p(dm), _p(hits), None, 0)
        return hits

    def intersect_any(self, om, dm):
        om, dm = _f32(om), _f32(dm)
        occ = np.empty(om.shape[0], np.uint32)
        self.lib().hs_intersect(self.h, om.shape[0], _p(om), _p(dm), None, _p(occ), 1)
        return occ

    def render_xyzaw(self, sensor=0, spp=None, seed=0, spp_begin=0, spp_total=None, counts=None):
        spp = self.data.spp if spp is None else spp
        spp_total = spp_begin + spp if spp_total is None else spp_total
        s = self.data.sensors[sensor]
        film = np.zeros((s["height"], s["width"], 5), np.float32)
        cnt = np.zeros(4, np.uint64)
        self.lib().hs_render(self.h, sensor, spp_begin, spp, spp_total, seed, _p(film), _p(cnt))
        if counts is not None:
            counts[:] = cnt
        return film

    def sample_emitter_direction(self, ref_p, s2):
        ref_p, s2 = _f32(ref_p), _f32(s2)
        out = np.empty((ref_p.shape[0], 11), np.float32)
        self.lib().hs_sample_emitter_direction(self.h, ref_p.shape[0], _p(ref_p), _p(s2), _p(out))
        return out

    def bsdf_sample(self, bsdf_id, wi, uv, s1, s2):
        wi, uv, s1, s2 = _f32(wi), _f32(uv), _f32(s1), _f32(s2)
        out = np.empty((wi.shape[0], 9), np.float32)
        self.lib().hs_bsdf_sample(self.h, bsdf_id, wi.shape[0], _p(wi), _p(uv), _p(s1), _p(s2), _p(out))
        return out

    def render_grad(self, dL_dimage, tex_id, sensor=0, spp=None, seed=0, spp_begin=0, spp_total=None):
        spp = self.data.spp if spp is None else spp
        spp_total = spp_begin + spp if spp_total is None else spp_total
        t = self.data.textures[tex_id]
        g = np.zeros_like(np.ascontiguousarray(t["data"], np.float32))
        dl = _f32(dL_dimage)
        self.lib().hs_render_grad(self.h, sensor, spp_begin, spp, spp_total, seed, _p(dl), tex_id, _p(g), -1, None, 0, 0.0)
        return g

    def render_grad_vertices(self, dL_dimage, mesh_id, sensor=0, spp=None, seed=0, spp_begin=0, spp_total=None, boundary=True,
                             edge_fraction=0.25):
        """d(sum(dL_dimage * image))/d(vertex positions of mesh_id): the same adjoint as mtsb_render_grad, run on the host."""
        spp = self.data.spp if spp is None else spp
        spp_total = spp_begin + spp if spp_total is None else spp_total
        g = np.zeros_like(_f32(self.data.meshes[mesh_id].positions))
        dl = _f32(dL_dimage)
        self.lib().hs_render_grad(self.h, sensor, spp_begin, spp, spp_total, seed, _p(dl), -1, None, mesh_id, _p(g), int(boundary),
                                  float(edge_fraction))
        return g


# ---- scenes ---------------------------------------------------------------------------------------
def random_rays(n, seed, radius=1.5, maxt=np.inf):
    rng = np.random.default_rng(seed)
    o = rng.normal(size=(n, 3))
    o = o / np.linalg.norm(o, axis=1, keepdims=True) * (rng.random((n, 1)) ** (1 / 3)) * radius
    d = rng.normal(size=(n, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    om = np.concatenate([o, np.zeros((n, 1))], axis=1).astype(np.float32)
    dm = np.concatenate([d, np.full((n, 1), maxt)], axis=1).astype(np.float32)
    return om, dm


def sphere_scene(n_theta=32, n_phi=64, res=(64, 48), spp=4):
    sd = SceneData()
    b = sd.add_diffuse([0.5, 0.4, 0.3])
    m = M.displaced_sphere(n_theta, n_phi)
    m.bsdf = b
    sd.add_mesh(m)
    sd.add_perspective(look_at([0, 0, -3.2], [0, 0, 0], [0, 1, 0]), 45.0, 0.01, 100.0, res[0], res[1], "box")
    sd.max_depth, sd.spp = 6, spp
    return sd


def cbox_scene(res=32, spp=8, rfilter="box"):
    """Cornell box from the committed fixture (generated by tests/golden/make_fixtures.py from the reference's XML)."""
    sd = SceneData.load_npz(os.path.join(ROOT, "tests", "golden", "cbox_scene.npz"))
    s = sd.sensors[0]
    from mitsuba2_b200.scenedata import perspective_sample_to_camera
    s["width"] = s["height"] = res
    s["sample_to_camera"] = perspective_sample_to_camera(s["fov_x"], s["near"], s["far"], 1.0).reshape(16)
    kinds = {"box": (A.RFILTER_BOX, 0.5), "tent": (A.RFILTER_TENT, 1.0), "gaussian": (A.RFILTER_GAUSSIAN, 2.0),
             "mitchell": (A.RFILTER_MITCHELL, 2.0), "catmullrom": (A.RFILTER_CATMULLROM, 2.0), "lanczos": (A.RFILTER_LANCZOS, 3.0)}
    s["rfilter"], s["radius"] = kinds[rfilter]
    s["stddev"], s["filter_c"] = (1.0 / 3.0, 1.0 / 3.0) if rfilter == "mitchell" else (0.5, 0.0)
    sd.spp = spp
    return sd


def synthetic_envmap(w=64, h=32, seed=7):
    """Smooth sky gradient + a bright 'sun' blob, deterministic."""
    v, u = np.meshgrid((np.arange(h) + 0.5) / h, (np.arange(w) + 0.5) / w, indexing="ij")
    sky = np.stack([0.3 + 0.4 * (1 - v), 0.4 + 0.4 * (1 - v), 0.6 + 0.4 * (1 - v)], axis=2)
    sun = 40.0 * np.exp(-(((u - 0.3) * 2) ** 2 + (v - 0.25) ** 2) / 0.004)
    rng = np.random.default_rng(seed)
    img = sky * (0.9 + 0.2 * rng.random((h, w, 1))) + sun[:, :, None] * np.array([1.0, 0.9, 0.7])
    return img.astype(np.float32)


def spectral(sd):
    """Switch a scene to the spectral mode (scalar_spectral semantics), using a small resolution-16 sRGB model table so
    that the CPU suite does not have to build the 64^3 one."""
    from mitsuba2_b200 import rgb2spec
    sd.mode, sd.rgb2spec_model = "spectral", rgb2spec.table(16)
    return sd


def material_scene(res=(48, 32), spp=8, envmap=True, area=True, named_conductor=False):
    """diffuse (textured) ground + roughconductor sphere + dielectric sphere + area light + envmap."""
    sd = SceneData()
    rng = np.random.default_rng(3)
    tex = sd.add_texture_bitmap(0.2 + 0.6 * rng.random((16, 16, 3)).astype(np.float32))
    ground = M.quad([-10, -1.2, -10], [-10, -1.2, 10], [10, -1.2, 10], [10, -1.2, -10])
    ground.bsdf = sd.add_diffuse(tex)
    sd.add_mesh(ground)
    s1 = M.displaced_sphere(24, 48)
    s1.bsdf = sd.add_roughconductor(material="Cu", alpha=0.1) if named_conductor else \
        sd.add_roughconductor((0.200438, 0.924033, 1.102212), (3.912949, 2.452848, 2.142188), 0.1)
    sd.add_mesh(s1)
    s2 = M.displaced_sphere(16, 32, radius=0.4, center=(1.6, -0.8, 0), displace=False)
    s2.bsdf = sd.add_dielectric()
    sd.add_mesh(s2)
    if area:
        light = M.quad([-0.5, 3, -0.5], [0.5, 3, -0.5], [0.5, 3, 0.5], [-0.5, 3, 0.5])
        light.bsdf = sd.add_diffuse(0.5)
        sd.add_mesh(light, area_radiance=(18.4, 15.6, 8.0))
    if envmap:
        sd.add_envmap(synthetic_envmap())
    sd.add_perspective(look_at([0, 0, -3.2], [0, 0, 0], [0, 1, 0]), 45.0, 0.01, 100.0, res[0], res[1], "box")
    sd.max_depth, sd.rr_depth, sd.spp = 6, 5, spp
    return sd


def n4_scene(res=(48, 32), spp=8):
    """SURVEY 8f N4 plugins: smooth `conductor` (Au), `plastic` (plain and nonlinear, textured base) and `roughdielectric`
    (GGX visible-normal sampling; anisotropic Beckmann with plain sampling) spheres on a diffuse ground under an area light
    and a constant background."""
    sd = SceneData()
    rng = np.random.default_rng(5)
    ground = M.quad([-10, -1.2, -10], [-10, -1.2, 10], [10, -1.2, 10], [10, -1.2, -10])
    ground.bsdf = sd.add_diffuse([0.5, 0.45, 0.4])
    sd.add_mesh(ground)
    s1 = M.displaced_sphere(16, 32, radius=0.7, center=(-1.2, -0.5, 0), displace=False)
    s1.bsdf = sd.add_conductor(material="Au")
    sd.add_mesh(s1)
    s2 = M.displaced_sphere(16, 32, radius=0.7, center=(0.3, -0.5, 0), displace=False)
    s2.bsdf = sd.add_plastic([0.1, 0.27, 0.36], int_ior=1.9)
    sd.add_mesh(s2)
    s3 = M.displaced_sphere(16, 32, radius=0.5, center=(1.6, -0.7, -0.4), displace=False)
    s3.bsdf = sd.add_plastic(sd.add_texture_bitmap(0.2 + 0.6 * rng.random((8, 8, 3)).astype(np.float32)), specular_reflectance=[0.9, 0.8, 0.7],
                             nonlinear=True)
    sd.add_mesh(s3)
    s4 = M.displaced_sphere(16, 32, radius=0.45, center=(-0.3, -0.75, -1.3), displace=False)
    s4.bsdf = sd.add_roughdielectric(alpha=0.1, distribution="ggx")
    sd.add_mesh(s4)
    s5 = M.displaced_sphere(16, 32, radius=0.4, center=(0.9, -0.8, -1.4), displace=False)
    s5.bsdf = sd.add_roughdielectric(alpha=0.15, alpha_v=0.05, sample_visible=False, specular_transmittance=[0.9, 0.95, 1.0])
    sd.add_mesh(s5)
    s6 = M.displaced_sphere(16, 32, radius=0.4, center=(1.9, -0.8, -1.2), displace=False)
    s6.bsdf = sd.add_roughplastic([0.5, 0.3, 0.2], alpha=0.1, nonlinear=True)
    sd.add_mesh(s6)
    leaf = M.quad([-2.4, -1.1, 0.8], [-1.0, -1.1, 0.8], [-1.0, 0.4, 0.8], [-2.4, 0.4, 0.8])       # `mask` over a two-sided diffuse BRDF
    yy, xx = np.mgrid[0:16, 0:16]
    leaf.bsdf = sd.add_mask(sd.make_two_sided(sd.add_diffuse([0.1, 0.8, 0.1])), (((xx // 2 + yy // 2) % 2) * 0.8 + 0.1).astype(np.float32))
    sd.add_mesh(leaf)
    light = M.quad([-0.5, 3, -0.5], [0.5, 3, -0.5], [0.5, 3, 0.5], [-0.5, 3, 0.5])
    light.bsdf = sd.add_diffuse(0.5)
    sd.add_mesh(light, area_radiance=(18.4, 15.6, 8.0))
    sd.add_constant([0.4, 0.5, 0.7])
    sd.add_perspective(look_at([0, 0, -3.2], [0, 0, 0], [0, 1, 0]), 45.0, 0.01, 100.0, res[0], res[1], "box")
    sd.max_depth, sd.rr_depth, sd.spp = 6, 5, spp
    return sd


# ---------------------------------------------------------------------------------------------------------------
# Reference header-only arithmetic (oracle/_ref/libmts_ref.so, built by `make -C oracle ref` where /root/reference
# exists) and the matching oracle entry points, driven through one table so that the golden-vector generator
# (tests/golden/make_ref_vectors.py) and the tests run exactly the same cases.
# ---------------------------------------------------------------------------------------------------------------
REF_SO = os.path.join(ROOT, "oracle", "_ref", "libmts_ref.so")


def _bind_math(l, prefix):
    f = C.c_float
    sig = {
        "pcg32_floats": (None, [C.c_uint64, vp, vp, C.c_uint32, vp]),
        "tea32": (C.c_uint32, [C.c_uint32, C.c_uint32, C.c_int]),
        "tea64": (C.c_uint64, [C.c_uint32, C.c_uint32, C.c_int]),
        "square_to_uniform_disk_concentric": (None, [f, f, vp]),
        "square_to_cosine_hemisphere": (None, [f, f, vp]),
        "square_to_uniform_triangle": (None, [f, f, vp]),
        "coordinate_system": (None, [vp, vp, vp]),
        "fresnel": (None, [f, f, vp]),
        "fresnel_conductor": (f, [f, f, f]),
        "microfacet_eval": (f, [C.c_int, f, f, vp]),
        "microfacet_pdf": (f, [C.c_int, f, f, C.c_int, vp, vp]),
        "microfacet_smith_g1": (f, [C.c_int, f, f, vp, vp]),
        "microfacet_sample": (None, [C.c_int, f, f, C.c_int, vp, f, f, vp]),
        "erfinv": (f, [f]),
        "perspective_sample_to_camera": (None, [f, f, f, f, vp]),
        "h2d_sample": (None, [vp, C.c_uint32, C.c_uint32, C.c_uint64, vp, vp]),
        "h2d_eval": (None, [vp, C.c_uint32, C.c_uint32, C.c_uint64, vp, vp]),
        "discrete_sample_reuse": (None, [vp, C.c_uint32, C.c_uint64, vp, vp, vp]),
        # spectral variants
        "sample_wavelength": (None, [f, vp, vp]),
        "srgb_model_eval": (f, [vp, f]),
        "regular_eval": (f, [vp, C.c_uint32, f, f, f]),
        "irregular_eval": (f, [vp, vp, C.c_uint32, f]),
        "spectrum_to_xyz": (None, [vp, vp, vp]),
    }
    out = {}
    for name, (res, args) in sig.items():
        fn = getattr(l, prefix + name)
        fn.restype = res
        fn.argtypes = args
        out[name] = fn
    return out


def oracle_math():
    return _bind_math(Oracle.lib(), "orc_")


def ref_math():
    """None when the reference harness has not been built (e.g. on the GPU box)."""
    if not os.path.exists(REF_SO):
        return None
    return _bind_math(C.CDLL(REF_SO), "ref_")


def _hemi(rng, n):
    v = rng.normal(size=(n, 3)).astype(np.float32)
    v /= np.linalg.norm(v, axis=1, keepdims=True)
    v[:, 2] = np.abs(v[:, 2])
    return v


def math_cases(fn, seed=11):
    """Run every unit case through the function table `fn` (oracle_math() or ref_math()); returns {name: ndarray}."""
    rng = np.random.default_rng(seed)
    R = {}
    st = rng.integers(0, 2**63, size=8, dtype=np.uint64)
    sq = rng.integers(0, 2**63, size=8, dtype=np.uint64)
    st[0], sq[0] = 0x853c49e6748fea9b, 0xda3e39cb94b95bdb          # PCG32 defaults (enoki random.h)
    out = np.zeros((8, 16), np.float32)
    fn["pcg32_floats"](8, _p(st), _p(sq), 16, _p(out))
    R["pcg32_floats"] = out
    v = rng.integers(0, 2**32, size=(32, 2), dtype=np.uint64)
    v[:8] = [[1, 1], [1, 2], [1, 3], [1, 4], [1, 5], [2, 1], [3, 1], [4, 1]]      # test_random.py:6-28
    R["tea32"] = np.array([fn["tea32"](int(a), int(b), 4) for a, b in v], np.uint32)
    R["tea64"] = np.array([fn["tea64"](int(a), int(b), 4) for a, b in v], np.uint64)
    s2 = rng.random((64, 2), dtype=np.float32)
    s2[:4] = [[0, 0], [0.5, 0.5], [1, 1], [0.25, 0.75]]
    for name, k in (("square_to_uniform_disk_concentric", 2), ("square_to_cosine_hemisphere", 3), ("square_to_uniform_triangle", 2)):
        o = np.zeros((64, k), np.float32)
        for i in range(64):
            fn[name](float(s2[i, 0]), float(s2[i, 1]), _p(o[i]))
        R[name] = o
    nrm = rng.normal(size=(32, 3)).astype(np.float32)
    nrm /= np.linalg.norm(nrm, axis=1, keepdims=True)
    nrm[:3] = [[0, 0, 1], [0, 0, -1], [1, 0, 0]]
    cs = np.zeros((32, 6), np.float32)
    for i in range(32):
        fn["coordinate_system"](_p(nrm[i]), _p(cs[i, :3]), _p(cs[i, 3:]))
    R["coordinate_system"] = cs
    fr = []
    for eta in (1.5, 1 / 1.5, 1.0, 1.33):
        for c in np.linspace(-1, 1, 41, dtype=np.float32):
            o = np.zeros(4, np.float32)
            fn["fresnel"](float(c), float(np.float32(eta)), _p(o))
            fr.append(o)
    R["fresnel"] = np.array(fr)
    R["fresnel_conductor"] = np.array([fn["fresnel_conductor"](float(c), float(np.float32(e)), float(np.float32(k)))
                                       for e, k in ((0.2, 3.9), (1.5, 0.0), (1.657, 9.22), (0.27, 3.6))
                                       for c in np.linspace(0, 1, 21, dtype=np.float32)], np.float32)
    wi, m = _hemi(rng, 16), _hemi(rng, 16)
    wi[0] = [0, 0, 1]
    ev, pd, g1, sm = [], [], [], []
    for t in (0, 1):
        for au, av in ((0.1, 0.1), (0.1, 0.3), (0.5, 0.5), (0.05, 0.2)):
            au, av = float(np.float32(au)), float(np.float32(av))
            for i in range(16):
                ev.append(fn["microfacet_eval"](t, au, av, _p(m[i])))
                g1.append(fn["microfacet_smith_g1"](t, au, av, _p(wi[i]), _p(m[i])))
                for vis in (0, 1):
                    pd.append(fn["microfacet_pdf"](t, au, av, vis, _p(wi[i]), _p(m[i])))
                    o = np.zeros(4, np.float32)
                    fn["microfacet_sample"](t, au, av, vis, _p(wi[i]), float(s2[i + 8, 0]), float(s2[i + 8, 1]), _p(o))
                    sm.append(o)
    R["microfacet_eval"], R["microfacet_pdf"] = np.array(ev, np.float32), np.array(pd, np.float32)
    R["microfacet_smith_g1"], R["microfacet_sample"] = np.array(g1, np.float32), np.array(sm)
    R["erfinv"] = np.array([fn["erfinv"](float(x)) for x in np.linspace(-0.999, 0.999, 101, dtype=np.float32)], np.float32)
    ps = []
    for fov, near, far, asp in ((45.0, 0.01, 100.0, 16 / 9), (39.3077, 0.001, 100.0, 1.0), (90.0, 1.0, 10.0, 0.5)):
        o = np.zeros(16, np.float32)
        fn["perspective_sample_to_camera"](float(np.float32(fov)), float(np.float32(near)), float(np.float32(far)), float(np.float32(asp)), _p(o))
        ps.append(o)
    R["perspective_sample_to_camera"] = np.array(ps)
    for tag, (w, h) in (("a", (17, 9)), ("b", (64, 32)), ("c", (33, 47))):
        img = (rng.random((h, w), dtype=np.float32) ** 4 * 10 + 0.01).astype(np.float32)
        s = rng.random((256, 2), dtype=np.float32)
        o3 = np.zeros((256, 3), np.float32)
        fn["h2d_sample"](_p(img), w, h, 256, _p(s), _p(o3))
        R["h2d_sample_" + tag] = o3
        o1 = np.zeros(256, np.float32)
        fn["h2d_eval"](_p(img), w, h, 256, _p(s), _p(o1))
        R["h2d_eval_" + tag] = o1
    pmf = rng.random(37, dtype=np.float32)
    pmf[[0, 5, 6, 36]] = 0
    u = rng.random(256, dtype=np.float32)
    idx, uo = np.zeros(256, np.uint32), np.zeros(256, np.float32)
    fn["discrete_sample_reuse"](_p(pmf), 37, 256, _p(u), _p(idx), _p(uo))
    R["discrete_idx"], R["discrete_u"] = idx, uo
    # ---- spectral variants (core/spectrum.h:147-217, 270-316; render/srgb.h:8-24; core/distr_1d.h:378-393, 655-677)
    us = np.concatenate([np.array([0.0, 0.25, 0.5, 0.75, 0.999999], np.float32), rng.random(123, dtype=np.float32)])
    sw = np.zeros((us.size, 8), np.float32)
    for i, u in enumerate(us):
        fn["sample_wavelength"](float(u), _p(sw[i, :4]), _p(sw[i, 4:]))
    R["sample_wavelength"] = sw
    coeffs = (rng.standard_normal((64, 3)) * np.array([1e-4, 0.05, 10.0])).astype(np.float32)
    coeffs[0], coeffs[1], coeffs[2] = (0, 0, np.inf), (0, 0, -np.inf), (0, 0, 0)
    lam = np.concatenate([np.array([360.0, 830.0, 538.0], np.float32), rng.uniform(360, 830, 13).astype(np.float32)])
    R["srgb_model_eval"] = np.array([[fn["srgb_model_eval"](_p(c), float(l)) for l in lam] for c in coeffs], np.float32)
    vals = rng.random(40, dtype=np.float32)
    nodes = np.sort(rng.uniform(300, 900, 40)).astype(np.float32)
    xs = np.concatenate([np.array([360.0, 830.0, 359.9, 830.1], np.float32), nodes[[0, 7, 39]], rng.uniform(280, 920, 200).astype(np.float32)])
    R["regular_eval"] = np.array([fn["regular_eval"](_p(vals), 40, 360.0, 830.0, float(x)) for x in xs], np.float32)
    R["irregular_eval"] = np.array([fn["irregular_eval"](_p(nodes), _p(vals), 40, float(x)) for x in xs], np.float32)
    sx = np.zeros((128, 3), np.float32)
    for i in range(128):
        v4, w4 = rng.random(4, dtype=np.float32), rng.uniform(350, 840, 4).astype(np.float32)
        fn["spectrum_to_xyz"](_p(v4), _p(w4), _p(sx[i]))
    R["spectrum_to_xyz"] = sx
    return R


def textured_cbox(res=16, spp=16, tex_res=6, seed=4, rr_off=True):
    """C4-style scene: Cornell box whose floor reflectance is a small RGB bitmap (the differentiated parameter).
    Returns (scene, texture id).  With `rr_off` Russian roulette never triggers, so the render is a polynomial in the
    texel values and central finite differences are exact up to rounding."""
    sd = cbox_scene(res=res, spp=spp, rfilter="box")
    rng = np.random.default_rng(seed)
    tex = sd.add_texture_bitmap((0.2 + 0.6 * rng.random((tex_res, tex_res, 3))).astype(np.float32))
    floor = next(m for m in sd.meshes if "floor" in m.name)
    p = floor.positions
    lo, hi = p.min(axis=0), p.max(axis=0)
    floor.texcoords = np.ascontiguousarray(((p - lo) / np.maximum(hi - lo, 1e-6))[:, [0, 2]], np.float32)
    floor.bsdf = sd.add_diffuse(tex)
    if rr_off:
        sd.rr_depth = sd.max_depth + 1
    return sd, tex


def z_test_pass_fraction(mean, reference, var_per_sample, n_mean, n_ref, significance=0.01):
    """Per-pixel two-sided z-test of src/librender/tests/test_renders.py:60-126 (variance floor 1e-4, Sidak-corrected
    significance); returns the fraction of pixel-channel values for which the null hypothesis is accepted."""
    from math import erf, sqrt
    var = np.maximum(var_per_sample, 1e-4)
    z = np.abs(mean - reference) / np.sqrt(var * (1.0 / n_mean + 1.0 / n_ref))
    cdf = 0.5 * (1.0 + np.vectorize(erf)(z / sqrt(2.0)))
    p = 2.0 * (1.0 - cdf)
    alpha = 1.0 - (1.0 - significance) ** (1.0 / (mean.size // mean.shape[-1]))
    return float(np.mean(p > alpha))


def n1_scene(res=(32, 24), spp=8, named_conductor=False):
    """SURVEY 8(f) N1 plugins: `twosided` BRDFs (an open quad seen from both sides, a rough-metal sphere), a `point`
    light and a `constant` background."""
    sd = SceneData()
    wall = M.quad([-2, -1, 1.5], [2, -1, 1.5], [2, 2, 1.5], [-2, 2, 1.5])
    wall.bsdf = sd.make_two_sided(sd.add_diffuse((0.7, 0.5, 0.3)))
    sd.add_mesh(wall)
    flip = M.quad([-2, -1.0, -3], [-2, -1.0, 1.5], [2, -1.0, 1.5], [2, -1.0, -3])      # normal points down: back side faces the camera
    flip.bsdf = sd.make_two_sided(sd.add_diffuse(0.6))
    sd.add_mesh(flip)
    s1 = M.displaced_sphere(16, 32, radius=0.6, center=(0.3, -0.3, 0.2))
    s1.bsdf = sd.make_two_sided(sd.add_roughconductor(material="Au", alpha=0.2) if named_conductor else
                                sd.add_roughconductor((0.2, 0.92, 1.1), (3.9, 2.45, 2.14), 0.2))
    sd.add_mesh(s1)
    sd.add_point([-1.0, 1.5, -1.0], (8.0, 7.0, 6.0))
    sd.add_constant((0.3, 0.4, 0.6))
    sd.add_perspective(look_at([0, 0.3, -3.2], [0, 0, 0], [0, 1, 0]), 45.0, 0.01, 100.0, res[0], res[1], "box")
    sd.max_depth, sd.rr_depth, sd.spp = 5, 3, spp
    return sd


# ---- vertex-gradient test scenes (tests/test_grad_cpu.py, tests/test_gpu_parity.py) -------------------------------------
def grid_mesh(nx, ny, size, origin, ax_u, ax_v):
    u, v = np.linspace(-0.5, 0.5, nx + 1), np.linspace(-0.5, 0.5, ny + 1)
    P = np.array([np.array(origin) + size * (a * np.array(ax_u) + b * np.array(ax_v)) for b in v for a in u], np.float32)
    F = []
    for j in range(ny):
        for i in range(nx):
            a = j * (nx + 1) + i
            F += [[a, a + 1, a + nx + 2], [a, a + nx + 2, a + nx + 1]]
    return P, np.array(F, np.uint32)


def vgrad_scene(kind, bsdf="diffuse", res=24, spp=64):
    """Scenes whose vertex gradients have no secondary-visibility component (see mitsuba2_b200/csrc/adjoint.h), mesh 0
    is the differentiated one.  plane*: one big plane filling the view (interior term only); quadenv: a small quad in
    front of a constant background (boundary term only); corner: plane + wall + area light outside the view."""
    sd = SceneData()
    if bsdf in ("diffuse", "diffuse_sky"):
        b = sd.add_diffuse([0.6, 0.5, 0.4])
    elif bsdf == "beckmann":
        b = sd.add_roughconductor(material="Cu", alpha=0.3)
    else:
        b = sd.add_roughconductor(material="Au", alpha=0.25, distribution="ggx")
    bw = sd.add_diffuse([0.3, 0.5, 0.7])
    if kind.startswith("plane") or kind == "corner":
        P, F = grid_mesh(4, 4, 8.0, [0, 0, 1.0], [1, 0, 0], [0, -1, 0])
        sd.add_mesh(Mesh(P, F, bsdf=b))
    if kind == "plane":
        sd.add_point([0.5, 1.0, -1.0], [8, 8, 8])
    if kind in ("planearea", "corner"):
        P3, F3 = grid_mesh(1, 1, 0.8, [-0.5, 1.5, -0.5], [1, 0, 0], [0, 0, 1])
        sd.add_mesh(Mesh(P3, F3, bsdf=bw), area_radiance=[20, 20, 20])
    if kind == "corner":
        P2, F2 = grid_mesh(2, 2, 3.0, [2.2, 0, -0.6], [0, 0, 1], [0, -1, 0])
        sd.add_mesh(Mesh(P2, F2, bsdf=bw))
    if kind == "wavy":      # smooth-shaded height field filling the view, distant point light: shading follows the normals only
        P, F = grid_mesh(12, 12, 8.0, [0, 0, 1.0], [1, 0, 0], [0, -1, 0])
        P[:, 2] += 0.15 * np.sin(2.0 * P[:, 0]) * np.cos(1.5 * P[:, 1])
        sd.add_mesh(Mesh(P, F, normals=M.compute_vertex_normals(P, F).astype(np.float32), bsdf=b))
        sd.add_point([500.0, 1000.0, -1000.0], [8e6, 8e6, 8e6])
    if kind == "texplane":  # a plane with a bitmap reflectance filling the view: in-plane motion slides the texture under the pixels
        P, F = grid_mesh(4, 4, 8.0, [0, 0, 1.0], [1, 0, 0], [0, -1, 0])
        uvs = ((P[:, :2] + 4.0) / 8.0).astype(np.float32)
        yy, xx = np.mgrid[0:12, 0:12]
        img = np.stack([0.5 + 0.4 * np.sin(0.9 * xx + 0.3 * yy), 0.5 + 0.4 * np.cos(0.5 * yy), 0.3 + 0.02 * xx], -1).astype(np.float32)
        sd.add_mesh(Mesh(P, F, texcoords=uvs, bsdf=sd.add_diffuse(sd.add_texture_bitmap(img))))
        sd.add_point([0.5, 1.0, -1.0], [8, 8, 8])
    if kind == "quadshadow":      # quad above a lit plane under a constant sky: its shadow on the plane moves with it
        P, F = grid_mesh(2, 2, 1.0, [0.1, 0.05, 0.0], [1, 0, 0], [0, -1, 0])
        sd.add_mesh(Mesh(P, F, bsdf=b))
        P2, F2 = grid_mesh(1, 1, 12.0, [0, 0, 2.0], [1, 0, 0], [0, -1, 0])
        sd.add_mesh(Mesh(P2, F2, bsdf=bw))
        sd.add_constant([1.0, 1.0, 1.0])
    if kind == "smoothsphere":    # smooth-shaded sphere lit from the camera's side on black: silhouette + the shading-normal terminator rim
        ms = M.displaced_sphere(20, 40, radius=0.8, center=(0.05, 0.02, 0.3), displace=False)
        ms.bsdf, ms.texcoords = b, None
        sd.add_mesh(ms)
        sd.add_point([0.6, 0.8, -3.0], [40, 40, 40])
    if kind == "indirectshadow":  # a wall lit by a floor that carries the moving quad's shadow: the upper image rows see the quad and
        # its shadow only through one diffuse bounce (boundary terms of the SECOND path vertex and of the first)
        P, F = grid_mesh(2, 2, 1.2, [0.1, -0.3, 0.2], [0, 0, 1], [1, 0, 0])          # horizontal quad, normal +y
        sd.add_mesh(Mesh(P, F, bsdf=b))
        P2, F2 = grid_mesh(1, 1, 8.0, [0, 0, 1.0], [1, 0, 0], [0, -1, 0])            # wall facing the camera
        sd.add_mesh(Mesh(P2, F2, bsdf=bw))
        P3, F3 = grid_mesh(1, 1, 12.0, [0, -1.0, 0.0], [0, 0, 1], [1, 0, 0])         # floor, normal +y
        sd.add_mesh(Mesh(P3, F3, bsdf=sd.add_diffuse([0.8, 0.8, 0.8])))
        if bsdf == "diffuse_sky":     # variant with a closed-form answer: uniform sky instead of the light panel
            sd.add_constant([1.0, 1.0, 1.0])
        else:
            P4, F4 = grid_mesh(1, 1, 1.5, [0.0, 2.5, 0.0], [1, 0, 0], [0, 0, 1])         # light panel facing down
            sd.add_mesh(Mesh(P4, F4, bsdf=bw), area_radiance=[30, 30, 30])
    if kind == "quadenv":
        P, F = grid_mesh(2, 2, 1.0, [0.1, 0.05, 0.0], [1, 0, 0], [0, -1, 0])
        sd.add_mesh(Mesh(P, F, bsdf=b))
        sd.add_constant([1.0, 1.0, 1.0])
    sd.add_perspective(look_at([0, 0, -3.0], [0, 0, 0], [0, 1, 0]), 40.0, 0.01, 100.0, res, res, "box")
    sd.max_depth, sd.rr_depth, sd.spp = (2 if kind in ("wavy", "texplane", "smoothsphere") else (3 if kind == "indirectshadow" else 4)), 100, spp
    return sd


def displacement_fields(P):
    """Global deformations used to compare per-vertex gradients with finite differences: d(loss)/d(tau) = sum_v g_v . D_v."""
    c, z = P.mean(0), np.zeros(len(P))
    return {"tx": np.tile([1.0, 0, 0], (len(P), 1)), "ty": np.tile([0, 1.0, 0], (len(P), 1)), "tz": np.tile([0, 0, 1.0], (len(P), 1)),
            "scale": P - c, "tilt": np.stack([z, z, P[:, 0] - c[0]], 1)}


def oracle_fd_vertices(sd, mesh_id, wgt, D, eps, spp, seed):
    """Central finite difference of sum(wgt * image) along the vertex displacement field D, rendered by the oracle."""
    orc = Oracle(sd)
    P0 = _f32(sd.meshes[mesh_id].positions).reshape(-1, 3)
    vals = []
    for sgn in (+1.0, -1.0):
        orc.update_vertices(mesh_id, (P0 + sgn * eps * D).astype(np.float32))
        vals.append(float((orc.render(spp=spp, seed=seed).astype(np.float64) * wgt).sum()))
    return (vals[0] - vals[1]) / (2 * eps)


# ---- the reference itself (oracle/_ref/mitsuba, built by oracle/build_reference.sh; test infrastructure) -----------------
class RefBuild:
    """The UNMODIFIED reference built from /root/reference: `dist/mitsuba` (CLI, scalar_rgb / scalar_spectral) and
    `ref_query` (oracle/ref_query.cpp linked against its libraries).  `available()` is False where it was not built."""
    DIR = os.path.join(ROOT, "oracle", "_ref", "mitsuba")
    DIST = os.path.join(DIR, "dist")
    CLI = os.path.join(DIST, "mitsuba")
    QUERY = os.path.join(DIR, "ref_query")

    @classmethod
    def available(cls, query=False):
        return os.path.exists(cls.CLI) and (not query or os.path.exists(cls.QUERY))

    @classmethod
    def _env(cls):
        return dict(os.environ, LD_LIBRARY_PATH=cls.DIST + os.pathsep + os.environ.get("LD_LIBRARY_PATH", ""))

    def __init__(self, sd, directory=None, spp=None):
        import tempfile
        from mitsuba2_b200 import export
        self.sd = sd
        self._tmp = None
        if directory is None:
            self._tmp = tempfile.TemporaryDirectory(prefix="mtsref_")
            directory = self._tmp.name
        self.dir = directory
        self.xml = export.write_mitsuba_scene(sd, directory, spp=spp)

    def render(self, spp, threads=None, variant="scalar_rgb", timeout=3600):
        """`mitsuba -m variant -t threads` -> (developed linear RGB image (H,W,3), seconds of the "Rendering finished"
        line = src/librender/integrator.cpp:171-173)."""
        import re
        import subprocess
        from mitsuba2_b200 import export, imageio
        export.write_mitsuba_scene(self.sd, self.dir, spp=spp)
        out = os.path.join(self.dir, "out.pfm")
        r = subprocess.run([self.CLI, "-m", variant, "-t", str(threads or os.cpu_count() or 1), "-o", out, self.xml],
                           env=self._env(), capture_output=True, text=True, timeout=timeout)
        if r.returncode != 0:
            raise RuntimeError("reference mitsuba failed: " + (r.stdout + r.stderr)[-2000:])
        m = re.search(r"Rendering finished\. \(took ([0-9.]+)(ms|s|m|h)\)", r.stdout + r.stderr)
        secs = float(m.group(1)) * {"ms": 1e-3, "s": 1.0, "m": 60.0, "h": 3600.0}[m.group(2)] if m else float("nan")
        return imageio.read_pfm(out), secs

    def _query(self, cmd, arr, width_out, threads=1, arg=0, timeout=3600):
        import subprocess
        fin, fout = os.path.join(self.dir, f"q_{cmd}_in.bin"), os.path.join(self.dir, f"q_{cmd}_out.bin")
        np.ascontiguousarray(arr, np.float32).tofile(fin)
        r = subprocess.run([self.QUERY, self.xml, cmd, fin, fout, str(threads), str(arg)], env=self._env(), capture_output=True,
                           text=True, timeout=timeout)
        if r.returncode != 0:
            raise RuntimeError("ref_query failed: " + (r.stdout + r.stderr)[-2000:])
        self.last_seconds = float(r.stdout.split("query_seconds")[1].split()[0])
        return np.fromfile(fout, np.float32).reshape(-1, width_out)

    def intersect(self, om, dm, naive=False, threads=None):
        """-> (n,16): t, uv, global prim bits, p, n, sh_n, wi  (`Scene::ray_intersect[_naive]`)"""
        rays = np.concatenate([_f32(om), _f32(dm)], axis=1)
        return self._query("naive" if naive else "intersect", rays, 16, threads or os.cpu_count() or 1)

    def ray_test(self, om, dm, threads=None):
        rays = np.concatenate([_f32(om), _f32(dm)], axis=1)
        return self._query("test", rays, 1, threads or os.cpu_count() or 1)[:, 0] > 0.5

    def bsdf(self, shape, wi, uv, s1, s2, wo):
        """-> (n,13): sample (wo, pdf, eta, weight rgb, delta), eval(wo) rgb, pdf(wo) of the BSDF of shape `shape`"""
        q = np.concatenate([_f32(wi), _f32(uv), _f32(s1).reshape(-1, 1), _f32(s2), _f32(wo)], axis=1)
        return self._query("bsdf", q, 13, 1, shape)

    def sample_emitter_direction(self, ref_p, s2):
        return self._query("emitter", np.concatenate([_f32(ref_p), _f32(s2)], axis=1), 11)

    def sensor_sample_ray(self, pos, sensor=0):
        r = self._query("sensor", _f32(pos), 8, 1, sensor)
        return r[:, :4].copy(), r[:, 4:].copy()


def attribute_scene(res=(48, 32), spp=8, per_face=False):
    """`mesh_attribute` textures (src/textures/mesh_attribute.cpp): a sphere whose diffuse reflectance is a per-vertex
    colour attribute (or per-face), and a ground quad whose reflectance is a per-face scalar attribute x scale."""
    sd = SceneData()
    rng = np.random.default_rng(12)
    s1 = M.displaced_sphere(20, 40)
    s1.attributes["vertex_color"] = (0.1 + 0.8 * rng.random((len(s1.positions), 3))).astype(np.float32)
    s1.attributes["face_color"] = (0.1 + 0.8 * rng.random((len(s1.faces), 3))).astype(np.float32)
    s1.bsdf = sd.add_diffuse(sd.add_texture_mesh_attribute(None, per_face=per_face, name="face_color" if per_face else "vertex_color"))
    s1.bsdf = sd.bind_mesh_attributes(s1)
    sd.add_mesh(s1)
    ground = M.quad([-10, -1.2, -10], [-10, -1.2, 10], [10, -1.2, 10], [10, -1.2, -10])
    ground.attributes["face_weight"] = np.array([[0.9], [0.3]], np.float32)
    ground.bsdf = sd.add_diffuse(sd.add_texture_mesh_attribute(None, per_face=True, scale=0.8, name="face_weight"))
    ground.bsdf = sd.bind_mesh_attributes(ground)
    sd.add_mesh(ground)
    light = M.quad([-0.5, 3, -0.5], [0.5, 3, -0.5], [0.5, 3, 0.5], [-0.5, 3, 0.5])
    light.bsdf = sd.add_diffuse(0.5)
    sd.add_mesh(light, area_radiance=(18.4, 15.6, 8.0))
    sd.add_constant((0.2, 0.25, 0.3))
    sd.add_perspective(look_at([0, 0, -3.2], [0, 0, 0], [0, 1, 0]), 45.0, 0.01, 100.0, res[0], res[1], "box")
    sd.max_depth, sd.rr_depth, sd.spp = 4, 5, spp
    return sd


def _trs(translate=(0, 0, 0), rotate_axis=(0, 0, 1), rotate_deg=0.0, scale=(1, 1, 1)):
    """4x4 translate * rotate(axis, angle) * scale"""
    a = np.asarray(rotate_axis, np.float64)
    a = a / np.linalg.norm(a)
    th = np.deg2rad(rotate_deg)
    K = np.array([[0, -a[2], a[1]], [a[2], 0, -a[0]], [-a[1], a[0], 0]])
    R = np.eye(3) + np.sin(th) * K + (1 - np.cos(th)) * (K @ K)
    m = np.eye(4)
    m[:3, :3] = R @ np.diag(np.broadcast_to(np.asarray(scale, np.float64), 3))
    m[:3, 3] = translate
    return m


def shapes_scene(res=(48, 32), spp=8, emitters="disk", tri_floor=True, named_conductor=False):
    """Analytic shapes as leaf primitives next to triangles (src/shapes/{sphere,disk,cylinder,rectangle}.cpp): a diffuse
    sphere, a rough-conductor sphere with flipped normals seen through an open cylinder, a bitmap-textured rectangle wall,
    a textured disk, a triangle floor, and an area light on a disk / sphere / cylinder / rectangle (`emitters`, a
    space-separated list)."""
    sd = SceneData()
    rng = np.random.default_rng(5)
    tex = sd.add_texture_bitmap((0.15 + 0.8 * rng.random((6, 8, 3))).astype(np.float32))
    if tri_floor:
        floor = M.quad([-6, -1, -6], [-6, -1, 6], [6, -1, 6], [6, -1, -6])
        floor.bsdf = sd.add_diffuse((0.5, 0.45, 0.4))
        sd.add_mesh(floor)
    else:
        sd.add_rectangle(_trs((0, -1, 0), (1, 0, 0), -90.0, (6, 6, 1)), bsdf=sd.add_diffuse((0.5, 0.45, 0.4)))
    sd.add_sphere((-1.2, -0.3, 0.4), 0.7, bsdf=sd.add_diffuse(tex))
    sd.add_sphere((1.3, -0.45, 0.2), 0.55, to_world=_trs((0, 0, 0), (0, 1, 0), 30.0), bsdf=sd.add_roughconductor(material="Cu", alpha=0.15) if named_conductor else sd.add_roughconductor((0.2, 0.92, 1.1), (3.9, 2.45, 2.14), 0.15))
    sd.add_cylinder((-0.3, -1.0, 1.6), (0.4, 0.3, 1.9), 0.35, bsdf=sd.make_two_sided(sd.add_diffuse((0.2, 0.5, 0.7))))
    sd.add_rectangle(_trs((0, 0.5, 3.0), (0, 1, 0), 180.0, (3.0, 1.5, 1)), bsdf=sd.add_diffuse(tex))
    sd.add_disk(_trs((0.2, -0.99, -0.8), (1, 0, 0), -90.0, (0.8, 0.5, 1)), bsdf=sd.add_diffuse(tex))
    em = emitters.split()
    lb = sd.add_diffuse(0.0)
    if "disk" in em:
        sd.add_disk(_trs((0, 2.5, 0.5), (1, 0, 0), 90.0, (0.6, 0.6, 1)), bsdf=lb, area_radiance=(14.0, 12.0, 9.0))
    if "sphere" in em:
        sd.add_sphere((2.2, 1.2, 1.0), 0.25, bsdf=lb, area_radiance=(20.0, 10.0, 10.0))
    if "cylinder" in em:
        sd.add_cylinder((-2.5, 0.2, 1.0), (-2.5, 1.2, 1.0), 0.12, bsdf=lb, area_radiance=(6.0, 12.0, 18.0))
    if "rectangle" in em:
        sd.add_rectangle(_trs((0, 2.2, -1.0), (1, 0, 0), 90.0, (0.5, 0.3, 1)), bsdf=lb, flip_normals=False, area_radiance=(12.0, 12.0, 12.0))
    sd.add_constant((0.08, 0.1, 0.12))
    sd.add_perspective(look_at([0, 0.6, -4.2], [0, -0.1, 0], [0, 1, 0]), 50.0, 0.01, 100.0, res[0], res[1], "box")
    sd.max_depth, sd.rr_depth, sd.spp = 5, 5, spp
    return sd


def set_rfilter(sd, kind, sensor=0):
    """Replace the reconstruction filter of a sensor (defaults of the reference's plugins: gaussian stddev 0.5 / radius 2,
    mitchell B = C = 1/3 / radius 2, src/rfilters/{gaussian,mitchell}.cpp)."""
    from mitsuba2_b200 import _abi as A
    s = sd.sensors[sensor]
    if kind == "gaussian":
        s.update(rfilter=A.RFILTER_GAUSSIAN, radius=2.0, stddev=0.5)
    elif kind == "mitchell":
        s.update(rfilter=A.RFILTER_MITCHELL, radius=2.0, stddev=1.0 / 3.0, filter_c=1.0 / 3.0)
    elif kind == "box":
        s.update(rfilter=A.RFILTER_BOX, radius=0.5)
    else:
        raise ValueError(kind)
    return sd
