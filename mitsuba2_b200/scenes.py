"""Synthetic benchmark scenes of BASELINE.json / SURVEY.md section 8(d), built procedurally (no files needed)."""
import numpy as np

from . import mesh as M
from .scenedata import SceneData, look_at

CU_ETA, CU_K = (0.200438, 0.924033, 1.102212), (3.912949, 2.452848, 2.142188)


def synthetic_envmap(w=1024, h=512, seed=7):
    """Stand-in for resources/data/envmap/museum.exr (not available on the GPU box): smooth sky gradient, a bright
    sun blob and low-amplitude PCG-free noise, deterministic."""
    v, u = np.meshgrid((np.arange(h) + 0.5) / h, (np.arange(w) + 0.5) / w, indexing="ij")
    sky = np.stack([0.3 + 0.4 * (1 - v), 0.4 + 0.4 * (1 - v), 0.6 + 0.4 * (1 - v)], axis=2)
    sun = 40.0 * np.exp(-(((u - 0.3) * 2) ** 2 + (v - 0.25) ** 2) / 0.004)
    rng = np.random.default_rng(seed)
    img = sky * (0.9 + 0.2 * rng.random((h, w, 1))) + sun[:, :, None] * np.array([1.0, 0.9, 0.7])
    return img.astype(np.float32)


def c3_envmap():
    """-> (rgb image, name): resources/data/envmap/museum.exr of the reference where tools/fetch_reference_data.py has
    copied it into mitsuba2_b200/data/ (SURVEY 8(d) names it for C3), else the synthetic sky.  MTSB_C3_ENVMAP=synthetic forces the latter."""
    import os
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "data", "museum_envmap.npy")
    if os.environ.get("MTSB_C3_ENVMAP", "") != "synthetic" and os.path.exists(path):
        return np.load(path), "museum.exr 1000x500 (reference resources/data/envmap)"
    return synthetic_envmap(), "synthetic sky 1024x512 (museum.exr not present)"


def c2_scene(n_theta=500, n_phi=1000, width=1920, height=1080):
    """C2: 1M-triangle displaced sphere, camera (0,0,-3.2)->0, fov 45 (x), near .01 far 100."""
    sd = SceneData()
    m = M.displaced_sphere(n_theta, n_phi)
    m.bsdf = sd.add_diffuse(0.5)
    sd.add_mesh(m)
    sd.add_perspective(look_at([0, 0, -3.2], [0, 0, 0], [0, 1, 0]), 45.0, 0.01, 100.0, width, height, "box")
    sd.max_depth, sd.rr_depth = 6, 5
    return sd


def c2_rays(sd, kind, n=None, seed=1):
    """coherent: one primary ray per pixel centre, row-major.  incoherent: origins uniform in the ball r=1.5,
    directions uniform on the sphere.  Shadow variants: maxt ~ U[0,3)."""
    s = sd.sensors[0]
    w, h = s["width"], s["height"]
    n = w * h if n is None else n
    if kind.startswith("coherent"):
        ys, xs = np.divmod(np.arange(n), w)
        px, py = (xs + 0.5) / w, (ys % h + 0.5) / h
        m = s["sample_to_camera"].reshape(4, 4).astype(np.float64)
        p = np.stack([px, py, np.zeros(n), np.ones(n)], axis=0)
        q = m @ p
        near = (q[:3] / q[3]).T
        d = near / np.linalg.norm(near, axis=1, keepdims=True)
        tw = s["to_world"].reshape(4, 4).astype(np.float64)
        dw = d @ tw[:3, :3].T
        om = np.concatenate([np.broadcast_to(tw[:3, 3], (n, 3)), (s["near"] / d[:, 2])[:, None]], axis=1)
        dm = np.concatenate([dw, (s["far"] / d[:, 2])[:, None]], axis=1)
    else:
        rng = np.random.default_rng(seed)
        o = rng.normal(size=(n, 3))
        o = o / np.linalg.norm(o, axis=1, keepdims=True) * (rng.random((n, 1)) ** (1 / 3)) * 1.5
        d = rng.normal(size=(n, 3))
        d /= np.linalg.norm(d, axis=1, keepdims=True)
        om = np.concatenate([o, np.zeros((n, 1))], axis=1)
        dm = np.concatenate([d, np.full((n, 1), np.inf)], axis=1)
    if kind.endswith("shadow"):
        dm[:, 3] = np.random.default_rng(seed + 1).random(n) * 3.0
    return om.astype(np.float32), dm.astype(np.float32)


def c3_scene(n_theta=1000, n_phi=2000, width=1920, height=1080, small=(158, 316), tex_res=512):
    """C3: 4M-triangle displaced sphere (roughconductor, Beckmann 0.1, Cu), 20x20 ground quad at y=-1.2 (diffuse,
    512^2 bitmap), ~100k-triangle dielectric sphere r=0.4 at (1.6,-0.8,0), 1x1 area light at y=3, envmap."""
    sd = SceneData()
    big = M.displaced_sphere(n_theta, n_phi)
    big.bsdf = sd.add_roughconductor(material="Cu", alpha=0.1)     # rgb mode: CU_ETA / CU_K; spectral mode: Cu.{eta,k}.spd
    sd.add_mesh(big)
    tex = 0.2 + 0.6 * M._pcg32_stream(tex_res * tex_res * 3, 3, 1).reshape(tex_res, tex_res, 3)
    ground = M.quad([-10, -1.2, -10], [-10, -1.2, 10], [10, -1.2, 10], [10, -1.2, -10])
    ground.bsdf = sd.add_diffuse(sd.add_texture_bitmap(tex))
    sd.add_mesh(ground)
    glass = M.displaced_sphere(small[0], small[1], radius=0.4, center=(1.6, -0.8, 0.0), displace=False)
    glass.bsdf = sd.add_dielectric(1.5046, 1.000277)
    sd.add_mesh(glass)
    light = M.quad([-0.5, 3, -0.5], [0.5, 3, -0.5], [0.5, 3, 0.5], [-0.5, 3, 0.5])
    light.bsdf = sd.add_diffuse(0.5)
    sd.add_mesh(light, area_radiance=(18.4, 15.6, 8.0))
    env, sd.envmap_name = c3_envmap()
    sd.add_envmap(env)
    sd.add_perspective(look_at([0, 0, -3.2], [0, 0, 0], [0, 1, 0]), 45.0, 0.01, 100.0, width, height, "box")
    sd.max_depth, sd.rr_depth, sd.spp = 6, 5, 1024
    return sd


def c4_scene(cbox_npz, res=512, spp=64, tex_res=512, sphere=(100, 100), seed=4, displace=True, smooth=True):
    """C4 (gradient pass): Cornell-box geometry (fixture `cbox_npz`) whose floor reflectance is a tex_res^2 RGB bitmap
    (values U[0.2,0.8)) and whose small box is replaced by a ~20k-triangle displaced sphere.  Returns (scene, tex id)."""
    sd = SceneData.load_npz(cbox_npz)
    s = sd.sensors[0]
    sd.sensors = []
    sd.add_perspective(s["to_world"].reshape(4, 4), s["fov_x"], s["near"], s["far"], res, res, "box")
    rng = np.random.default_rng(seed)
    tex = sd.add_texture_bitmap((0.2 + 0.6 * rng.random((tex_res, tex_res, 3))).astype(np.float32))
    floor = next(m for m in sd.meshes if "floor" in m.name)
    p = floor.positions
    lo, hi = p.min(axis=0), p.max(axis=0)
    floor.texcoords = np.ascontiguousarray(((p - lo) / np.maximum(hi - lo, 1e-6))[:, [0, 2]], np.float32)
    floor.bsdf = sd.add_diffuse(tex)
    for i, m in enumerate(sd.meshes):
        if "smallbox" in m.name:
            c = 0.5 * (m.positions.min(axis=0) + m.positions.max(axis=0))
            r = 0.5 * float((m.positions.max(axis=0) - m.positions.min(axis=0)).min())
            sph = M.displaced_sphere(sphere[0], sphere[1], radius=r, center=tuple(c), displace=displace)
            if not smooth:          # diagnostic variants (tools/c4_fd_check.py): flat shading / undisplaced convex sphere
                sph.normals = None
            sph.bsdf, sph.name = m.bsdf, "c4_sphere"
            sd.meshes[i] = sph
    sd.spp = spp
    return sd, tex
