"""Host-side flattened scene description (numpy) and its conversion to ``mtsb_scene_desc``.

This is the Python mirror of what ``Scene::Scene`` collects (src/librender/scene.cpp:22-99):
shapes (triangle meshes), their BSDFs / area emitters, environment emitter, sensors and
the integrator properties.  The same object feeds the CUDA library (product) and, in the
tests only, the CPU oracle.
"""
import ctypes as C
import math

import numpy as np

from . import _abi as A


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _ptr(a, ty):
    return a.ctypes.data_as(ty) if a is not None else ty()


def perspective_sample_to_camera(x_fov_deg, near, far, aspect, rel_offset=(0.0, 0.0), rel_size=(1.0, 1.0)):
    """``PerspectiveCamera::update_camera_transforms`` (src/sensors/perspective.cpp:106-133): inverse of
    scale(1/rel_size) * translate(-rel_offset) * scale(-.5,-.5*aspect,1) * translate(-1,-1/aspect,0) *
    perspective(fov,near,far) (``Transform::perspective``, include/mitsuba/core/transform.h:203-220); rel_offset /
    rel_size = the film's crop window relative to the full film (aspect = full film width / height)."""
    tan = math.tan(math.radians(x_fov_deg * 0.5))
    inv_persp = np.array([[tan, 0, 0, 0], [0, tan, 0, 0], [0, 0, 0, 1],
                          [0, 0, (near - far) / (far * near), 1.0 / near]], dtype=np.float64)
    inv_trans = np.array([[1, 0, 0, 1], [0, 1, 0, 1.0 / aspect], [0, 0, 1, 0], [0, 0, 0, 1]], dtype=np.float64)
    inv_scale = np.diag([-2.0, -2.0 / aspect, 1.0, 1.0])
    inv_crop = np.array([[rel_size[0], 0, 0, rel_offset[0]], [0, rel_size[1], 0, rel_offset[1]], [0, 0, 1, 0], [0, 0, 0, 1]], dtype=np.float64)
    return (inv_persp @ inv_trans @ inv_scale @ inv_crop).astype(np.float32)


def parse_fov(fov=None, fov_axis="x", focal_length="50mm", aspect=1.0):
    """``parse_fov`` (src/librender/sensor.cpp:119-173) -> horizontal fov in degrees."""
    if fov is not None:
        fov_axis = fov_axis.lower()
        if fov_axis == "smaller":
            fov_axis = "y" if aspect > 1 else "x"
        elif fov_axis == "larger":
            fov_axis = "x" if aspect > 1 else "y"
    else:
        f = focal_length[:-2] if str(focal_length).endswith("mm") else focal_length
        value = float(f)
        fov = 2.0 * math.degrees(math.atan(math.sqrt(36 * 36 + 24 * 24) / (2.0 * value)))
        fov_axis = "diagonal"
    if fov_axis == "x":
        result = fov
    elif fov_axis == "y":
        result = math.degrees(2.0 * math.atan(math.tan(0.5 * math.radians(fov)) * aspect))
    elif fov_axis == "diagonal":
        diagonal = 2.0 * math.tan(0.5 * math.radians(fov))
        width = diagonal / math.sqrt(1.0 + 1.0 / (aspect * aspect))
        result = math.degrees(2.0 * math.atan(width * 0.5))
    else:
        raise ValueError("The 'fov_axis' parameter must be set to one of 'smaller', 'larger', 'diagonal', 'x', or 'y'!")
    if result <= 0.0 or result >= 180.0:
        raise ValueError("The horizontal field of view must be in the range [0, 180]!")
    return result


def look_at(origin, target, up):
    """``Transform::look_at`` (include/mitsuba/core/transform.h): camera-to-world, +z forward, left-handed image x."""
    origin, target, up = (np.asarray(v, dtype=np.float64) for v in (origin, target, up))
    d = target - origin
    d /= np.linalg.norm(d)
    left = np.cross(up, d)
    left /= np.linalg.norm(left)
    new_up = np.cross(d, left)
    m = np.eye(4)
    m[:3, 0], m[:3, 1], m[:3, 2], m[:3, 3] = left, new_up, d, origin
    return m


class Mesh:
    def __init__(self, positions, faces, normals=None, texcoords=None, bsdf=0, emitter=-1, name="mesh"):
        self.device = hasattr(positions, "is_cuda") and positions.is_cuda
        if self.device:       # CUDA torch tensors: handed to the library as device pointers (mtsb_mesh_desc.mem = MTSB_MEM_DEVICE)
            import torch
            self.positions = positions.to(torch.float32).reshape(-1, 3).contiguous()
            self.faces = faces.to(device=positions.device, dtype=torch.int32).reshape(-1, 3).contiguous()
            self.normals = normals.to(torch.float32).reshape(-1, 3).contiguous() if normals is not None else None
            self.texcoords = texcoords.to(torch.float32).reshape(-1, 2).contiguous() if texcoords is not None else None
            self.bsdf, self.emitter, self.name, self.attributes = int(bsdf), int(emitter), name, {}
            return
        self.positions = _f32(positions).reshape(-1, 3)
        self.faces = np.ascontiguousarray(faces, dtype=np.uint32).reshape(-1, 3)
        self.normals = _f32(normals).reshape(-1, 3) if normals is not None else None
        self.texcoords = _f32(texcoords).reshape(-1, 2) if texcoords is not None else None
        self.bsdf, self.emitter, self.name = int(bsdf), int(emitter), name
        self.attributes = {}       # "vertex_<name>" / "face_<name>" -> (n, size) float32 (Mesh::m_mesh_attributes, mesh.h)


class SceneData:
    """Flattened scene.  Lists of plain dicts + :class:`Mesh`; ``to_desc()`` builds the C struct."""

    def __init__(self):
        self.meshes, self.bsdfs, self.textures, self.emitters, self.sensors = [], [], [], [], []
        self.shapes = []                # analytic shapes (sphere / disk / cylinder / rectangle): BVH leaf primitives
        self.max_depth, self.rr_depth = -1, 5
        self.spp = 4
        # "rgb" (scalar_rgb semantics) or "spectral" (scalar_spectral): the same authored scene serves both; to_desc()
        # performs the conversions the reference's plugins do at load time in a spectral variant (RGB -> model
        # coefficients, emitters -> coefficients x D65, named conductors -> tabulated eta/k)
        self.mode = "rgb"
        self.rgb2spec_model = None      # optional (scale, data) table for rgb2spec.fetch (tests use a small one)

    # ---- builders -------------------------------------------------------------------------
    def add_texture_rgb(self, rgb):
        rgb = [float(rgb)] * 3 if np.isscalar(rgb) else [float(x) for x in rgb]
        self.textures.append({"type": A.TEX_CONSTANT, "value": rgb})
        return len(self.textures) - 1

    def add_texture_spectrum(self, wavelengths, values):
        """``irregular`` spectrum (src/spectra/irregular.cpp); spectral mode only."""
        w, v = _f32(wavelengths).reshape(-1), _f32(values).reshape(-1)
        assert w.shape == v.shape and w.size >= 2 and np.all(np.diff(w) > 0)
        self.textures.append({"type": A.TEX_IRREGULAR, "data": np.ascontiguousarray(np.concatenate([w, v]))})
        return len(self.textures) - 1

    def add_texture_regular(self, lambda_min, lambda_max, values):
        """``regular`` spectrum (src/spectra/regular.cpp); spectral mode only."""
        self.textures.append({"type": A.TEX_REGULAR, "data": _f32(values).reshape(-1), "lambda_min": float(lambda_min), "lambda_max": float(lambda_max)})
        return len(self.textures) - 1

    def add_texture_uniform(self, value):
        """``uniform`` spectrum (src/spectra/uniform.cpp); spectral mode only."""
        self.textures.append({"type": A.TEX_UNIFORM, "value": [float(value)] * 3})
        return len(self.textures) - 1

    def add_texture_coeff(self, coeff, d65_scale=0.0):
        """Constant texture given directly as sRGB-model coefficients (spectral mode; what ``srgb`` stores, srgb.cpp:44-48)."""
        self.textures.append({"type": A.TEX_CONSTANT, "value": [float(x) for x in coeff], "coeff": True, "d65_scale": float(d65_scale)})
        return len(self.textures) - 1

    def _tex(self, x):
        """texture id | rgb triple / scalar | {"wavelengths", "values"} -> texture id"""
        if isinstance(x, (int, np.integer)) and not isinstance(x, bool):
            return int(x)
        if isinstance(x, dict):
            return self.add_texture_spectrum(x["wavelengths"], x["values"])
        return self.add_texture_rgb(x)

    def add_texture_bitmap(self, data, to_uv=None):
        data = _f32(data)
        if data.ndim == 2:
            data = data[:, :, None]
        assert data.ndim == 3 and data.shape[2] in (1, 3)
        self.textures.append({"type": A.TEX_BITMAP, "data": np.ascontiguousarray(data),
                              "to_uv": _f32(np.eye(3) if to_uv is None else to_uv).reshape(9)})
        return len(self.textures) - 1

    def bind_mesh_attributes(self, mesh):
        """A ``mesh_attribute`` texture declared by name only is bound to the attribute buffer of the mesh whose BSDF uses it
        (``Mesh::eval_attribute`` looks the name up in the shape that was hit): returns the BSDF id to use for `mesh` -- a
        copy with per-mesh textures when any of its textures is an unbound attribute, else ``mesh.bsdf``."""
        b = self.bsdfs[mesh.bsdf]
        keys = [k for k, v in b.items() if (k.endswith("_tex")) and isinstance(v, int) and v >= 0
                and self.textures[v]["type"] == A.TEX_MESH_ATTRIBUTE and self.textures[v].get("data") is None]
        if not keys:
            return mesh.bsdf
        nb = dict(b)
        for k in keys:
            t = self.textures[b[k]]
            attrs = getattr(mesh, "attributes", {})
            if t["name"] not in attrs:
                raise RuntimeError(f"Invalid attribute requested {t['name']}.")      # mesh.cpp:530-531
            nb[k] = self.add_texture_mesh_attribute(attrs[t["name"]], t["per_face"], t["scale"], t["name"])
        self.bsdfs.append(nb)
        return len(self.bsdfs) - 1

    def add_texture_mesh_attribute(self, data, per_face=False, scale=1.0, name=None):
        """``mesh_attribute`` texture (src/textures/mesh_attribute.cpp): `data` (n, 1|3) = one element per vertex (barycentric
        interpolation) or per face of the mesh whose BSDF uses it.  Spectral mode: 3-float elements are taken as rgb and
        converted to sRGB-model coefficients here (the reference expects the caller to have done that, mesh_attribute.cpp:54-58)."""
        if data is None:          # declared by name (XML): bound per mesh by bind_mesh_attributes
            self.textures.append({"type": A.TEX_MESH_ATTRIBUTE, "data": None, "per_face": bool(per_face), "scale": float(scale), "name": name})
            return len(self.textures) - 1
        d = _f32(data)
        d = d.reshape(-1, 1) if d.ndim == 1 else d.reshape(d.shape[0], -1)
        if d.shape[1] not in (1, 3):
            raise RuntimeError(f'eval_attribute(): Attribute "{name}" requested but had size {d.shape[1]}.')      # mesh.cpp:543
        self.textures.append({"type": A.TEX_MESH_ATTRIBUTE, "data": np.ascontiguousarray(d), "per_face": bool(per_face), "scale": float(scale), "name": name})
        return len(self.textures) - 1

    def add_diffuse(self, reflectance=0.5):
        tex = reflectance if isinstance(reflectance, int) and not isinstance(reflectance, bool) else self.add_texture_rgb(reflectance)
        self.bsdfs.append({"type": A.BSDF_DIFFUSE, "reflectance_tex": tex})
        return len(self.bsdfs) - 1

    def add_roughconductor(self, eta=None, k=None, alpha=0.1, alpha_v=None, distribution="beckmann", sample_visible=True,
                           specular_reflectance=None, material=None):
        """``eta``/``k``: rgb triples, spectra ({"wavelengths", "values"}) or texture ids; or ``material`` = a named
        conductor (roughconductor.cpp:155 -> complex_ior_from_file): rgb mode uses the pre-integrated colour, spectral
        mode the tabulated optical constants."""
        if material is not None:
            from ._ior_data import CONDUCTORS_SPD
            from .xml import CONDUCTORS_RGB
            name = material.lower()
            spd = CONDUCTORS_SPD.get(name)
            eta_id = self.add_texture_rgb(CONDUCTORS_RGB[name][0])
            k_id = self.add_texture_rgb(CONDUCTORS_RGB[name][1])
            self.textures[eta_id]["material"] = self.textures[k_id]["material"] = material      # kept for export.py
            if spd is not None:      # spectral alternative of the same two textures
                self.textures[eta_id]["spectrum"] = (spd["wavelengths"], spd["eta"])
                self.textures[k_id]["spectrum"] = (spd["wavelengths"], spd["k"])
            eta, k = eta_id, k_id
        self.bsdfs.append({"type": A.BSDF_ROUGHCONDUCTOR, "eta_tex": self._tex(eta),
                           "k_tex": self._tex(k), "alpha_u": float(alpha),
                           "alpha_v": float(alpha if alpha_v is None else alpha_v),
                           "distribution": A.MF_GGX if distribution.lower() == "ggx" else A.MF_BECKMANN,
                           "sample_visible": int(bool(sample_visible)),
                           "specular_reflectance_tex": -1 if specular_reflectance is None else self.add_texture_rgb(specular_reflectance)})
        return len(self.bsdfs) - 1

    def add_dielectric(self, int_ior=1.5046, ext_ior=1.000277, specular_reflectance=None, specular_transmittance=None):
        self.bsdfs.append({"type": A.BSDF_DIELECTRIC, "eta": float(int_ior) / float(ext_ior),
                           "specular_reflectance_tex": -1 if specular_reflectance is None else self.add_texture_rgb(specular_reflectance),
                           "specular_transmittance_tex": -1 if specular_transmittance is None else self.add_texture_rgb(specular_transmittance)})
        return len(self.bsdfs) - 1

    def add_conductor(self, eta=None, k=None, specular_reflectance=None, material=None):
        """Smooth ``conductor`` (src/bsdfs/conductor.cpp): delta reflection weighted by the conductor Fresnel term."""
        idx = self.add_roughconductor(eta, k, specular_reflectance=specular_reflectance, material=material)
        b = self.bsdfs[idx]
        self.bsdfs[idx] = {"type": A.BSDF_CONDUCTOR, "eta_tex": b["eta_tex"], "k_tex": b["k_tex"],
                           "specular_reflectance_tex": b["specular_reflectance_tex"]}
        return idx

    def _texture_mean(self, tex_id):
        """``Texture::mean`` of an rgb variant: srgb.cpp:52-57 (mean of the colour), bitmap.cpp:333 (mean of the texels)."""
        t = self.textures[tex_id]
        return float(np.mean(t["value"])) if t["type"] == A.TEX_CONSTANT else float(np.mean(t["data"]))

    def add_plastic(self, diffuse_reflectance=0.5, int_ior=1.49, ext_ior=1.000277, specular_reflectance=None, nonlinear=False):
        """Smooth ``plastic`` (src/bsdfs/plastic.cpp): dielectric coating (delta reflection) over a diffuse base; defaults
        `polypropylene` / `air` (:138-141).  The lobe-selection weight uses the textures' rgb means in both colour modes."""
        d = self._tex(diffuse_reflectance)
        s_id = -1 if specular_reflectance is None else self._tex(specular_reflectance)
        d_mean, s_mean = self._texture_mean(d), (1.0 if s_id < 0 else self._texture_mean(s_id))
        self.bsdfs.append({"type": A.BSDF_PLASTIC, "reflectance_tex": d, "specular_reflectance_tex": s_id,
                           "eta": float(int_ior) / float(ext_ior), "nonlinear": int(bool(nonlinear)),
                           "specular_sampling_weight": s_mean / (d_mean + s_mean)})
        return len(self.bsdfs) - 1

    def add_roughplastic(self, diffuse_reflectance=0.5, int_ior=1.49, ext_ior=1.000277, alpha=0.1, distribution="beckmann",
                         sample_visible=True, specular_reflectance=None, nonlinear=False):
        """``roughplastic`` (src/bsdfs/roughplastic.cpp): rough dielectric coating over a diffuse base.  The library computes the
        transmittance tables of the coating when the scene is created."""
        d = self._tex(diffuse_reflectance)
        s_id = -1 if specular_reflectance is None else self._tex(specular_reflectance)
        d_mean, s_mean = self._texture_mean(d), (1.0 if s_id < 0 else self._texture_mean(s_id))
        self.bsdfs.append({"type": A.BSDF_ROUGHPLASTIC, "reflectance_tex": d, "specular_reflectance_tex": s_id,
                           "eta": float(int_ior) / float(ext_ior), "alpha_u": float(alpha), "alpha_v": float(alpha),
                           "distribution": A.MF_GGX if distribution.lower() == "ggx" else A.MF_BECKMANN,
                           "sample_visible": int(bool(sample_visible)), "nonlinear": int(bool(nonlinear)),
                           "specular_sampling_weight": s_mean / (d_mean + s_mean)})
        return len(self.bsdfs) - 1

    def add_roughdielectric(self, int_ior=1.5046, ext_ior=1.000277, alpha=0.1, alpha_v=None, distribution="beckmann", sample_visible=True,
                            specular_reflectance=None, specular_transmittance=None):
        """``roughdielectric`` (src/bsdfs/roughdielectric.cpp): microfacet reflection and refraction."""
        self.bsdfs.append({"type": A.BSDF_ROUGHDIELECTRIC, "eta": float(int_ior) / float(ext_ior), "alpha_u": float(alpha),
                           "alpha_v": float(alpha if alpha_v is None else alpha_v),
                           "distribution": A.MF_GGX if distribution.lower() == "ggx" else A.MF_BECKMANN,
                           "sample_visible": int(bool(sample_visible)),
                           "specular_reflectance_tex": -1 if specular_reflectance is None else self.add_texture_rgb(specular_reflectance),
                           "specular_transmittance_tex": -1 if specular_transmittance is None else self.add_texture_rgb(specular_transmittance)})
        return len(self.bsdfs) - 1

    def add_mask(self, nested_bsdf_id, opacity=0.5):
        """``mask`` adapter (src/bsdfs/mask.cpp) around an existing BSDF: with probability 1 - opacity the ray passes straight
        through.  `opacity`: a scalar or an image (H,W) / (H,W,3); stored as a single-channel bitmap, i.e. what `eval_1` of
        the reference's texture returns (luminance for colour images, bitmap.cpp:215-231)."""
        o = np.asarray(opacity, np.float32)
        if o.ndim == 0:
            o = np.full((2, 2), float(o), np.float32)
        if o.ndim == 3:
            o = o[:, :, :3] @ np.array([0.212671, 0.715160, 0.072169], np.float32) if o.shape[2] >= 3 else o[:, :, 0]
        b = dict(self.bsdfs[nested_bsdf_id])
        b["opacity_tex"] = self.add_texture_bitmap(np.ascontiguousarray(o[:, :, None], np.float32))
        self.bsdfs.append(b)
        return len(self.bsdfs) - 1

    def add_mesh(self, mesh, area_radiance=None):
        self.meshes.append(mesh)
        mid = len(self.meshes) - 1
        if area_radiance is not None:
            self.emitters.append(dict(self._radiance(area_radiance), type=A.EMITTER_AREA, mesh_id=mid))
            mesh.emitter = len(self.emitters) - 1
        return mid

    # ---- analytic shapes (src/shapes/{sphere,disk,cylinder,rectangle}.cpp) ----------------------------------
    def add_shape(self, kind, to_world=None, bsdf=0, flip_normals=False, area_radiance=None, name=None):
        """One analytic primitive.  ``to_world`` is the plugin's final object-to-world matrix (see add_sphere /
        add_cylinder for the parameters the plugins fold into it).  Global primitive index = number of mesh faces +
        position in ``shapes``; an area emitter on it is addressed as mesh ``len(meshes) + shape id`` in the C ABI."""
        kinds = {"sphere": A.SHAPE_SPHERE, "disk": A.SHAPE_DISK, "cylinder": A.SHAPE_CYLINDER, "rectangle": A.SHAPE_RECTANGLE}
        sh = {"type": kinds[kind] if isinstance(kind, str) else int(kind), "to_world": _f32(np.eye(4) if to_world is None else to_world).reshape(4, 4),
              "flip_normals": bool(flip_normals), "bsdf": int(bsdf), "emitter": -1, "name": name or str(kind)}
        self.shapes.append(sh)
        sid = len(self.shapes) - 1
        if area_radiance is not None:
            self.emitters.append(dict(self._radiance(area_radiance), type=A.EMITTER_AREA, shape_id=sid))
            sh["emitter"] = len(self.emitters) - 1
        return sid

    def add_sphere(self, center=(0.0, 0.0, 0.0), radius=1.0, to_world=None, **kw):
        """sphere.cpp:100-110: to_world * translate(center) * scale(radius)"""
        m = np.eye(4) if to_world is None else np.asarray(to_world, np.float64).reshape(4, 4)
        t = np.eye(4); t[:3, 3] = center
        return self.add_shape("sphere", m @ t @ np.diag([radius, radius, radius, 1.0]), **kw)

    def add_disk(self, to_world=None, **kw):
        return self.add_shape("disk", to_world, **kw)

    def add_rectangle(self, to_world=None, **kw):
        return self.add_shape("rectangle", to_world, **kw)

    def add_cylinder(self, p0=(0.0, 0.0, 0.0), p1=(0.0, 0.0, 1.0), radius=1.0, to_world=None, **kw):
        """cylinder.cpp:84-97: to_world * translate(p0) * to_frame(Frame(p1 - p0)) * scale(radius, radius, 1); the frame is
        built from the raw difference vector as the plugin does (Frame3f(v) -> coordinate_system(v), frame.h / vector.h)."""
        m = np.eye(4) if to_world is None else np.asarray(to_world, np.float64).reshape(4, 4)
        n = np.asarray(p1, np.float64) - np.asarray(p0, np.float64)
        sign = np.copysign(1.0, n[2])
        a = -1.0 / (sign + n[2])
        b = n[0] * n[1] * a
        s_ = np.array([np.copysign(1.0, n[2]) * n[0] * n[0] * a + 1.0, np.copysign(1.0, n[2]) * b, -np.copysign(1.0, n[2]) * n[0]])
        t_ = np.array([b, sign + n[1] * n[1] * a, -n[1]])
        fr = np.eye(4); fr[:3, 0], fr[:3, 1], fr[:3, 2] = s_, t_, n
        tr = np.eye(4); tr[:3, 3] = p0
        return self.add_shape("cylinder", m @ tr @ fr @ np.diag([radius, radius, 1.0, 1.0]), **kw)

    def _radiance(self, value):
        """rgb triple / scalar -> {"radiance"}; a spectrum ({"wavelengths", "values"}, spectral mode) -> + {"radiance_tex"}"""
        if isinstance(value, dict):
            # src/libcore/xml.cpp:1086-1145 (within_emitter): values are scaled by MTS_CIE_Y_NORMALIZATION; spectral variants
            # keep the spectrum, the others pre-integrate it to an (unclamped above) colour
            from .xml import spectrum_to_rgb, CIE_Y_NORMALIZATION
            vals = np.asarray(value["values"], np.float32) * np.float32(CIE_Y_NORMALIZATION)
            return {"radiance": spectrum_to_rgb(value["wavelengths"], vals, bounded=False),
                    "radiance_tex": self.add_texture_spectrum(value["wavelengths"], vals)}
        return {"radiance": [float(value)] * 3 if np.isscalar(value) else [float(x) for x in value]}

    def make_two_sided(self, bsdf_id):
        """``twosided`` adapter with the same BRDF on both sides (src/bsdfs/twosided.cpp); returns the id."""
        if self.bsdfs[bsdf_id]["type"] in (A.BSDF_DIELECTRIC, A.BSDF_ROUGHDIELECTRIC):
            raise ValueError("Only materials without a transmission component can be nested!")
        self.bsdfs[bsdf_id]["two_sided"] = 1
        return bsdf_id

    def add_constant(self, radiance):
        """``constant`` environment emitter (src/emitters/constant.cpp)."""
        self.emitters.append(dict(self._radiance(radiance), type=A.EMITTER_CONSTANT))
        return len(self.emitters) - 1

    def add_point(self, position, intensity):
        """``point`` light (src/emitters/point.cpp)."""
        self.emitters.append(dict(self._radiance(intensity), type=A.EMITTER_POINT, position=[float(x) for x in position]))
        return len(self.emitters) - 1

    def add_envmap(self, rgba, to_world=None, scale=1.0):
        rgba = _f32(rgba)
        if rgba.shape[2] == 3:
            rgba = np.concatenate([rgba, np.ones(rgba.shape[:2] + (1,), np.float32)], axis=2)
        self.emitters.append({"type": A.EMITTER_ENVMAP, "rgba": np.ascontiguousarray(rgba),
                              "to_world": _f32(np.eye(4) if to_world is None else to_world).reshape(16), "scale": float(scale)})
        return len(self.emitters) - 1

    def add_perspective(self, to_world, fov_x, near, far, width, height, rfilter="gaussian", stddev=0.5, radius=None,
                        B=1.0 / 3.0, C=1.0 / 3.0, lobes=3, crop=None):
        """`width` x `height` = the full film; `crop` = (offset_x, offset_y, crop_width, crop_height) renders only that
        window (src/librender/film.cpp:16-44): the device film then has the crop's size."""
        kinds = {"box": A.RFILTER_BOX, "tent": A.RFILTER_TENT, "gaussian": A.RFILTER_GAUSSIAN, "mitchell": A.RFILTER_MITCHELL,
                 "catmullrom": A.RFILTER_CATMULLROM, "lanczos": A.RFILTER_LANCZOS}
        if radius is None:
            radius = {"box": 0.5, "tent": 1.0, "gaussian": 4.0 * stddev, "mitchell": 2.0, "catmullrom": 2.0, "lanczos": float(lobes)}[rfilter]
        cw, ch, rel_off, rel_size = int(width), int(height), (0.0, 0.0), (1.0, 1.0)
        if crop is not None:
            ox, oy, cw, ch = (int(v) for v in crop)
            if ox < 0 or oy < 0 or cw <= 0 or ch <= 0 or ox + cw > width or oy + ch > height:
                raise ValueError("invalid crop window")
            rel_off, rel_size = (ox / width, oy / height), (cw / width, ch / height)
        self.sensors.append({"to_world": _f32(to_world).reshape(16),
                             "sample_to_camera": perspective_sample_to_camera(fov_x, near, far, width / height, rel_off, rel_size).reshape(16),
                             "near": float(near), "far": float(far), "width": cw, "height": ch,
                             "rfilter": kinds[rfilter], "radius": float(radius),
                             "stddev": float(B if rfilter == "mitchell" else stddev), "filter_c": float(C), "fov_x": float(fov_x)})
        return len(self.sensors) - 1

    # ---- C struct -------------------------------------------------------------------------
    def to_desc(self):
        """Returns (SceneDesc, keepalive) -- keepalive must outlive the call that consumes the desc."""
        keep = []
        md = (A.MeshDesc * max(1, len(self.meshes)))()
        for i, m in enumerate(self.meshes):
            keep.append(m)
            if getattr(m, "device", False):
                import torch
                torch.cuda.current_stream(m.positions.device).synchronize()
                cast = lambda t, ty: C.cast(C.c_void_p(t.data_ptr() if t is not None else None), ty)      # noqa: E731
                md[i].faces, md[i].positions = cast(m.faces, A.c_u32_p), cast(m.positions, A.c_float_p)
                md[i].normals, md[i].texcoords = cast(m.normals, A.c_float_p), cast(m.texcoords, A.c_float_p)
                md[i].mem = A.MEM_DEVICE
            else:
                md[i].faces = _ptr(m.faces, A.c_u32_p)
                md[i].positions = _ptr(m.positions, A.c_float_p)
                md[i].normals = _ptr(m.normals, A.c_float_p)
                md[i].texcoords = _ptr(m.texcoords, A.c_float_p)
            md[i].face_count, md[i].vertex_count = m.faces.shape[0], m.positions.shape[0]
            md[i].bsdf_id, md[i].emitter_id = m.bsdf, m.emitter
        bd = (A.BsdfDesc * max(1, len(self.bsdfs)))()
        for i, b in enumerate(self.bsdfs):
            bd[i].type = b["type"]
            for key in ("reflectance_tex", "eta_tex", "k_tex", "specular_reflectance_tex", "specular_transmittance_tex"):
                setattr(bd[i], key, b.get(key, -1))
            bd[i].alpha_u, bd[i].alpha_v = b.get("alpha_u", 0.1), b.get("alpha_v", 0.1)
            bd[i].distribution, bd[i].sample_visible = b.get("distribution", 0), b.get("sample_visible", 1)
            bd[i].eta = b.get("eta", 1.0)
            bd[i].two_sided = int(b.get("two_sided", 0))
            bd[i].nonlinear = int(b.get("nonlinear", 0))
            bd[i].specular_sampling_weight = float(b.get("specular_sampling_weight", 0.5))
            bd[i].opacity_tex_plus1 = int(b.get("opacity_tex", -1)) + 1
        spectral = self.mode == "spectral"
        if self.mode not in ("rgb", "spectral"):
            raise ValueError("SceneData.mode must be 'rgb' or 'spectral'")
        if spectral:
            from . import rgb2spec

            def fetch(rgb):
                return rgb2spec.fetch(rgb, model=self.rgb2spec_model)
        texs = [dict(t) if not (t["type"] == A.TEX_MESH_ATTRIBUTE and t.get("data") is None) else {"type": A.TEX_CONSTANT, "value": [0.0, 0.0, 0.0]}
                for t in self.textures]      # a mesh_attribute declared by name that no shape bound: placeholder, never referenced
        emit_tex = {}
        if spectral:
            # what the texture plugins do when they are instantiated in a spectral variant
            for t in texs:
                if t["type"] == A.TEX_CONSTANT and "spectrum" in t:          # ior.h:100-115 -> `irregular`
                    w, v = t["spectrum"]
                    t["type"], t["data"] = A.TEX_IRREGULAR, np.ascontiguousarray(np.concatenate([_f32(w), _f32(v)]))
                elif t["type"] == A.TEX_CONSTANT and not t.get("coeff"):     # srgb.cpp:28-48
                    c = np.asarray(t["value"], dtype=np.float32)
                    if np.any(c < 0) or np.any(c > 1):
                        raise ValueError(f"Invalid RGB reflectance value {c}, must be in the range [0, 1]!")
                    t["value"] = [float(x) for x in fetch(c)]
                elif t["type"] == A.TEX_MESH_ATTRIBUTE and t["data"].shape[1] == 3:
                    t["data"] = np.ascontiguousarray(fetch(np.clip(t["data"], 0.0, 1.0)).astype(np.float32))
                elif t["type"] == A.TEX_BITMAP and t["data"].shape[2] == 3:  # bitmap.cpp:120-135 (texels -> coefficients)
                    t["data"] = np.ascontiguousarray(fetch(np.clip(t["data"], 0.0, 1.0)))
            for i, e in enumerate(self.emitters):                            # srgb_d65.cpp:36-52
                if e["type"] != A.EMITTER_ENVMAP:
                    if "radiance_tex" in e:
                        emit_tex[i] = e["radiance_tex"]
                    else:
                        coeff, scale = rgb2spec.emitter_coeff_scale(np.asarray(e["radiance"], dtype=np.float32), self.rgb2spec_model)
                        texs.append({"type": A.TEX_CONSTANT, "value": [float(x) for x in coeff], "d65_scale": float(scale)})
                        emit_tex[i] = len(texs) - 1
        else:
            used = {v for b in self.bsdfs for k, v in b.items() if k.endswith("_tex") and isinstance(v, int) and v >= 0}
            for i, t in enumerate(texs):
                if t["type"] not in (A.TEX_CONSTANT, A.TEX_BITMAP, A.TEX_MESH_ATTRIBUTE) or t.get("coeff"):
                    if i in used:
                        raise ValueError("spectrum textures need mode='spectral' (rgb variants take pre-integrated colours)")
                    texs[i] = {"type": A.TEX_CONSTANT, "value": [0.0, 0.0, 0.0]}     # an emitter's spectrum: rgb mode uses its colour
        td = (A.TextureDesc * max(1, len(texs)))()
        for i, t in enumerate(texs):
            td[i].type = t["type"]
            td[i].d65_scale = t.get("d65_scale", 0.0)
            if t["type"] in (A.TEX_CONSTANT, A.TEX_UNIFORM):
                td[i].value = (C.c_float * 3)(*t["value"])
            elif t["type"] == A.TEX_MESH_ATTRIBUTE:
                d = t["data"]
                keep.append(d)
                td[i].width, td[i].channels = d.shape
                td[i].height = 1 if t["per_face"] else 0
                td[i].data = _ptr(d, A.c_float_p)
                td[i].value = (C.c_float * 3)(t["scale"], 0.0, 0.0)
            elif t["type"] == A.TEX_BITMAP:
                d = t["data"]
                keep.append(d)
                td[i].height, td[i].width, td[i].channels = d.shape
                td[i].data = _ptr(d, A.c_float_p)
                td[i].to_uv = (C.c_float * 9)(*t["to_uv"])
            else:
                d = t["data"]
                keep.append(d)
                td[i].width = d.size // (2 if t["type"] == A.TEX_IRREGULAR else 1)
                td[i].data = _ptr(d, A.c_float_p)
                td[i].lambda_min, td[i].lambda_max = t.get("lambda_min", 0.0), t.get("lambda_max", 0.0)
        ed = (A.EmitterDesc * max(1, len(self.emitters)))()
        for i, e in enumerate(self.emitters):
            ed[i].type = e["type"]
            ed[i].radiance_tex = emit_tex.get(i, -1)
            if e["type"] == A.EMITTER_AREA:
                ed[i].radiance = (C.c_float * 3)(*e["radiance"])
                ed[i].mesh_id = e["mesh_id"] if "mesh_id" in e else len(self.meshes) + e["shape_id"]
            elif e["type"] == A.EMITTER_CONSTANT:
                ed[i].radiance = (C.c_float * 3)(*e["radiance"])
                ed[i].mesh_id = -1
            elif e["type"] == A.EMITTER_POINT:
                ed[i].radiance = (C.c_float * 3)(*e["radiance"])
                ed[i].position = (C.c_float * 3)(*e["position"])
                ed[i].mesh_id = -1
            else:
                keep.append(e["rgba"])
                ed[i].height, ed[i].width = e["rgba"].shape[:2]
                ed[i].rgba = _ptr(e["rgba"], A.c_float_p)
                ed[i].to_world = (C.c_float * 16)(*e["to_world"])
                ed[i].scale = e["scale"]
                if spectral:                                                 # envmap.cpp:98-110
                    coeff, scale = rgb2spec.emitter_coeff_scale(e["rgba"][:, :, :3], self.rgb2spec_model)
                    cs = np.ascontiguousarray(np.concatenate([coeff, scale[:, :, None]], axis=2).astype(np.float32))
                    keep.append(cs)
                    ed[i].coeff_scale = _ptr(cs, A.c_float_p)
        sd = (A.SensorDesc * max(1, len(self.sensors)))()
        for i, s in enumerate(self.sensors):
            sd[i].to_world = (C.c_float * 16)(*s["to_world"])
            sd[i].sample_to_camera = (C.c_float * 16)(*s["sample_to_camera"])
            sd[i].near_clip, sd[i].far_clip = s["near"], s["far"]
            sd[i].film_width, sd[i].film_height = s["width"], s["height"]
            sd[i].rfilter, sd[i].rfilter_radius, sd[i].rfilter_stddev = s["rfilter"], s["radius"], s["stddev"]
            sd[i].rfilter_c = s.get("filter_c", 1.0 / 3.0)
        desc = A.SceneDesc()
        desc.n_meshes, desc.n_bsdfs, desc.n_textures = len(self.meshes), len(self.bsdfs), len(texs)
        desc.mode = A.MODE_SPECTRAL if spectral else A.MODE_RGB
        desc.n_emitters, desc.n_sensors = len(self.emitters), len(self.sensors)
        desc.meshes, desc.bsdfs, desc.textures, desc.emitters, desc.sensors = md, bd, td, ed, sd
        desc.integrator.max_depth, desc.integrator.rr_depth = self.max_depth, self.rr_depth
        shd = (A.ShapeDesc * max(1, len(self.shapes)))()
        for i, sh in enumerate(self.shapes):
            shd[i].type, shd[i].flip_normals = sh["type"], int(sh["flip_normals"])
            shd[i].to_world = (C.c_float * 16)(*_f32(sh["to_world"]).reshape(16))
            shd[i].bsdf_id, shd[i].emitter_id = sh["bsdf"], sh["emitter"]
        desc.n_shapes, desc.shapes = len(self.shapes), shd
        keep += [md, bd, td, ed, sd, shd]
        return desc, keep

    # ---- (de)serialisation for fixtures -------------------------------------------------------
    def save_npz(self, path):
        import json
        arrays, meta = {}, {"max_depth": self.max_depth, "rr_depth": self.rr_depth, "spp": self.spp,
                            "shapes": [dict(sh, to_world=np.asarray(sh["to_world"], np.float64).reshape(16).tolist()) for sh in self.shapes],
                            "meshes": [],
                            "bsdfs": self.bsdfs, "textures": [], "emitters": [], "sensors": []}
        for i, m in enumerate(self.meshes):
            arrays[f"m{i}_p"], arrays[f"m{i}_f"] = m.positions, m.faces
            if m.normals is not None:
                arrays[f"m{i}_n"] = m.normals
            if m.texcoords is not None:
                arrays[f"m{i}_t"] = m.texcoords
            meta["meshes"].append({"bsdf": m.bsdf, "emitter": m.emitter, "name": m.name})
        for i, t in enumerate(self.textures):
            if t["type"] == A.TEX_BITMAP:
                arrays[f"t{i}_d"], arrays[f"t{i}_uv"] = t["data"], t["to_uv"]
                meta["textures"].append({"type": t["type"]})
            else:
                meta["textures"].append(t)
        for i, e in enumerate(self.emitters):
            if e["type"] == A.EMITTER_ENVMAP:
                arrays[f"e{i}_d"], arrays[f"e{i}_w"] = e["rgba"], e["to_world"]
                meta["emitters"].append({"type": e["type"], "scale": e["scale"]})
            else:
                meta["emitters"].append(e)
        for i, s in enumerate(self.sensors):
            arrays[f"s{i}_w"], arrays[f"s{i}_c"] = s["to_world"], s["sample_to_camera"]
            meta["sensors"].append({k: v for k, v in s.items() if k not in ("to_world", "sample_to_camera")})
        arrays["meta"] = np.frombuffer(json.dumps(meta).encode(), dtype=np.uint8)
        np.savez_compressed(path, **arrays)

    @staticmethod
    def load_npz(path):
        import json
        z = np.load(path)
        meta = json.loads(bytes(z["meta"]).decode())
        sd = SceneData()
        sd.max_depth, sd.rr_depth, sd.spp = meta["max_depth"], meta["rr_depth"], meta["spp"]
        sd.bsdfs = meta["bsdfs"]
        sd.shapes = [dict(sh, to_world=_f32(sh["to_world"]).reshape(4, 4)) for sh in meta.get("shapes", [])]
        for i, m in enumerate(meta["meshes"]):
            sd.meshes.append(Mesh(z[f"m{i}_p"], z[f"m{i}_f"], z[f"m{i}_n"] if f"m{i}_n" in z else None,
                                  z[f"m{i}_t"] if f"m{i}_t" in z else None, m["bsdf"], m["emitter"], m["name"]))
        for i, t in enumerate(meta["textures"]):
            sd.textures.append({"type": t["type"], "data": _f32(z[f"t{i}_d"]), "to_uv": _f32(z[f"t{i}_uv"])} if t["type"] == A.TEX_BITMAP else t)
        for i, e in enumerate(meta["emitters"]):
            sd.emitters.append({"type": e["type"], "rgba": _f32(z[f"e{i}_d"]), "to_world": _f32(z[f"e{i}_w"]), "scale": e["scale"]}
                               if e["type"] == A.EMITTER_ENVMAP else e)
        for i, s in enumerate(meta["sensors"]):
            s = dict(s)
            s["to_world"], s["sample_to_camera"] = _f32(z[f"s{i}_w"]), _f32(z[f"s{i}_c"])
            sd.sensors.append(s)
        return sd
