"""Write a :class:`SceneData` as a Mitsuba 2 scene directory (scene.xml + binary PLY meshes + PFM bitmaps) that the
reference's own loader reads (`xml::load_file`, src/libcore/xml.cpp; the reference's Python-side writer of the same
format is src/python/python/xml.py).  This is how the synthetic benchmark scenes reach the reference `mitsuba` binary
(CPU baseline, statistical image parity) -- the product itself never needs it.

Order matters: shapes are written in mesh order (global primitive index = prefix sum over `Scene::m_shapes`,
kdtree.h:2382-2400) and non-area emitters are interleaved so that `Scene::m_emitters` (scene.cpp:23-48, document order of
unnamed children) equals `SceneData.emitters`.
"""
import os

import numpy as np

from . import _abi as A
from .imageio import write_pfm
from .mesh import write_ply

_FILTERS = {A.RFILTER_BOX: "box", A.RFILTER_TENT: "tent", A.RFILTER_GAUSSIAN: "gaussian", A.RFILTER_MITCHELL: "mitchell",
            A.RFILTER_CATMULLROM: "catmullrom", A.RFILTER_LANCZOS: "lanczos"}


def _vec(v):
    return ", ".join(repr(float(x)) for x in v)


def _matrix(m):
    return " ".join(repr(float(x)) for x in np.asarray(m, np.float64).reshape(16))


class _Writer:
    def __init__(self, sd, directory):
        self.sd, self.dir, self.lines, self.ind = sd, directory, [], 1

    def w(self, s):
        self.lines.append("    " * self.ind + s)

    def texture(self, name, tex_id, emitter=False):
        t = self.sd.textures[tex_id]
        if t["type"] == A.TEX_CONSTANT:
            v = t["value"]
            if all(0.0 <= x <= 1.0 for x in v) or emitter or name in ("eta", "k"):      # xml.cpp:81-86 is_unbounded_spectrum
                self.w(f'<rgb name="{name}" value="{_vec(v)}"/>')
            elif v[0] == v[1] == v[2]:
                self.w(f'<spectrum name="{name}" value="{float(v[0])!r}"/>')
            else:
                raise ValueError(f"colour {v} outside [0,1] cannot be written as a reflectance")
        elif t["type"] == A.TEX_BITMAP:
            fn = f"tex_{tex_id}.pfm"
            path = os.path.join(self.dir, fn)
            if not os.path.exists(path):
                d = t["data"]
                write_pfm(path, d if d.shape[2] == 3 else d[:, :, 0])
            self.w(f'<texture type="bitmap" name="{name}">')
            self.w(f'    <string name="filename" value="{fn}"/>')
            uv = np.asarray(t["to_uv"], np.float64).reshape(3, 3)
            if not np.allclose(uv, np.eye(3)):
                m = np.eye(4)
                m[:2, :2], m[:2, 3] = uv[:2, :2], uv[:2, 2]
                self.w(f'    <transform name="to_uv"><matrix value="{_matrix(m)}"/></transform>')
            self.w("</texture>")
        elif t["type"] == A.TEX_MESH_ATTRIBUTE:
            self.w(f'<texture type="mesh_attribute" name="{name}"><string name="name" value="{t["name"]}"/>'
                   f'<float name="scale" value="{t["scale"]!r}"/></texture>')
        else:
            raise ValueError("tabulated spectra are not exported")

    def bsdf(self, b):
        ty = b["type"]
        two = bool(b.get("two_sided", 0))
        if b.get("opacity_tex", -1) >= 0:
            raise ValueError("mask adapter not exported")
        if two:
            self.w('<bsdf type="twosided">')
            self.ind += 1
        if ty == A.BSDF_DIFFUSE:
            self.w('<bsdf type="diffuse">')
            self.ind += 1
            self.texture("reflectance", b["reflectance_tex"])
        elif ty in (A.BSDF_ROUGHCONDUCTOR, A.BSDF_CONDUCTOR):
            self.w(f'<bsdf type="{"roughconductor" if ty == A.BSDF_ROUGHCONDUCTOR else "conductor"}">')
            self.ind += 1
            mat = self.sd.textures[b["eta_tex"]].get("material")
            if mat:
                self.w(f'<string name="material" value="{mat}"/>')
            else:
                self.texture("eta", b["eta_tex"])
                self.texture("k", b["k_tex"])
            if ty == A.BSDF_ROUGHCONDUCTOR:
                self.w(f'<string name="distribution" value="{"ggx" if b.get("distribution", 0) == A.MF_GGX else "beckmann"}"/>')
                if b["alpha_u"] == b["alpha_v"]:
                    self.w(f'<float name="alpha" value="{b["alpha_u"]!r}"/>')
                else:
                    self.w(f'<float name="alpha_u" value="{b["alpha_u"]!r}"/>')
                    self.w(f'<float name="alpha_v" value="{b["alpha_v"]!r}"/>')
                self.w(f'<boolean name="sample_visible" value="{"true" if b.get("sample_visible", 1) else "false"}"/>')
            if b.get("specular_reflectance_tex", -1) >= 0:
                self.texture("specular_reflectance", b["specular_reflectance_tex"])
        elif ty in (A.BSDF_DIELECTRIC, A.BSDF_ROUGHDIELECTRIC):
            self.w(f'<bsdf type="{"dielectric" if ty == A.BSDF_DIELECTRIC else "roughdielectric"}">')
            self.ind += 1
            self.w(f'<float name="int_ior" value="{float(np.float32(b["eta"]))!r}"/>')     # only the ratio matters (dielectric.cpp:170-176)
            self.w('<float name="ext_ior" value="1.0"/>')
            if ty == A.BSDF_ROUGHDIELECTRIC:
                self.w(f'<string name="distribution" value="{"ggx" if b.get("distribution", 0) == A.MF_GGX else "beckmann"}"/>')
                self.w(f'<float name="alpha" value="{b["alpha_u"]!r}"/>')
                self.w(f'<boolean name="sample_visible" value="{"true" if b.get("sample_visible", 1) else "false"}"/>')
            for key in ("specular_reflectance", "specular_transmittance"):
                if b.get(key + "_tex", -1) >= 0:
                    self.texture(key, b[key + "_tex"])
        elif ty in (A.BSDF_PLASTIC, A.BSDF_ROUGHPLASTIC):
            self.w(f'<bsdf type="{"plastic" if ty == A.BSDF_PLASTIC else "roughplastic"}">')
            self.ind += 1
            self.w(f'<float name="int_ior" value="{float(np.float32(b["eta"]))!r}"/>')
            self.w('<float name="ext_ior" value="1.0"/>')
            self.texture("diffuse_reflectance", b["reflectance_tex"])
            self.w(f'<boolean name="nonlinear" value="{"true" if b.get("nonlinear", 0) else "false"}"/>')
            if ty == A.BSDF_ROUGHPLASTIC:
                self.w(f'<string name="distribution" value="{"ggx" if b.get("distribution", 0) == A.MF_GGX else "beckmann"}"/>')
                self.w(f'<float name="alpha" value="{b["alpha_u"]!r}"/>')
                self.w(f'<boolean name="sample_visible" value="{"true" if b.get("sample_visible", 1) else "false"}"/>')
            if b.get("specular_reflectance_tex", -1) >= 0:
                self.texture("specular_reflectance", b["specular_reflectance_tex"])
        else:
            raise ValueError(f"BSDF type {ty} not exported")
        self.ind -= 1
        self.w("</bsdf>")
        if two:
            self.ind -= 1
            self.w("</bsdf>")

    def emitter(self, e, i):
        ty = e["type"]
        if ty == A.EMITTER_ENVMAP:
            fn = f"envmap_{i}.pfm"
            write_pfm(os.path.join(self.dir, fn), e["rgba"][:, :, :3])
            self.w('<emitter type="envmap">')
            self.w(f'    <string name="filename" value="{fn}"/>')
            self.w(f'    <float name="scale" value="{e["scale"]!r}"/>')
            self.w(f'    <transform name="to_world"><matrix value="{_matrix(e["to_world"])}"/></transform>')
            self.w("</emitter>")
        elif ty == A.EMITTER_CONSTANT:
            self.w(f'<emitter type="constant"><rgb name="radiance" value="{_vec(e["radiance"])}"/></emitter>')
        elif ty == A.EMITTER_POINT:
            self.w(f'<emitter type="point"><rgb name="intensity" value="{_vec(e["radiance"])}"/>'
                   f'<point name="position" x="{e["position"][0]!r}" y="{e["position"][1]!r}" z="{e["position"][2]!r}"/></emitter>')
        else:
            raise ValueError("area emitters are written with their shape")


def write_mitsuba_scene(sd, directory, spp=None, sensor=0, file_format="pfm", variant_hint="scalar_rgb", integrator="path"):
    """-> path of scene.xml.  Meshes go to mesh_<i>.ply (normals and texcoords included when the mesh has them; a mesh
    without normals is loaded with face_normals=true, which is what `normals is None` means on the device)."""
    os.makedirs(directory, exist_ok=True)
    w = _Writer(sd, directory)
    spp = int(sd.spp if spp is None else spp)
    w.lines.append(f'<!-- written by mitsuba2_b200.export for {variant_hint} -->')
    w.lines.append('<scene version="2.1.0">')
    w.w(f'<integrator type="{integrator}"><integer name="max_depth" value="{int(sd.max_depth)}"/>'
        f'<integer name="rr_depth" value="{int(sd.rr_depth)}"/></integrator>')
    s = sd.sensors[sensor]
    w.w('<sensor type="perspective">')
    w.ind += 1
    w.w('<string name="fov_axis" value="x"/>')
    w.w(f'<float name="fov" value="{float(s["fov_x"])!r}"/>')
    w.w(f'<float name="near_clip" value="{float(s["near"])!r}"/>')
    w.w(f'<float name="far_clip" value="{float(s["far"])!r}"/>')
    w.w(f'<transform name="to_world"><matrix value="{_matrix(s["to_world"])}"/></transform>')
    w.w(f'<sampler type="independent"><integer name="sample_count" value="{spp}"/></sampler>')
    w.w('<film type="hdrfilm">')
    w.ind += 1
    w.w(f'<integer name="width" value="{int(s["width"])}"/>')
    w.w(f'<integer name="height" value="{int(s["height"])}"/>')
    w.w(f'<string name="file_format" value="{file_format}"/>')
    w.w('<string name="pixel_format" value="rgb"/>')
    w.w('<string name="component_format" value="float32"/>')
    f = _FILTERS[s["rfilter"]]
    if f == "gaussian":
        w.w(f'<rfilter type="gaussian"><float name="stddev" value="{float(s["stddev"])!r}"/></rfilter>')
    elif f == "mitchell":
        w.w(f'<rfilter type="mitchell"><float name="B" value="{float(s["stddev"])!r}"/><float name="C" value="{float(s.get("filter_c", 1 / 3))!r}"/></rfilter>')
    elif f == "lanczos":
        w.w(f'<rfilter type="lanczos"><integer name="lobes" value="{int(round(s["radius"]))}"/></rfilter>')
    else:
        w.w(f'<rfilter type="{f}"/>')
    w.ind -= 1
    w.w("</film>")
    w.ind -= 1
    w.w("</sensor>")
    # shapes in mesh order, other emitters interleaved to keep the emitter order
    pending = [i for i, e in enumerate(sd.emitters) if e["type"] != A.EMITTER_AREA]
    for mi_, m in enumerate(sd.meshes):
        if m.emitter >= 0:
            while pending and pending[0] < m.emitter:
                k = pending.pop(0)
                w.emitter(sd.emitters[k], k)
        fn = f"mesh_{mi_}.ply"
        write_ply(os.path.join(directory, fn), m)
        w.w('<shape type="ply">')
        w.ind += 1
        w.w(f'<string name="filename" value="{fn}"/>')
        if m.normals is None:
            w.w('<boolean name="face_normals" value="true"/>')
        w.bsdf(sd.bsdfs[m.bsdf])
        if m.emitter >= 0:
            e = sd.emitters[m.emitter]
            w.w(f'<emitter type="area"><rgb name="radiance" value="{_vec(e["radiance"])}"/></emitter>')
        w.ind -= 1
        w.w("</shape>")
    # analytic shapes after the meshes (global primitive index = mesh faces, then one primitive per shape); to_world already
    # carries center / radius / p0 / p1, so the plugins' own parameters stay at their defaults
    names = {A.SHAPE_SPHERE: "sphere", A.SHAPE_DISK: "disk", A.SHAPE_CYLINDER: "cylinder", A.SHAPE_RECTANGLE: "rectangle"}
    for sh in getattr(sd, "shapes", []):
        if sh["emitter"] >= 0:
            while pending and pending[0] < sh["emitter"]:
                k = pending.pop(0)
                w.emitter(sd.emitters[k], k)
        w.w(f'<shape type="{names[sh["type"]]}">')
        w.ind += 1
        w.w(f'<transform name="to_world"><matrix value="{_matrix(sh["to_world"])}"/></transform>')
        if sh["flip_normals"]:
            w.w('<boolean name="flip_normals" value="true"/>')
        w.bsdf(sd.bsdfs[sh["bsdf"]])
        if sh["emitter"] >= 0:
            e = sd.emitters[sh["emitter"]]
            w.w(f'<emitter type="area"><rgb name="radiance" value="{_vec(e["radiance"])}"/></emitter>')
        w.ind -= 1
        w.w("</shape>")
    for k in pending:
        w.emitter(sd.emitters[k], k)
    w.lines.append("</scene>")
    path = os.path.join(directory, "scene.xml")
    with open(path, "w") as fh:
        fh.write("\n".join(w.lines) + "\n")
    return path
