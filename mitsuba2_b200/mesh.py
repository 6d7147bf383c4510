"""Mesh loading (host side): OBJ / PLY readers with the reference's import semantics, angle-weighted
vertex normals, and the synthetic benchmark meshes of SURVEY.md section 8(d).

References: src/shapes/obj.cpp:117-339 (per-(v,vt,vn) vertex de-duplication, fan triangulation, to_world baked
at parse, vt.y flipped, normals auto-computed when absent unless face_normals), src/shapes/ply.cpp,
src/librender/mesh.cpp:199-276 (recompute_vertex_normals).
"""
import struct

import numpy as np

from .scenedata import Mesh


def unit_angle(a, b):
    """enoki::unit_angle -- numerically robust angle between unit vectors (rows)."""
    dot = np.sum(a * b, axis=1)
    diff = b - np.where(dot[:, None] < 0, -a, a)
    temp = 2.0 * np.arcsin(np.minimum(1.0, 0.5 * np.linalg.norm(diff, axis=1)))
    return np.where(dot >= 0, temp, np.pi - temp)


def compute_vertex_normals(positions, faces):
    """``Mesh::recompute_vertex_normals`` (src/librender/mesh.cpp:199-276): face normals weighted by the face angle."""
    p = np.asarray(positions, dtype=np.float32).reshape(-1, 3)
    f = np.asarray(faces, dtype=np.int64).reshape(-1, 3)
    v = [p[f[:, k]] for k in range(3)]
    n = np.cross(v[1] - v[0], v[2] - v[0]).astype(np.float32)
    l2 = np.sum(n * n, axis=1)
    ok = l2 > 0
    n = np.where(ok[:, None], n / np.sqrt(np.where(ok, l2, 1.0))[:, None], 0.0)
    acc = np.zeros((p.shape[0], 3), dtype=np.float64)
    with np.errstate(invalid="ignore", divide="ignore"):
        for k in range(3):
            d0 = v[(k + 1) % 3] - v[k]
            d1 = v[(k + 2) % 3] - v[k]
            d0 = d0 / np.linalg.norm(d0, axis=1, keepdims=True)
            d1 = d1 / np.linalg.norm(d1, axis=1, keepdims=True)
            ang = np.where(ok, unit_angle(d0, d1), 0.0)
            contrib = np.nan_to_num(n * ang[:, None])
            for c in range(3):
                acc[:, c] += np.bincount(f[:, k], weights=contrib[:, c], minlength=p.shape[0])
    ln = np.linalg.norm(acc, axis=1)
    out = np.where(ln[:, None] != 0, acc / np.where(ln == 0, 1.0, ln)[:, None], np.array([1.0, 0.0, 0.0]))
    return out.astype(np.float32)


def _apply(to_world, pts, vector=False, normal=False):
    if to_world is None:
        return pts
    m = np.asarray(to_world, dtype=np.float64)
    if normal:
        r = pts.astype(np.float64) @ np.linalg.inv(m[:3, :3])      # n' = (M^-1)^T n
        return (r / np.maximum(np.linalg.norm(r, axis=1, keepdims=True), 1e-30)).astype(np.float32)
    r = pts.astype(np.float64) @ m[:3, :3].T
    if not vector:
        r = r + m[:3, 3]
    return r.astype(np.float32)


def load_obj(path, to_world=None, face_normals=False, flip_tex_coords=True, flip_normals=False):
    vs, vts, vns = [], [], []
    key_to_index, out_v, out_t, out_n, faces = {}, [], [], [], []
    with open(path, "r") as fh:
        for line in fh:
            tok = line.split()
            if not tok:
                continue
            if tok[0] == "v":
                vs.append([float(x) for x in tok[1:4]])
            elif tok[0] == "vt":
                u, v = float(tok[1]), float(tok[2]) if len(tok) > 2 else 0.0
                vts.append([u, 1.0 - v if flip_tex_coords else v])
            elif tok[0] == "vn":
                vns.append([float(x) for x in tok[1:4]])
            elif tok[0] == "f":
                idx = []
                for t in tok[1:]:
                    parts = t.split("/")
                    vi = int(parts[0])
                    ti = int(parts[1]) if len(parts) > 1 and parts[1] else 0
                    ni = int(parts[2]) if len(parts) > 2 and parts[2] else 0
                    vi = vi - 1 if vi > 0 else len(vs) + vi
                    ti = ti - 1 if ti > 0 else (len(vts) + ti if ti < 0 else -1)
                    ni = ni - 1 if ni > 0 else (len(vns) + ni if ni < 0 else -1)
                    key = (vi, ti, ni)
                    if key not in key_to_index:
                        key_to_index[key] = len(out_v)
                        out_v.append(vs[vi])
                        out_t.append(vts[ti] if ti >= 0 else None)
                        out_n.append(vns[ni] if ni >= 0 else None)
                    idx.append(key_to_index[key])
                for k in range(1, len(idx) - 1):          # fan triangulation (obj.cpp:255-264)
                    faces.append([idx[0], idx[k], idx[k + 1]])
    pos = _apply(to_world, np.asarray(out_v, dtype=np.float32).reshape(-1, 3))
    faces = np.asarray(faces, dtype=np.uint32).reshape(-1, 3)
    has_t = len(out_t) > 0 and all(t is not None for t in out_t)
    has_n = len(out_n) > 0 and all(n is not None for n in out_n)
    tex = np.asarray(out_t, dtype=np.float32) if has_t else None
    nrm = None
    if not face_normals:
        if has_n:
            nrm = _apply(to_world, np.asarray(out_n, dtype=np.float32), normal=True)
            if flip_normals:
                nrm = -nrm
        else:
            nrm = compute_vertex_normals(pos, faces)
    return Mesh(pos, faces, nrm, tex, name=str(path))


_PLY_TYPES = {"char": "i1", "uchar": "u1", "short": "i2", "ushort": "u2", "int": "i4", "uint": "u4", "float": "f4",
              "double": "f8", "int8": "i1", "uint8": "u1", "int16": "i2", "uint16": "u2", "int32": "i4", "uint32": "u4",
              "float32": "f4", "float64": "f8"}


def load_ply(path, to_world=None, face_normals=False, flip_normals=False):
    with open(path, "rb") as fh:
        assert fh.readline().strip() == b"ply"
        fmt, elements = None, []
        while True:
            line = fh.readline().decode("ascii", "replace").strip()
            tok = line.split()
            if not tok:
                continue
            if tok[0] == "format":
                fmt = tok[1]
            elif tok[0] == "element":
                elements.append({"name": tok[1], "count": int(tok[2]), "props": []})
            elif tok[0] == "property":
                if tok[1] == "list":
                    elements[-1]["props"].append(("list", tok[2], tok[3], tok[4]))
                else:
                    elements[-1]["props"].append((tok[1], tok[2]))
            elif tok[0] == "end_header":
                break
        data = {}
        endian = "<" if fmt == "binary_little_endian" else ">"
        for el in elements:
            has_list = any(p[0] == "list" for p in el["props"])
            if fmt == "ascii":
                rows = [fh.readline().split() for _ in range(el["count"])]
                if has_list:
                    data[el["name"]] = [[int(x) for x in r[1:1 + int(r[0])]] for r in rows]
                else:
                    arr = np.asarray(rows, dtype=np.float64)
                    data[el["name"]] = {p[1]: arr[:, i] for i, p in enumerate(el["props"])}
            elif not has_list:
                dt = np.dtype([(p[1], endian + _PLY_TYPES[p[0]]) for p in el["props"]])
                arr = np.frombuffer(fh.read(dt.itemsize * el["count"]), dtype=dt)
                data[el["name"]] = {p[1]: arr[p[1]] for p in el["props"]}
            else:
                assert el["props"][0][0] == "list" and all(p[0] != "list" for p in el["props"][1:]), "PLY: one index list, then scalar properties"
                _, cty, ity, _ = el["props"][0]
                extra = el["props"][1:]
                cdt, idt = np.dtype(endian + _PLY_TYPES[cty]), np.dtype(endian + _PLY_TYPES[ity])
                raw = fh.read()
                # fast path: all triangles (per-face attributes follow the index list, src/shapes/ply.cpp:262-291)
                rec = np.dtype([("c", cdt), ("i", idt, (3,))] + [(p[1], endian + _PLY_TYPES[p[0]]) for p in extra])
                if len(raw) >= rec.itemsize * el["count"]:
                    arr = np.frombuffer(raw[:rec.itemsize * el["count"]], dtype=rec)
                    if np.all(arr["c"] == 3):
                        data[el["name"]] = arr["i"].astype(np.uint32)
                        data[el["name"] + "_props"] = {p[1]: arr[p[1]] for p in extra}
                        continue
                assert not extra, "PLY: per-face attributes are supported for triangle meshes only"
                out, off = [], 0
                for _ in range(el["count"]):
                    c = int(np.frombuffer(raw, cdt, 1, off)[0]); off += cdt.itemsize
                    out.append(np.frombuffer(raw, idt, c, off).astype(np.int64).tolist()); off += idt.itemsize * c
                data[el["name"]] = out
    v = data["vertex"]
    pos = np.stack([v["x"], v["y"], v["z"]], axis=1).astype(np.float32)
    fc = data["face"]
    if isinstance(fc, np.ndarray):
        faces = fc
    else:
        tris = []
        for poly in fc:
            for k in range(1, len(poly) - 1):
                tris.append([poly[0], poly[k], poly[k + 1]])
        faces = np.asarray(tris, dtype=np.uint32)
    pos = _apply(to_world, pos)
    nrm = tex = None
    if "nx" in v and not face_normals:
        nrm = _apply(to_world, np.stack([v["nx"], v["ny"], v["nz"]], axis=1).astype(np.float32), normal=True)
        if flip_normals:
            nrm = -nrm
    elif not face_normals:
        nrm = compute_vertex_normals(pos, faces)
    for a, b in (("u", "v"), ("s", "t"), ("texture_u", "texture_v")):
        if a in v:
            tex = np.stack([v[a], v[b]], axis=1).astype(np.float32)
    m = Mesh(pos, faces, nrm, tex, name=str(path))
    m.attributes.update(_ply_attributes("vertex_", v, {"x", "y", "z", "nx", "ny", "nz", "u", "v", "s", "t", "texture_u", "texture_v"}))
    m.attributes.update(_ply_attributes("face_", data.get("face_props", {}), set()))
    return m


def _ply_attributes(kind, props, reserved):
    """``PLYMesh::find_other_fields`` (src/shapes/ply.cpp:594-690): r/g/b(/a) or red/green/blue -> "<kind>color"; other fields
    group by a common prefix with postfixes from one of [x y z w], [r g b a], [0 1 2 3], [1 2 3 4] -> "<kind><prefix>"."""
    out, names = {}, list(props)
    ren = {"red": "r", "green": "g", "blue": "b", "alpha": "a"}
    if all(k in props for k in ("red", "green", "blue")) and not all(k in props for k in "rgb"):
        props = {ren.get(k, k): val for k, val in props.items()}
        names = list(props)
    if all(k in props for k in "rgb"):
        cols = [props[k] for k in ("r", "g", "b")] + ([props["a"]] if "a" in props else [])
        out[kind + "color"] = np.stack(cols, axis=1).astype(np.float32)
    reserved = set(reserved) | {"r", "g", "b", "a"}
    groups = {}
    for nme in names:
        if nme in reserved or "_" not in nme:
            continue
        prefix, post = nme.rsplit("_", 1)
        if len(post) == 1 and post in "xyzwrgba01234":
            groups.setdefault(prefix, []).append(nme)
    for prefix, members in groups.items():
        out[kind + prefix] = np.stack([props[k] for k in members], axis=1).astype(np.float32)
    return out


def load_serialized(path, to_world=None, shape_index=0, face_normals=False, flip_normals=False):
    """``SerializedMesh`` (src/shapes/serialized.cpp:177-330): header 0x041C + version 3/4, optional offset dictionary
    at the end of the file (``shape_index``), then a zlib stream with flags, [name (v4)], uint64 vertex / face counts,
    positions, [normals], [texcoords], [colors (skipped)] in single or double precision, uint32 face indices."""
    import struct
    import zlib
    raw = open(path, "rb").read()
    fmt, version = struct.unpack_from("<hh", raw, 0)
    if fmt != 0x041C:
        raise ValueError(f'Error while loading serialized file "{path}": encountered an invalid file format!')
    if version not in (3, 4):
        raise ValueError(f'Error while loading serialized file "{path}": encountered an incompatible file version!')
    start = 4
    if shape_index != 0:
        count = struct.unpack_from("<I", raw, len(raw) - 4)[0]
        if shape_index < 0 or shape_index >= count:
            raise ValueError(f"Unable to unserialize mesh, shape index is out of range! (requested {shape_index} out of 0..{count - 1})")
        if version == 4:
            start = struct.unpack_from("<Q", raw, len(raw) - 8 * (count - shape_index) - 4)[0] + 4
        else:
            start = struct.unpack_from("<I", raw, len(raw) - 4 * (count - shape_index + 1))[0] + 4
    data = zlib.decompressobj().decompress(raw[start:])
    off = 0
    flags = struct.unpack_from("<I", data, off)[0]; off += 4
    name = str(path)
    if version == 4:
        end = data.index(b"\0", off)
        name = data[off:end].decode("utf-8", "replace") or name
        off = end + 1
    nv, nf = struct.unpack_from("<QQ", data, off); off += 16
    dt = np.dtype("<f8") if flags & 0x2000 else np.dtype("<f4")

    def take(dim):
        nonlocal off
        a = np.frombuffer(data, dt, nv * dim, off).reshape(nv, dim).astype(np.float32)
        off += nv * dim * dt.itemsize
        return a

    pos = take(3)
    nrm = take(3) if flags & 0x0001 else None
    tex = take(2) if flags & 0x0002 else None
    if flags & 0x0008:
        take(3)
    faces = np.frombuffer(data, "<u4", nf * 3, off).reshape(nf, 3).astype(np.uint32)
    pos = _apply(to_world, pos)
    if face_normals:
        nrm = None
    elif nrm is not None:
        nrm = _apply(to_world, nrm, normal=True)
        if flip_normals:
            nrm = -nrm
    else:
        nrm = compute_vertex_normals(pos, faces)
    return Mesh(pos, faces, nrm, tex, name=name)


def write_ply(path, mesh):
    """Binary little-endian PLY (so the reference could load the identical mesh)."""
    p, f = mesh.positions, mesh.faces
    props = ["x", "y", "z"] + (["nx", "ny", "nz"] if mesh.normals is not None else []) + (["u", "v"] if mesh.texcoords is not None else [])
    cols = [p] + ([mesh.normals] if mesh.normals is not None else []) + ([mesh.texcoords] if mesh.texcoords is not None else [])
    fprops, fcols = [], []
    for name, a in getattr(mesh, "attributes", {}).items():       # "<kind><prefix>" -> fields prefix_0 .. prefix_{n-1}
        kind, prefix = name.split("_", 1)
        fields = ["%s_%d" % (prefix, k) for k in range(a.shape[1])]
        if kind == "vertex":
            props += fields; cols.append(a)
        else:
            fprops += fields; fcols.append(a)
    vert = np.concatenate(cols, axis=1).astype("<f4")
    with open(path, "wb") as fh:
        hdr = "ply\nformat binary_little_endian 1.0\nelement vertex %d\n" % p.shape[0]
        hdr += "".join("property float %s\n" % q for q in props)
        hdr += "element face %d\nproperty list uchar int vertex_indices\n" % f.shape[0]
        hdr += "".join("property float %s\n" % q for q in fprops) + "end_header\n"
        fh.write(hdr.encode())
        fh.write(vert.tobytes())
        rec = np.empty(f.shape[0], dtype=[("c", "u1"), ("i", "<i4", (3,))] + [(q, "<f4") for q in fprops])
        rec["c"], rec["i"] = 3, f
        if fcols:
            fa = np.concatenate(fcols, axis=1)
            for k, q in enumerate(fprops):
                rec[q] = fa[:, k]
        fh.write(rec.tobytes())


# ---- synthetic meshes ----------------------------------------------------------------
def _pcg32_stream(n, initstate, initseq):
    """n PCG32 float32 draws (enoki/random.h) -- numpy, vectorised by jumping is not needed: sequential generator in
    64-bit integer arithmetic, chunked for speed."""
    mult = np.uint64(0x5851f42d4c957f2d)
    inc = np.uint64((int(initseq) << 1 | 1) & 0xFFFFFFFFFFFFFFFF)
    mask = 0xFFFFFFFFFFFFFFFF
    state = 0
    state = (state * int(mult) + int(inc)) & mask
    state = (state + int(initstate)) & mask
    state = (state * int(mult) + int(inc)) & mask
    # LCG skip-ahead: state_k = A_k*state + C_k ; build per-index coefficients by doubling
    idx = np.arange(n, dtype=np.uint64)
    acc_mult = np.ones(n, dtype=np.uint64)
    acc_plus = np.zeros(n, dtype=np.uint64)
    cur_mult, cur_plus = mult, inc
    with np.errstate(over="ignore"):
        bit = 0
        while (n - 1) >> bit:
            sel = ((idx >> np.uint64(bit)) & np.uint64(1)).astype(bool)
            acc_mult = np.where(sel, acc_mult * cur_mult, acc_mult)
            acc_plus = np.where(sel, acc_plus * cur_mult + cur_plus, acc_plus)
            cur_plus = (cur_mult + np.uint64(1)) * cur_plus
            cur_mult = cur_mult * cur_mult
            bit += 1
        old = acc_mult * np.uint64(state) + acc_plus
        xs = (((old >> np.uint64(18)) ^ old) >> np.uint64(27)).astype(np.uint32)
        rot = (old >> np.uint64(59)).astype(np.uint32)
        out = (xs >> rot) | (xs << ((np.uint32(32) - rot) & np.uint32(31)))
    return ((out >> np.uint32(9)) | np.uint32(0x3f800000)).view(np.float32) - np.float32(1.0)


def displaced_sphere(n_theta, n_phi, radius=1.0, center=(0, 0, 0), displace=True, seed=0x5EED):
    """UV sphere of n_theta x n_phi quads (2 triangles each), r = 1 + 0.05 sin(8 theta) sin(8 phi) + 0.01 xi,
    xi in [-1,1) from PCG32(initstate=seed, initseq=1), one draw per vertex in row-major order; uv = (phi/2pi, theta/pi)."""
    nt, nphi = n_theta + 1, n_phi + 1
    theta = np.linspace(0.0, np.pi, nt, dtype=np.float64)[:, None]
    phi = np.linspace(0.0, 2.0 * np.pi, nphi, dtype=np.float64)[None, :]
    r = np.ones((nt, nphi))
    if displace:
        xi = _pcg32_stream(nt * nphi, seed, 1).astype(np.float64).reshape(nt, nphi) * 2.0 - 1.0
        xi[:, -1] = xi[:, 0]                       # keep the seam closed
        xi[0, :], xi[-1, :] = xi[0, 0], xi[-1, 0]  # and the poles
        r = 1.0 + 0.05 * np.sin(8.0 * theta) * np.sin(8.0 * phi) + 0.01 * xi
    r = r * radius
    x = r * np.sin(theta) * np.cos(phi)
    z = r * np.sin(theta) * np.sin(phi)
    y = r * np.cos(theta) * np.ones_like(phi)
    pos = np.stack([x, y, z], axis=2).reshape(-1, 3) + np.asarray(center, dtype=np.float64)
    uv = np.stack([np.broadcast_to(phi / (2 * np.pi), (nt, nphi)), np.broadcast_to(theta / np.pi, (nt, nphi))], axis=2).reshape(-1, 2)
    i, j = np.meshgrid(np.arange(n_theta), np.arange(n_phi), indexing="ij")
    a = (i * nphi + j).ravel(); b = a + 1; c = a + nphi; d = c + 1
    faces = np.concatenate([np.stack([a, b, c], axis=1), np.stack([b, d, c], axis=1)], axis=0).astype(np.uint32)   # outward normals
    pos = pos.astype(np.float32)
    return Mesh(pos, faces, compute_vertex_normals(pos, faces), uv.astype(np.float32), name="displaced_sphere")


def quad(p0, p1, p2, p3, uv=True):
    pos = np.asarray([p0, p1, p2, p3], dtype=np.float32)
    faces = np.asarray([[0, 1, 2], [0, 2, 3]], dtype=np.uint32)
    tex = np.asarray([[0, 0], [1, 0], [1, 1], [0, 1]], dtype=np.float32) if uv else None
    return Mesh(pos, faces, None, tex, name="quad")
