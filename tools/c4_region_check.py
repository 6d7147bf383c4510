#!/usr/bin/env python
"""Where on the film does the vertex adjoint of the C4 scene disagree with finite differences?  (GPU; companion of
tools/c4_fd_check.py.)  The loss is the mean of the image over a MASK -- pixels well inside the sphere's silhouette, a band
around the silhouette, and everything else (floor / walls with the sphere's shadow and its colour bleeding) -- so each of the
adjoint's terms (interior, film boundary, boundary seen from x1) is tested where it is the only thing that matters.
Prints one line per mask and boundary mode: d loss / d field (adjoint +- standard error over the passes) / FD."""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def dilate(m, k):
    out = m.copy()
    for _ in range(k):
        p = np.pad(out, 1)
        out = p[1:-1, 1:-1] | p[:-2, 1:-1] | p[2:, 1:-1] | p[1:-1, :-2] | p[1:-1, 2:]
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--res", type=int, default=512); ap.add_argument("--spp", type=int, default=64); ap.add_argument("--passes", type=int, default=16)
    ap.add_argument("--fd-spp", type=int, default=2048); ap.add_argument("--max-depth", type=int, default=2); ap.add_argument("--convex", action="store_true")
    ap.add_argument("--band", type=int, default=4); ap.add_argument("--modes", default="0,1,2"); ap.add_argument("--edge-fraction", type=float, default=0.25)
    ap.add_argument("--variant", default="full"); ap.add_argument("--flat", action="store_true")
    ap.add_argument("--eps", type=float, default=0.125, help="finite-difference step in mean edge lengths"); ap.add_argument("--richardson", type=int, default=1)
    a = ap.parse_args()
    import mitsuba2_b200 as mi
    from mitsuba2_b200 import scenes
    mi.set_variant("cuda_b200_rgb")
    sd, tex = scenes.c4_scene(os.path.join(ROOT, "tests", "golden", "cbox_scene.npz"), res=a.res, spp=a.spp, displace=not a.convex, smooth=not a.flat)
    sd.max_depth = a.max_depth
    mid = next(i for i, m in enumerate(sd.meshes) if m.name == "c4_sphere")
    if a.variant == "black_sphere":
        sd.meshes[mid].bsdf = sd.add_diffuse(0.0)
    elif a.variant == "black_room":
        blk = sd.add_diffuse(0.0)
        for i, m in enumerate(sd.meshes):
            if i != mid and m.emitter < 0:
                m.bsdf = blk
    sc = mi.Scene(sd, device=0)
    P0 = np.asarray(sd.meshes[mid].positions, np.float32).reshape(-1, 3).copy()
    c = P0.mean(axis=0)
    radius = float(np.linalg.norm(P0 - c, axis=1).mean())
    f = np.asarray(sd.meshes[mid].faces)
    edge = float(np.linalg.norm(P0[f[:, 0]] - P0[f[:, 1]], axis=1).mean())
    # coverage of the sphere: primary rays through the pixel centres
    yy, xx = np.mgrid[0:a.res, 0:a.res]
    pos = np.stack([(xx.ravel() + 0.5) / a.res, (yy.ravel() + 0.5) / a.res], axis=1).astype(np.float32)
    o, d = sc.sensor_sample_ray(pos)
    hits = sc.intersect_closest_raw(o, d)
    prim = hits[:, 3].copy().view(np.uint32).astype(np.int64)
    lo = int(sc.primitive_map[mid]); hi = int(sc.primitive_map[mid + 1])
    cover = ((prim >= lo) & (prim < hi) & np.isfinite(hits[:, 0])).reshape(a.res, a.res)
    inside = ~dilate(~cover, a.band)
    band = dilate(cover, a.band) & ~inside
    outside = ~dilate(cover, a.band)
    masks = {"inside_sphere": inside, "silhouette_band": band, "rest_of_room": outside}
    fields = {"translate_x": np.tile([1.0, 0, 0], (len(P0), 1)), "translate_y": np.tile([0, 1.0, 0], (len(P0), 1)), "scale": (P0 - c) / radius}
    # finite differences: one pair of renders per field, seed and epsilon serves all masks; central differences have an
    # O(eps^2) error that is NOT small here (the silhouette rim of a smooth-shaded mesh is dark, shadows sweep a noise texture),
    # so two epsilons are Richardson-extrapolated: fd(0) = fd(e) + (fd(e) - fd(2e)) / 3
    fd = {n: {m: [] for m in masks} for n in fields}
    for n, D in fields.items():
        for s_ in (7, 8):
            per_eps = {}
            for e in (a.eps, 2 * a.eps):
                eps = e * edge
                sc.update_vertices(mid, (P0 + eps * D).astype(np.float32)); ip = sc.render(spp=a.fd_spp, seed=s_).astype(np.float64)
                sc.update_vertices(mid, (P0 - eps * D).astype(np.float32)); im = sc.render(spp=a.fd_spp, seed=s_).astype(np.float64)
                per_eps[e] = {m: float(((ip - im) * mk[..., None]).sum() / (2 * eps) / ip.size) for m, mk in masks.items()}
            for m in masks:
                f1, f2 = per_eps[a.eps][m], per_eps[2 * a.eps][m]
                fd[n][m].append(f1 + (f1 - f2) / 3.0 if a.richardson else f1)
    sc.update_vertices(mid, P0)
    out = {}
    for mode in [int(x) for x in a.modes.split(",")]:
        sc.grad_config(boundary=mode, edge_fraction=a.edge_fraction)
        for m, mk in masks.items():
            dl = (np.repeat(mk[..., None], 3, axis=2) / float(mk.size * 3)).astype(np.float32)
            per = []
            for k in range(a.passes):
                _, gv = sc.render_grad(dl, spp=a.spp, seed=100 + k, mesh_id=mid)
                per.append(np.asarray(gv, np.float64).reshape(-1, 3))
            row = {}
            for n, D in fields.items():
                v = np.array([float((g * D).sum()) for g in per])
                row[n] = {"adjoint": float(v.mean()), "se": float(v.std(ddof=1) / np.sqrt(len(v))), "fd": float(np.mean(fd[n][m])), "fd_sd": float(np.std(fd[n][m]))}
            out[f"b{mode}_{m}"] = row
            print(f"b{mode} {m:16s} " + "  ".join(f"{n}: {r['adjoint'] * 1e6:+.3f}+-{r['se'] * 1e6:.3f} / fd {r['fd'] * 1e6:+.3f}+-{r['fd_sd'] * 1e6:.3f}" for n, r in row.items()), flush=True)
    print(json.dumps({"config": vars(a), "pixels": {m: int(mk.sum()) for m, mk in masks.items()}, "results": out}))


if __name__ == "__main__":
    main()
