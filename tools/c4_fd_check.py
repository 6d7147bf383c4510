#!/usr/bin/env python
"""SURVEY 8(d) protocol for the gradient pass on the C4 configuration (run on a GPU):
cbox with a 512x512 albedo texture on the floor and a 20k-triangle displaced sphere, 512 x 512, L2 loss against a target
rendered from a perturbed texture / mesh.  The adjoint (mtsb_render_grad, 64 spp per pass, averaged over `--passes` seeds)
is compared with central finite differences of the renderer itself at `--fd-spp` spp and identical seeds on 64 random
texel channels and 32 random vertex coordinates (plus 3 global displacement fields), with an epsilon sweep for the vertices.
The primal renderer used for the differences is the one whose images are pinned against the oracle / the reference
(tests/test_gpu_bench_parity.py); the CPU oracle cannot render 2 x 96 x 512^2 x 4096 samples in bounded time.
Prints one JSON object (also used by bench.py's parity gates with smaller settings)."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def run(res=512, spp=64, passes=8, fd_spp=4096, n_tex=64, n_vtx=32, eps_list=(0.25, 0.5, 1.0), device=0, seed=11, tex_res=512, log=None,
        max_depth=None, boundary=2, edge_fraction=0.25, rr_off=False, same_samples=False, variant="full", displace=True, smooth=True, sphere=(100, 100)):
    import mitsuba2_b200 as mi
    from mitsuba2_b200 import scenes
    mi.set_variant("cuda_b200_rgb")
    sd, tex = scenes.c4_scene(os.path.join(ROOT, "tests", "golden", "cbox_scene.npz"), res=res, spp=spp, tex_res=tex_res, displace=displace, smooth=smooth, sphere=sphere)
    mid = next(i for i, m in enumerate(sd.meshes) if m.name == "c4_sphere")
    if variant == "black_sphere":        # only what the sphere does to the rest of the room (its shadow, its silhouette) is left
        sd.meshes[mid].bsdf = sd.add_diffuse(0.0)
    elif variant == "black_room":        # only the sphere's own shading (and its silhouette against black) is left
        blk = sd.add_diffuse(0.0)
        for i, m in enumerate(sd.meshes):
            if i != mid and m.emitter < 0:
                m.bsdf = blk
    if max_depth is not None:
        sd.max_depth = max_depth
    if rr_off:
        sd.rr_depth = 1000        # Russian roulette makes same-seed differences jump when a survival test flips
    sc = mi.Scene(sd, device=device)
    sc.grad_config(boundary=boundary, edge_fraction=edge_fraction)
    rng = np.random.default_rng(seed)
    T0 = sd.textures[tex]["data"].copy()
    P0 = np.asarray(sd.meshes[mid].positions, np.float32).reshape(-1, 3).copy()
    c = P0.mean(axis=0)
    radius = float(np.linalg.norm(P0 - c, axis=1).mean())
    # target: smoother, darker texture and a sphere scaled by 1.05 and shifted
    Tt = np.clip(0.8 * T0 + 0.05, 0, 1).astype(np.float32)
    Pt = (c + 1.05 * (P0 - c) + radius * np.array([0.05, 0.03, -0.04], np.float32)).astype(np.float32)
    sc.update_texture(tex, Tt); sc.update_vertices(mid, Pt)
    target = sc.render(spp=max(256, fd_spp // 4), seed=999)
    sc.update_texture(tex, T0); sc.update_vertices(mid, P0)
    cur = sc.render(spp=max(256, fd_spp // 4), seed=998)
    dl = (2.0 * (cur - target) / cur.size).astype(np.float32)
    out = {"config": {"variant": variant, "max_depth": sd.max_depth, "rr_depth": sd.rr_depth, "boundary": boundary, "edge_fraction": edge_fraction, "same_samples": same_samples, "res": res, "adjoint_spp": spp, "adjoint_passes": passes, "fd_spp": fd_spp, "texels": n_tex, "vertex_coords": n_vtx,
                      "sphere_triangles": int(len(sd.meshes[mid].faces)), "loss": float(np.mean((cur - target) ** 2))}}
    # ---- adjoint
    t0 = time.time()
    gt = np.zeros(T0.size, np.float64); gv = np.zeros(P0.shape, np.float64)
    if same_samples:      # the adjoint on exactly the sample set of the differences (seed 7, fd_spp): continuous parameters then agree sample by sample
        passes, spp = 1, fd_spp
    per_pass = []
    for k in range(passes):
        a, b = sc.render_grad(dl, spp=spp, seed=7 if same_samples else 100 + k, tex_id=tex, mesh_id=mid)
        gt += np.asarray(a, np.float64) / passes; gv += np.asarray(b, np.float64) / passes
        per_pass.append(np.asarray(b, np.float64))
    gt = gt.reshape(T0.shape)
    out["adjoint_seconds"] = time.time() - t0

    def loss(seed_):
        return float(np.sum(np.float64(dl) * sc.render(spp=fd_spp, seed=seed_)))

    # ---- texture: the image is linear in a texel up to the bilinear footprint, one epsilon suffices
    h, w, _ = T0.shape
    # the film is summed with float atomics, so two renders of the same scene differ by ~1e-7 per pixel and a single texel's
    # difference (1e-10 .. 1e-9 here) drowns in that: perturb BLOCKS of 8 x 8 texels (signal x 64) chosen where the adjoint
    # has weight (the floor region the camera and the first bounces see) and uniformly
    B = 8
    gb = gt[: h // B * B, : w // B * B].reshape(h // B, B, w // B, B, -1).sum(axis=(1, 3))        # adjoint of a block = sum over its texels
    flat = np.abs(gb).reshape(-1) + 1e-300
    pick = np.concatenate([rng.choice(flat.size, n_tex // 2, replace=False, p=flat / flat.sum()), rng.choice(flat.size, n_tex - n_tex // 2, replace=False)])
    fd_t, an_t = [], []
    for k, idx in enumerate(pick):
        by, bx, ch = np.unravel_index(idx, gb.shape)
        T = T0.copy(); T[by * B:(by + 1) * B, bx * B:(bx + 1) * B, ch] += 0.1; sc.update_texture(tex, T); lp = loss(7)
        T[by * B:(by + 1) * B, bx * B:(bx + 1) * B, ch] -= 0.2; sc.update_texture(tex, T); lm = loss(7)
        fd_t.append((lp - lm) / 0.2); an_t.append(gb[by, bx, ch])
        if log: log(f"texel block {k}: fd {fd_t[-1]:+.4e} adjoint {an_t[-1]:+.4e}")
    sc.update_texture(tex, T0)
    fd_t, an_t = np.array(fd_t), np.array(an_t)
    out["texture"] = {"rel_l2": float(np.linalg.norm(fd_t - an_t) / np.linalg.norm(fd_t)), "cosine": float(fd_t @ an_t / (np.linalg.norm(fd_t) * np.linalg.norm(an_t)))}
    # ---- vertices: global fields (translate x / y, scale) and single coordinates; epsilon sweep (fraction of the mean edge length)
    f = np.asarray(sd.meshes[mid].faces)
    edge = float(np.linalg.norm(P0[f[:, 0]] - P0[f[:, 1]], axis=1).mean())
    fields = {"translate_x": np.tile([1.0, 0, 0], (len(P0), 1)), "translate_y": np.tile([0, 1.0, 0], (len(P0), 1)), "scale": (P0 - c) / radius}
    res_v = {}
    for name, D in fields.items():
        an = float((gv * D).sum())
        pp = np.array([float((g * D).sum()) for g in per_pass])
        an_se = float(pp.std(ddof=1) / np.sqrt(len(pp))) if len(pp) > 1 else float("nan")
        sweep = {}
        for e in eps_list:
            eps = e * edge
            vals = []
            for s_ in (7, 8):
                sc.update_vertices(mid, (P0 + eps * D).astype(np.float32)); lp = loss(s_)
                sc.update_vertices(mid, (P0 - eps * D).astype(np.float32)); lm = loss(s_)
                vals.append((lp - lm) / (2 * eps))
            sweep[str(e)] = [float(np.mean(vals)), float(np.std(vals))]
        res_v[name] = {"adjoint": an, "adjoint_std_error": an_se, "fd_by_eps": sweep}
        # central differences carry an O(eps^2) error that is not small here (measured: the full scene's d/dx changes by 60 %
        # between eps = 0.25 and 1 edge lengths): Richardson-extrapolate the two smallest steps when they are a factor 2 apart
        es = sorted(eps_list)
        if len(es) >= 2 and abs(es[1] / es[0] - 2.0) < 1e-6:
            f1, f2 = sweep[str(es[0])][0], sweep[str(es[1])][0]
            res_v[name]["fd_richardson"] = f1 + (f1 - f2) / 3.0
        if log: log(f"{name}: adjoint {an:+.4e} fd {sweep}")
    out["vertex_fields"] = res_v
    e_mid = eps_list[len(eps_list) // 2]
    fd_f = np.array([res_v[n]["fd_by_eps"][str(e_mid)][0] for n in fields]); an_f = np.array([res_v[n]["adjoint"] for n in fields])
    out["vertex_fields_rel_l2"] = float(np.linalg.norm(fd_f - an_f) / np.linalg.norm(fd_f))
    if all("fd_richardson" in res_v[n] for n in fields):
        fd_r = np.array([res_v[n]["fd_richardson"] for n in fields])
        out["vertex_fields_rel_l2_richardson"] = float(np.linalg.norm(fd_r - an_f) / np.linalg.norm(fd_r))
    if n_vtx > 0:
        # single coordinates of vertices that matter (|adjoint| weighted) and uniform ones
        wv = np.abs(gv).reshape(-1)
        pickv = np.concatenate([rng.choice(wv.size, n_vtx // 2, replace=False, p=wv / wv.sum()), rng.choice(wv.size, n_vtx - n_vtx // 2, replace=False)])
        fd_v, an_v = [], []
        eps = 1.0 * edge
        for k, idx in enumerate(pickv):
            vi, ax = divmod(int(idx), 3)
            P = P0.copy(); P[vi, ax] += eps; sc.update_vertices(mid, P); lp = loss(7)
            P[vi, ax] -= 2 * eps; sc.update_vertices(mid, P); lm = loss(7)
            fd_v.append((lp - lm) / (2 * eps)); an_v.append(gv[vi, ax])
            if log: log(f"vertex {k}: fd {fd_v[-1]:+.4e} adjoint {an_v[-1]:+.4e}")
        sc.update_vertices(mid, P0)
        fd_v, an_v = np.array(fd_v), np.array(an_v)
        out["vertex_coords"] = {"rel_l2": float(np.linalg.norm(fd_v - an_v) / np.linalg.norm(fd_v)), "cosine": float(fd_v @ an_v / (np.linalg.norm(fd_v) * np.linalg.norm(an_v))),
                                "eps_over_edge": 1.0}
    return out


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--res", type=int, default=512); ap.add_argument("--spp", type=int, default=64); ap.add_argument("--passes", type=int, default=8)
    ap.add_argument("--fd-spp", type=int, default=4096); ap.add_argument("--texels", type=int, default=64); ap.add_argument("--vertex-coords", type=int, default=32)
    ap.add_argument("--quiet", action="store_true"); ap.add_argument("--max-depth", type=int, default=None); ap.add_argument("--boundary", type=int, default=2)
    ap.add_argument("--edge-fraction", type=float, default=0.25); ap.add_argument("--rr-off", action="store_true"); ap.add_argument("--same-samples", action="store_true"); ap.add_argument("--variant", default="full")
    ap.add_argument("--convex", action="store_true", help="undisplaced sphere"); ap.add_argument("--flat", action="store_true", help="no vertex normals")
    ap.add_argument("--sphere-res", type=int, default=100)
    a = ap.parse_args()
    r = run(a.res, a.spp, a.passes, a.fd_spp, a.texels, a.vertex_coords, log=None if a.quiet else lambda s: print(s, file=sys.stderr, flush=True),
            max_depth=a.max_depth, boundary=a.boundary, edge_fraction=a.edge_fraction, rr_off=a.rr_off, same_samples=a.same_samples, variant=a.variant, displace=not a.convex, smooth=not a.flat,
            sphere=(a.sphere_res, a.sphere_res))
    print(json.dumps(r))
