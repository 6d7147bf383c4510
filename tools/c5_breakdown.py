#!/usr/bin/env python
"""Where one C5 iteration spends its time (GPU): vertex update (normals + BVH rebuild), primal render, gradient pass, at the
example's size (256 x 256, 32 spp, 545-vertex sphere in the Cornell box).  Wall clock around synchronised calls."""
import argparse
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--res", type=int, default=256); ap.add_argument("--spp", type=int, default=32); ap.add_argument("--sphere", type=int, default=16)
    ap.add_argument("--reps", type=int, default=20)
    a = ap.parse_args()
    import mitsuba2_b200 as mi
    from mitsuba2_b200 import scenes
    mi.set_variant("cuda_b200_rgb")
    sd, tex = scenes.c4_scene(os.path.join(ROOT, "tests", "golden", "cbox_scene.npz"), res=a.res, spp=a.spp, tex_res=64, sphere=(a.sphere, 2 * a.sphere))
    mid = next(i for i, m in enumerate(sd.meshes) if m.name == "c4_sphere")
    sc = mi.Scene(sd, device=0)
    P = torch.from_numpy(np.asarray(sd.meshes[mid].positions, np.float32).reshape(-1, 3).copy()).cuda()
    dl = torch.ones(a.res, a.res, 3, device="cuda") / (a.res * a.res)

    def timed(fn):
        fn(); torch.cuda.synchronize()
        t = time.perf_counter()
        for _ in range(a.reps):
            fn()
        torch.cuda.synchronize()
        return (time.perf_counter() - t) / a.reps * 1e3

    def primal():
        sc.film_clear(0); sc.render_accumulate(0, a.spp, 1, 0, a.spp); return sc.film_develop_tensor(0)

    out = {"update_vertices_ms": timed(lambda: sc.update_vertices(mid, P)), "primal_ms": timed(primal),
           "grad_vertices_ms": timed(lambda: sc.render_grad(dl, 0, a.spp, 1, -1, mid)),
           "grad_texture_ms": timed(lambda: sc.render_grad(dl, 0, a.spp, 1, tex, -1)), "build": sc.build_stats()}
    print(out)


if __name__ == "__main__":
    main()
