#!/usr/bin/env python
"""Small driver for ncu captures (profiles/README.md): one short invocation of a hot path, nothing timed.

    python tools/prof_run.py spectral   # C3 in the spectral mode, 2 x 2 spp
    python tools/prof_run.py grad       # C4 gradient pass (texture + vertex positions), 8 spp
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np                      # noqa: E402
import mitsuba2_b200 as mi              # noqa: E402
from mitsuba2_b200 import scenes        # noqa: E402


def main():
    what = sys.argv[1] if len(sys.argv) > 1 else "spectral"
    mi.set_variant("cuda_b200_rgb")
    if what in ("spectral_time", "rgb_time"):      # stage timing of the C3 step (not a bench value: no clock sampling)
        sd = scenes.c3_scene()
        if what == "spectral_time":
            sd.mode = "spectral"
        sc = mi.Scene(sd, device=0)
        sc.set_profiling(True)
        for i in range(6):
            sc.film_clear()
            st = sc.render_accumulate(spp=16, seed=i, spp_begin=0, spp_total=16)
            if i >= 3:
                print({k: round(st[k], 3) for k in ("ms_total", "ms_closest", "ms_shade", "ms_any")})
    elif what == "grad_time":      # C4 gradient pass (texture + vertices), stage timing
        import torch
        sd, tex = scenes.c4_scene(os.path.join(ROOT, "tests", "golden", "cbox_scene.npz"))
        mid = next(i for i, m in enumerate(sd.meshes) if m.name == "c4_sphere")
        sc = mi.Scene(sd, device=0)
        sc.set_profiling(True)
        dl = torch.randn(512 * 512 * 3, device="cuda:0")
        for i in range(5):
            sc.render_grad(dl, spp=64, seed=i, tex_id=tex, mesh_id=mid)
            if i >= 2:
                print({k: round(sc.last_stats[k], 3) for k in ("ms_total", "ms_closest", "ms_shade", "ms_any")})
    elif what == "update_time":    # cost of one parameter update of the C4 sphere (vertex upload, normals, BVH rebuild) + a small render
        import time
        import torch
        sd, tex = scenes.c4_scene(os.path.join(ROOT, "tests", "golden", "cbox_scene.npz"), res=256, spp=32)
        mid = next(i for i, m in enumerate(sd.meshes) if m.name == "c4_sphere")
        sc = mi.Scene(sd, device=0)
        P = torch.from_numpy(np.asarray(sd.meshes[mid].positions, np.float32).copy()).cuda()
        dl = torch.randn(256 * 256 * 3, device="cuda:0")
        for name, fn in (("update_vertices(device tensor)", lambda: sc.update_vertices(mid, P)),
                         ("render 256^2 x 32", lambda: sc.render(spp=32, seed=1)),
                         ("render_grad vertices 256^2 x 32", lambda: sc.render_grad(dl, spp=32, seed=1, mesh_id=mid))):
            fn(); torch.cuda.synchronize()
            t0 = time.perf_counter()
            for _ in range(5):
                fn()
            torch.cuda.synchronize()
            print(f"{name}: {(time.perf_counter() - t0) / 5 * 1e3:.2f} ms")
    elif what in ("c2", "c2_time"):      # C2 intersection launches: coherent closest, incoherent closest, coherent any, incoherent any
        import torch
        sd = scenes.c2_scene()
        sc = mi.Scene(sd, device=0)
        for kind in ("coherent", "incoherent"):
            om, dm = scenes.c2_rays(sd, kind)
            _, dms = scenes.c2_rays(sd, kind + "_shadow")
            o_t, d_t, ds_t = (torch.from_numpy(x).cuda() for x in (om, dm, dms))
            reps = 12 if what == "c2_time" else 1
            for mode, fn, dd in (("closest", sc.intersect_closest_raw, d_t), ("any", sc.intersect_any_raw, ds_t)):
                ms = []
                for _ in range(reps):
                    fn(o_t, dd)
                    ms.append(sc.last_stats["ms_total"])
                print(kind, mode, f"{om.shape[0] / (float(np.median(ms)) * 1e-3) / 1e6:.1f} Mrays/s")
    elif what == "spectral":
        sd = scenes.c3_scene()
        sd.mode = "spectral"
        sc = mi.Scene(sd, device=0)
        for i in range(2):
            sc.film_clear()
            sc.render_accumulate(spp=2, seed=i, spp_begin=0, spp_total=2)
    else:
        sd, tex = scenes.c4_scene(os.path.join(ROOT, "tests", "golden", "cbox_scene.npz"))
        mid = next(i for i, m in enumerate(sd.meshes) if m.name == "c4_sphere")
        sc = mi.Scene(sd, device=0)
        dl = np.random.default_rng(0).normal(size=(512, 512, 3)).astype(np.float32)
        sc.render_grad(dl, spp=8, seed=1, tex_id=tex, mesh_id=mid)
    print("done", what)


if __name__ == "__main__":
    main()
