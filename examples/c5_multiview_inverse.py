#!/usr/bin/env python
"""C5: multi-view inverse rendering, the views dealt round-robin to the ranks (8 views, one per rank on 8 GPUs).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 examples/c5_multiview_inverse.py

Every rank holds the whole scene, renders ITS views with all samples, computes the L2 image loss against the target
of each view, and back-propagates through `mtsb_render_grad`; the parameter gradients are summed over the ranks with one NCCL
all-reduce and every rank applies the identical Adam update (parameters never need a broadcast).

--param vertices (default, BASELINE config C5): Adam on the vertex positions of the sphere, started from a shrunk and
shifted copy of the target shape.  --param texture: Adam on the floor's albedo texture.
"""
import argparse
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import mitsuba2_b200 as mi                      # noqa: E402
from mitsuba2_b200 import scenes, distributed   # noqa: E402
from mitsuba2_b200.scenedata import look_at     # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=100)
    ap.add_argument("--res", type=int, default=256)
    ap.add_argument("--spp", type=int, default=16)
    ap.add_argument("--tex", type=int, default=64)
    ap.add_argument("--param", choices=("vertices", "texture"), default="vertices")
    ap.add_argument("--sphere", type=int, default=16, help="latitude bands of the sphere (longitude = 2x)")
    ap.add_argument("--lr", type=float, default=None)
    ap.add_argument("--views", type=int, default=None, help="number of sensors (default: max(8, ranks)); rank r owns views r, r + ranks, ...")
    ap.add_argument("--smooth", type=int, default=8, help="Jacobi sweeps of gradient smoothing over the mesh (vertices mode)")
    args = ap.parse_args()
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    mi.set_variant("cuda_b200_rgb")
    sd, tex = scenes.c4_scene(os.path.join(ROOT, "tests", "golden", "cbox_scene.npz"), res=args.res, spp=args.spp, tex_res=args.tex,
                              sphere=(args.sphere, 2 * args.sphere))
    mid = next(i for i, m in enumerate(sd.meshes) if m.name == "c4_sphere")
    # the views orbit the camera around the box centre; this rank's sensors are its share of them
    n_views = args.views or max(8, world)
    mine = list(range(rank, n_views, world))
    s0 = sd.sensors[0]
    c = np.array([278.0, 273.0, 280.0])
    sd.sensors = []
    for v in mine:
        ang = (v - (n_views - 1) / 2) * 0.12
        eye = c + np.array([np.sin(ang) * 1080.0, 0.0, -np.cos(ang) * 1080.0])
        sd.add_perspective(look_at(eye, c, [0, 1, 0]), s0["fov_x"], s0["near"], s0["far"], args.res, args.res, "box")
    scene = mi.Scene(sd, device=local)
    scene.update_config(refit=True, rebuild_every=16)      # Adam moves the vertices a little per step: refit the BVH, rebuild now and then
    from mitsuba2_b200 import autodiff
    targets = [torch.from_numpy(scene.render(sensor=k, spp=32 * args.spp, seed=12345).copy()).cuda(local) for k in range(len(mine))]
    if args.param == "texture":
        truth = torch.from_numpy(sd.textures[tex]["data"].copy()).cuda(local)
        param = torch.full_like(truth, 0.5, requires_grad=True)
        lr = args.lr or 0.02
    else:
        truth = torch.from_numpy(np.asarray(sd.meshes[mid].positions, np.float32).reshape(-1, 3).copy()).cuda(local)
        c0 = truth.mean(0, keepdim=True)
        size = float((truth.max(0).values - truth.min(0).values).max())
        start = c0 + 0.8 * (truth - c0) + torch.tensor([[0.06 * size, 0.04 * size, 0.0]], device=truth.device)
        param = start.clone().requires_grad_(True)
        lr = args.lr or 0.004 * size
        # per-vertex Monte Carlo gradients are noisy (a vertex sees a handful of pixels): low-pass them over the mesh
        # before Adam (a few Jacobi sweeps of g_i <- mean of g over the closed 1-ring), as mesh-based inverse rendering
        # commonly does; the gradient itself is what mtsb_render_grad returns
        fc = torch.from_numpy(np.asarray(sd.meshes[mid].faces, np.int64).reshape(-1, 3)).cuda(local)
        ii = torch.cat([fc[:, 0], fc[:, 1], fc[:, 2], fc[:, 1], fc[:, 2], fc[:, 0]])
        jj = torch.cat([fc[:, 1], fc[:, 2], fc[:, 0], fc[:, 0], fc[:, 1], fc[:, 2]])
        nv = truth.shape[0]
        adj = torch.sparse_coo_tensor(torch.stack([ii, jj]), torch.ones(ii.numel(), device=ii.device), (nv, nv)).coalesce()
        adj = torch.sparse_coo_tensor(adj.indices(), torch.ones_like(adj.values()), (nv, nv))
        deg = torch.sparse.sum(adj, 1).to_dense().unsqueeze(1)

        def smooth(gv, sweeps=args.smooth):
            for _ in range(sweeps):
                gv = (gv + torch.sparse.mm(adj, gv)) / (1.0 + deg)
            return gv
    def converged_loss():
        """L2 loss of this rank's views at 32x the samples (the per-iteration loss is dominated by Monte Carlo noise), summed over ranks."""
        with torch.no_grad():
            if args.param == "texture":
                scene.update_texture(tex, param.detach().cpu().numpy())
            else:
                scene.update_vertices(mid, param.detach().cpu().numpy())
            tot = torch.zeros(1, device=param.device)
            for k, target in enumerate(targets):
                tot += ((torch.from_numpy(scene.render(sensor=k, spp=32 * args.spp, seed=777).copy()).to(param.device) - target) ** 2).mean()
            if world > 1:
                dist.all_reduce(tot)
        return tot.item() / n_views

    opt = torch.optim.Adam([param], lr=lr)
    loss_start = converged_loss()
    t0 = torch.cuda.Event(enable_timing=True); t1 = torch.cuda.Event(enable_timing=True)
    t0.record()
    for it in range(args.iters):
        opt.zero_grad()
        for k, target in enumerate(targets):           # gradients of this rank's views accumulate in param.grad
            if args.param == "texture":
                img = autodiff.render_torch(scene, param, tex, spp=args.spp, seed=it, sensor=k, shard=False)
            else:
                img = autodiff.render_torch(scene, spp=args.spp, seed=it, sensor=k, shard=False, vertices=param, mesh_id=mid)
            loss = ((img - target) ** 2).mean()
            loss.backward()
        if world > 1:                              # sum the per-view gradients; identical update everywhere
            distributed.allreduce_gradients([param.grad])
        if args.param == "vertices":
            param.grad.copy_(smooth(param.grad))
        opt.step()
        if args.param == "texture":
            with torch.no_grad():
                param.clamp_(0.01, 0.99)
        if rank == 0 and (it % 10 == 0 or it == args.iters - 1):
            err = (param.detach() - truth).abs().mean().item()
            extra = ""
            if args.param == "vertices":               # vertices may slide along the surface: also the distance to the nearest target vertex
                seen = (truth - c0)[:, 2] < -0.3 * (truth - c0).norm(dim=1)     # the cameras orbit on the -z side: the back is never observed
                dist_seen = torch.cdist(param.detach()[seen], truth).min(dim=1).values.mean().item()
                extra = f"  camera-facing vertices: mean distance to the target's vertices {dist_seen:.4f}"
            print(f"iter {it:4d}  loss(view {mine[-1]}) {loss.item():.6f}  mean |{args.param} error| {err:.4f}{extra}", flush=True)
    t1.record(); torch.cuda.synchronize()
    loss_end = converged_loss()
    if rank == 0:
        print(f"mean L2 image loss over the {n_views} views at {32 * args.spp} spp: {loss_start:.3e} -> {loss_end:.3e}", flush=True)
        print(f"{n_views} views on {world} rank(s): {args.iters} iterations in {t0.elapsed_time(t1) / 1e3:.2f} s ({t0.elapsed_time(t1) / args.iters:.1f} ms / iteration)", flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
