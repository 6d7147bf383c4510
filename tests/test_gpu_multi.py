"""Multi-GPU (needs >= 2 devices; skipped otherwise): spp-sharded render + NCCL film all-reduce equals the one-GPU render
of the union of the sample indices; the texture and vertex gradient all-reduces equal the one-GPU gradients."""
import os
import socket

import numpy as np
import pytest

import helpers as H

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_dir):
    import torch
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        import mitsuba2_b200 as mi
        from mitsuba2_b200 import autodiff
        mi.set_variant("cuda_b200_rgb")
        sd, tex = H.textured_cbox(res=24, spp=8, tex_res=4)
        sc = mi.Scene(sd, device=rank)
        img = sc.render_distributed(spp=8, seed=3)
        np.save(os.path.join(out_dir, f"img{rank}.npy"), img)
        p = torch.from_numpy(sd.textures[tex]["data"].copy()).cuda(rank).requires_grad_(True)
        out = autodiff.render_torch(sc, p, tex, spp=8, seed=3)
        wgt = torch.from_numpy(np.random.default_rng(0).normal(size=(24, 24, 3)).astype(np.float32)).cuda(rank)
        (out * wgt).sum().backward()
        np.save(os.path.join(out_dir, f"grad{rank}.npy"), p.grad.cpu().numpy())
        # vertex positions: samples sharded over the ranks, gradient all-reduced
        sdv = H.vgrad_scene("quadshadow", "diffuse", res=16, spp=8)
        scv = mi.Scene(sdv, device=rank)
        scv.grad_config(boundary=2, edge_fraction=0.5)
        v = torch.from_numpy(np.asarray(sdv.meshes[0].positions, np.float32).reshape(-1, 3).copy()).cuda(rank).requires_grad_(True)
        outv = autodiff.render_torch(scv, spp=8, seed=4, vertices=v, mesh_id=0)
        wv = torch.from_numpy(np.random.default_rng(1).uniform(0.5, 1.5, (16, 16, 3)).astype(np.float32)).cuda(rank)
        (outv * wv).sum().backward()
        np.save(os.path.join(out_dir, f"vgrad{rank}.npy"), v.grad.cpu().numpy())
    finally:
        dist.destroy_process_group()


def test_two_gpu_render_and_gradient_equal_single_gpu(tmp_path):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import torch.multiprocessing as mp
    import mitsuba2_b200 as mi
    mp.spawn(_worker, args=(2, _free_port(), str(tmp_path)), nprocs=2, join=True)
    mi.set_variant("cuda_b200_rgb")
    sd, tex = H.textured_cbox(res=24, spp=8, tex_res=4)
    sc = mi.Scene(sd, device=0)
    ref = sc.render(spp=8, seed=3)
    i0, i1 = np.load(tmp_path / "img0.npy"), np.load(tmp_path / "img1.npy")
    np.testing.assert_array_equal(i0, i1)
    np.testing.assert_allclose(i0, ref, rtol=1e-4, atol=1e-5)
    wgt = np.random.default_rng(0).normal(size=(24, 24, 3)).astype(np.float32)
    g, _ = sc.render_grad(wgt, spp=8, seed=3, tex_id=tex)
    g0, g1 = np.load(tmp_path / "grad0.npy"), np.load(tmp_path / "grad1.npy")
    np.testing.assert_array_equal(g0, g1)
    np.testing.assert_allclose(g0.ravel(), g, rtol=1e-3, atol=1e-4 * np.abs(g).max())
    sdv = H.vgrad_scene("quadshadow", "diffuse", res=16, spp=8)
    scv = mi.Scene(sdv, device=0)
    scv.grad_config(boundary=2, edge_fraction=0.5)
    wv = np.random.default_rng(1).uniform(0.5, 1.5, (16, 16, 3)).astype(np.float32)
    _, gv = scv.render_grad(wv, spp=8, seed=4, mesh_id=0)
    v0, v1 = np.load(tmp_path / "vgrad0.npy"), np.load(tmp_path / "vgrad1.npy")
    np.testing.assert_array_equal(v0, v1)
    np.testing.assert_allclose(v0, gv, rtol=1e-3, atol=1e-4 * np.abs(gv).max())
