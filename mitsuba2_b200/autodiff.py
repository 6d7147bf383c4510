"""Differentiable rendering front end -- mirrors ``mitsuba.python.autodiff`` (src/python/python/autodiff.py):
``render_torch`` (:380-482) exposes the renderer as a torch autograd function so that ``torch.optim.Adam`` (the
reference's own ``Adam``/``SGD`` classes, :240-377, are thin re-implementations of the same updates) can drive inverse
rendering.  The backward pass is ``mtsb_render_grad`` (replay adjoint), not a tape.

Only bitmap textures that enter a BSDF multiplicatively (diffuse ``reflectance``, roughconductor
``specular_reflectance``) are differentiable in this round; vertex positions raise.
"""
from __future__ import annotations

import numpy as np


def render_torch(scene, texture, tex_id, spp=None, seed=0, sensor=0, group=None, shard=True):
    """Render `scene` with bitmap texture `tex_id` set to the torch tensor `texture` (H,W,C); returns the developed
    linear-RGB image (H,W,3) as a tensor on the texture's device, differentiable w.r.t. `texture`.

    With an initialised process group and `shard=True` every rank renders its share of the samples of the SAME view and
    both the image and the texture gradient are all-reduced (distributed.py), so all ranks see identical values.
    `shard=False` renders locally (multi-view optimisation: one view per rank, the caller reduces the gradients)."""
    import torch
    from . import distributed as D

    class _Render(torch.autograd.Function):
        @staticmethod
        def forward(ctx, tex):
            t = tex.detach().contiguous().float()
            scene.update_texture(tex_id, t if t.is_cuda else t.cpu().numpy())
            rank, world = D.world_info(group) if shard else (0, 1)
            total = int(scene.data.spp if spp is None else spp)
            begin, count = D.sample_share(total, rank, world)
            ctx.share = (begin, count, total)
            if world > 1:
                img = scene.render_distributed(sensor, total, seed, group)
            else:
                img = scene.render(sensor, total, seed)
            return torch.from_numpy(np.ascontiguousarray(img)).to(tex.device)

        @staticmethod
        def backward(ctx, grad_img):
            begin, count, total = ctx.share
            g = grad_img.detach().contiguous().float()
            gt = None
            if count > 0:
                gt, _ = scene.render_grad(g if g.is_cuda else g.cpu().numpy(), sensor, count, seed, tex_id, -1, begin, total)
            if gt is None:
                gt = np.zeros(int(np.prod(texture.shape)), np.float32)
            gt = gt if isinstance(gt, torch.Tensor) else torch.from_numpy(gt)
            gt = gt.to(grad_img.device).reshape(texture.shape)
            if shard:
                D.allreduce_gradients([gt], group)
            return gt

    return _Render.apply(texture)
