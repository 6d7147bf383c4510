"""Differentiable rendering front end -- mirrors ``mitsuba.python.autodiff`` (src/python/python/autodiff.py):
``render_torch`` (:380-482) exposes the renderer as a torch autograd function so that ``torch.optim.Adam`` (the
reference's own ``Adam``/``SGD`` classes, :240-377, are thin re-implementations of the same updates) can drive inverse
rendering.  The backward pass is ``mtsb_render_grad`` (replay adjoint), not a tape.

Differentiable parameters (the reference exposes them through ``traverse``, src/python/python/util.py:14-179):
* a bitmap texture that enters a BSDF multiplicatively -- ``<bsdf>.reflectance.data`` of ``diffuse``,
  ``specular_reflectance`` of ``roughconductor`` (src/textures/bitmap.cpp:295-331);
* the vertex positions of a mesh -- ``<mesh>.vertex_positions_buf`` (src/librender/mesh.cpp:781-804): interior term +
  primary-visibility boundary term (csrc/adjoint.h lists what is not covered).
"""
from __future__ import annotations

import numpy as np


def render_torch(scene, texture=None, tex_id=-1, spp=None, seed=0, sensor=0, group=None, shard=True, vertices=None, mesh_id=-1):
    """Render `scene` with bitmap texture `tex_id` set to the torch tensor `texture` (H,W,C) and / or the vertex
    positions of mesh `mesh_id` set to `vertices` (V,3); returns the developed linear-RGB image (H,W,3) as a tensor on
    the parameters' device, differentiable w.r.t. both.

    With an initialised process group and `shard=True` every rank renders its share of the samples of the SAME view and
    the image and the gradients are all-reduced (distributed.py), so all ranks see identical values.
    `shard=False` renders locally (multi-view optimisation: one view per rank, the caller reduces the gradients)."""
    import torch
    from . import distributed as D

    if texture is None and vertices is None:
        raise ValueError("render_torch: pass a texture and / or vertices to differentiate")
    use_tex, use_vtx = texture is not None, vertices is not None
    if use_tex and tex_id < 0 or use_vtx and mesh_id < 0:
        raise ValueError("render_torch: tex_id / mesh_id missing")
    dev = (texture if use_tex else vertices).device

    def _host(t):
        return t if t.is_cuda else t.cpu().numpy()

    def _stamp(t):
        # storage, in-place version counter and a checksum (writes through `.data` do not bump the counter)
        return (t.data_ptr(), t._version, tuple(t.shape), float(t.detach().double().sum()))

    class _Render(torch.autograd.Function):
        @staticmethod
        def forward(ctx, tex, vtx):
            # several views of one iteration pass the same tensors: the scene is updated (texture upload; vertex normals +
            # BVH rebuild) only when a parameter changed since the update it last made from it
            for use, key, t, update in ((use_tex, ("texture", tex_id), tex, scene.update_texture), (use_vtx, ("vertices", mesh_id), vtx, scene.update_vertices)):
                if not use:
                    continue
                stamp = _stamp(t)
                if scene._param_stamps.get(key) != stamp:
                    update(key[1], _host(t.detach().contiguous().float()))
                    scene._param_stamps[key] = stamp
            rank, world = D.world_info(group) if shard else (0, 1)
            total = int(scene.data.spp if spp is None else spp)
            begin, count = D.sample_share(total, rank, world)
            ctx.share = (begin, count, total)
            if world > 1:
                img = scene.render_distributed(sensor, total, seed, group)
            elif dev.type == "cuda" and dev.index == scene.device:
                # parameters live on the scene's GPU: the developed image stays there too (no host round trip)
                scene.film_clear(sensor)
                scene.render_accumulate(sensor, total, seed, 0, total)
                return scene.film_develop_tensor(sensor)
            else:
                img = scene.render(sensor, total, seed)
            return torch.from_numpy(np.ascontiguousarray(img)).to(dev)

        @staticmethod
        def backward(ctx, grad_img):
            begin, count, total = ctx.share
            g = grad_img.detach().contiguous().float()
            gt = gv = None
            if count > 0:
                gt, gv = scene.render_grad(_host(g), sensor, count, seed, tex_id if use_tex else -1,
                                           mesh_id if use_vtx else -1, begin, total)
            out = []
            for want, val, like in ((use_tex, gt, texture), (use_vtx, gv, vertices)):
                if not want:
                    out.append(None)
                    continue
                if val is None:
                    val = np.zeros(int(np.prod(like.shape)), np.float32)
                val = val if isinstance(val, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(val))
                out.append(val.to(grad_img.device).reshape(like.shape))
            if shard:
                D.allreduce_gradients([o for o in out if o is not None], group)
            return out[0], out[1]

    dummy = torch.zeros(1, device=dev)
    return _Render.apply(texture if use_tex else dummy, vertices if use_vtx else dummy)
