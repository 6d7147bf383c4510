"""Device scene: Python mirror of the reference's ``Scene`` / ``Integrator`` / ``Film`` entry points for the
hot path, calling the CUDA library through the C ABI (include/mtsb200.h).

Reference interfaces mirrored (names and argument meaning):
  ``Scene.ray_intersect(ray)`` / ``Scene.ray_test(ray)``  -- src/librender/scene.cpp:108-139
  ``Integrator.render(scene, sensor)``                    -- src/librender/integrator.cpp:52-176
  ``Film.bitmap(raw)`` / ``Film.develop()``               -- src/films/hdrfilm.cpp:249-342
"""
import ctypes as C

import numpy as np

from . import _abi as A
from .scenedata import SceneData

_ctx_cache = {}


def _context(device):
    if device not in _ctx_cache:
        h = C.c_void_p()
        A.check(A.lib().mtsb_ctx_create(int(device), C.byref(h)))
        _ctx_cache[device] = h
    return _ctx_cache[device]


def _is_torch_cuda(x):
    return hasattr(x, "is_cuda") and x.is_cuda


def _arg(x, dtype, n_expected=None, device=None):
    """-> (pointer, mem, keepalive) for a numpy array or a CUDA torch tensor."""
    if _is_torch_cuda(x):
        import torch
        want = {np.float32: torch.float32, np.uint32: torch.int32, np.int32: torch.int32}.get(dtype, torch.float32)
        if x.dtype != want and not (dtype == np.uint32 and x.dtype == torch.uint32):
            raise TypeError(f"CUDA tensor of dtype {x.dtype} passed where {want} is required (the library reads raw device memory)")
        if device is not None and x.device.index != int(device):
            raise ValueError(f"CUDA tensor lives on cuda:{x.device.index}, the scene on cuda:{device}")
        x = x.contiguous()
        if n_expected is not None and x.numel() != n_expected:
            raise ValueError(f"expected {n_expected} elements, got {x.numel()}")
        # the library works on its own non-blocking stream: everything torch has queued for this tensor must be done first
        torch.cuda.current_stream(x.device).synchronize()
        return C.c_void_p(x.data_ptr()), A.MEM_DEVICE, x
    a = np.ascontiguousarray(x, dtype=dtype)
    if n_expected is not None and a.size != n_expected:
        raise ValueError(f"expected {n_expected} elements, got {a.size}")
    return C.c_void_p(a.ctypes.data), A.MEM_HOST, a


class Scene:
    def __init__(self, data: SceneData, device=0):
        self.data = data
        self.device = device
        self._param_stamps = {}     # autodiff.render_torch: (kind, id) -> stamp of the tensor the scene was last updated from
        self._ctx = _context(device)
        self._h = C.c_void_p()
        desc, keep = data.to_desc()
        A.check(A.lib().mtsb_scene_create(self._ctx, C.byref(desc), C.byref(self._h)))
        del keep
        self.last_stats = None
        if getattr(data, "moment", False):      # <integrator type="moment"> around `path` (src/integrators/moment.cpp)
            for i in range(len(data.sensors)):
                self.enable_moments(True, i)
        prim_counts = [m.faces.shape[0] for m in data.meshes] + [1] * len(getattr(data, "shapes", []))
        self.primitive_map = np.concatenate([[0], np.cumsum(prim_counts)]).astype(np.int64)

    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            A.lib().mtsb_scene_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_profiling(self, on=True):
        A.check(A.lib().mtsb_set_profiling(self._ctx, int(bool(on))))

    # ---- statistics ---------------------------------------------------------------------------
    def build_stats(self):
        st = A.Stats()
        A.check(A.lib().mtsb_scene_stats(self._h, C.byref(st)))
        return st.as_dict()

    # ---- Scene::ray_intersect / ray_test --------------------------------------------------------
    @staticmethod
    def pack_rays(o, d, mint=0.0, maxt=np.inf):
        o = np.asarray(o, dtype=np.float32).reshape(-1, 3)
        d = np.asarray(d, dtype=np.float32).reshape(-1, 3)
        n = o.shape[0]
        om = np.empty((n, 4), np.float32)
        dm = np.empty((n, 4), np.float32)
        om[:, :3], dm[:, :3] = o, d
        om[:, 3], dm[:, 3] = mint, maxt
        return om, dm

    def intersect_closest_raw(self, o_mint, d_maxt, hits=None, naive=False):
        """SoA entry: o_mint/d_maxt (n,4) float32 numpy or CUDA torch; returns hits (n,4) = t,u,v,prim bits."""
        po, mem, k0 = _arg(o_mint, np.float32, device=self.device)
        pd, mem2, k1 = _arg(d_maxt, np.float32, device=self.device)
        assert mem == mem2
        n = (k0.numel() if mem == A.MEM_DEVICE else k0.size) // 4
        if hits is None:
            if mem == A.MEM_DEVICE:
                import torch
                hits = torch.empty((n, 4), dtype=torch.float32, device=k0.device)
            else:
                hits = np.empty((n, 4), np.float32)
        ph = C.c_void_p(hits.data_ptr()) if mem == A.MEM_DEVICE else C.c_void_p(hits.ctypes.data)
        st = A.Stats()
        if naive:
            A.check(A.lib().mtsb_intersect_naive(self._h, n, po, pd, ph, mem))
        else:
            A.check(A.lib().mtsb_intersect_closest(self._h, n, po, pd, ph, mem, C.byref(st)))
        self.last_stats = st.as_dict()
        return hits

    def intersect_any_raw(self, o_mint, d_maxt, out=None):
        po, mem, k0 = _arg(o_mint, np.float32, device=self.device)
        pd, mem2, k1 = _arg(d_maxt, np.float32, device=self.device)
        assert mem == mem2
        n = (k0.numel() if mem == A.MEM_DEVICE else k0.size) // 4
        if out is None:
            if mem == A.MEM_DEVICE:
                import torch
                out = torch.empty((n,), dtype=torch.int32, device=k0.device)
            else:
                out = np.empty((n,), np.uint32)
        ph = C.c_void_p(out.data_ptr()) if mem == A.MEM_DEVICE else C.c_void_p(out.ctypes.data)
        st = A.Stats()
        A.check(A.lib().mtsb_intersect_any(self._h, n, po, pd, ph, mem, C.byref(st)))
        self.last_stats = st.as_dict()
        return out

    def ray_intersect(self, o, d, mint=0.0, maxt=np.inf):
        """Mirror of ``Scene::ray_intersect``: returns a dict with ``t`` (inf on miss), ``prim_index`` (within the
        shape), ``shape_index``, ``prim_uv`` (barycentrics b1,b2) and ``is_valid``."""
        om, dm = self.pack_rays(o, d, mint, maxt)
        hits = self.intersect_closest_raw(om, dm)
        gprim = hits[:, 3].copy().view(np.uint32).astype(np.int64)
        valid = np.isfinite(hits[:, 0])
        shape = np.searchsorted(self.primitive_map, np.where(valid, gprim, 0), side="right") - 1
        return {"t": hits[:, 0].copy(), "prim_uv": hits[:, 1:3].copy(), "is_valid": valid,
                "global_prim_index": np.where(valid, gprim, -1),
                "shape_index": np.where(valid, shape, -1),
                "prim_index": np.where(valid, gprim - self.primitive_map[np.clip(shape, 0, None)], -1)}

    def ray_test(self, o, d, mint=0.0, maxt=np.inf):
        om, dm = self.pack_rays(o, d, mint, maxt)
        return self.intersect_any_raw(om, dm).astype(bool)

    # ---- Integrator::render ---------------------------------------------------------------------
    def render_accumulate(self, sensor=0, spp=None, seed=0, spp_begin=0, spp_total=None):
        """Adds ``spp`` samples per pixel (sample indices [spp_begin, spp_begin+spp)) to the device film."""
        spp = self.data.spp if spp is None else int(spp)
        spp_total = spp_begin + spp if spp_total is None else int(spp_total)
        st = A.Stats()
        A.check(A.lib().mtsb_render(self._h, sensor, spp_begin, spp, spp_total, int(seed), C.byref(st)))
        self.last_stats = st.as_dict()
        return self.last_stats

    def render(self, sensor=0, spp=None, seed=0, samples_per_pass=None, timeout=None, progress=None):
        """``Integrator.render`` + ``Film.bitmap()``: clears the film, renders, returns linear RGB (H,W,3) float32.

        ``samples_per_pass`` / ``timeout`` are the properties of ``SamplingIntegrator`` (src/librender/integrator.cpp:35-36,
        60-66): the job runs as spp / samples_per_pass passes over the film (``progress(pass_index, n_passes, scene)`` is called
        after each, e.g. to develop a preview), and a job that exceeds ``timeout`` seconds stops early and returns the partial
        film, as the reference does.  A concurrent :meth:`cancel` raises :class:`~mitsuba2_b200._abi.MtsbError` (code -4)."""
        import time
        spp = int(self.data.spp if spp is None else spp)
        per_pass = spp if samples_per_pass in (None, -1) else min(int(samples_per_pass), spp)
        if per_pass <= 0 or spp % per_pass != 0:
            raise ValueError(f"sample_count ({spp}) must be a multiple of samples_per_pass ({samples_per_pass}).")
        n_passes, t0 = spp // per_pass, time.perf_counter()
        self.film_clear(sensor)
        try:
            for k in range(n_passes):
                if timeout is not None and timeout > 0:
                    left = timeout - (time.perf_counter() - t0)
                    if left <= 0:
                        break
                    self.set_timeout(left)
                self.render_accumulate(sensor, per_pass, seed, k * per_pass, spp)
                if progress is not None:
                    progress(k, n_passes, self)
        finally:
            if timeout is not None:
                self.set_timeout(-1.0)
        return self.film_develop(sensor)

    def cancel(self):
        """``Integrator.cancel()`` (include/mitsuba/render/integrator.h:51): may be called from another thread."""
        A.check(A.lib().mtsb_cancel(self._ctx))

    def set_timeout(self, seconds):
        A.check(A.lib().mtsb_set_timeout(self._ctx, float(seconds)))

    def film_clear(self, sensor=0):
        A.check(A.lib().mtsb_film_clear(self._h, sensor))

    def _film_shape(self, sensor):
        s = self.data.sensors[sensor]
        return s["height"], s["width"]

    def film_xyzaw(self, sensor=0):
        h, w = self._film_shape(sensor)
        out = np.empty((h, w, 5), np.float32)
        A.check(A.lib().mtsb_film_read(self._h, sensor, C.c_void_p(out.ctypes.data), A.MEM_HOST))
        return out

    def enable_moments(self, on=True, sensor=0):
        """Second-moment AOVs of the ``moment`` integrator (src/integrators/moment.cpp:55-90) for later renders."""
        A.check(A.lib().mtsb_film_enable_moments(self._h, sensor, int(bool(on))))

    def film_moments(self, sensor=0):
        """(H,W,3) sums of w X^2, w Y^2, w Z^2 -- the ``m2_<nested>.X/Y/Z`` channels before the division by W."""
        h, w = self._film_shape(sensor)
        out = np.empty((h, w, 3), np.float32)
        A.check(A.lib().mtsb_film_read_moments(self._h, sensor, C.c_void_p(out.ctypes.data), A.MEM_HOST))
        return out

    def render_moments(self, sensor=0, spp=None, seed=0):
        """``moment(path)``: -> (E[xyz], E[xyz^2]) per pixel, float64 (H,W,3) each, what the reference's render tests feed to
        their z-test (src/librender/tests/test_renders.py:60-126: var = E[x^2] - E[x]^2)."""
        self.enable_moments(True, sensor)
        self.film_clear(sensor)
        self.render_accumulate(sensor, spp, seed)
        f = self.film_xyzaw(sensor).astype(np.float64)
        w = np.maximum(f[:, :, 4:5], 1e-12)
        return f[:, :, :3] / w, self.film_moments(sensor).astype(np.float64) / w

    def film_accumulate(self, xyzaw, sensor=0):
        p, mem, keep = _arg(xyzaw, np.float32, device=self.device)
        A.check(A.lib().mtsb_film_accumulate(self._h, sensor, p, mem))

    def film_develop(self, sensor=0):
        h, w = self._film_shape(sensor)
        out = np.empty((h, w, 3), np.float32)
        A.check(A.lib().mtsb_film_develop(self._h, sensor, C.c_void_p(out.ctypes.data), A.MEM_HOST))
        return out

    def film_develop_tensor(self, sensor=0):
        """`film_develop` into a torch CUDA tensor (H,W,3) on this scene's device: no host round trip (autodiff.render_torch)."""
        import torch
        h, w = self._film_shape(sensor)
        out = torch.empty((h, w, 3), dtype=torch.float32, device=torch.device("cuda", self.device))
        torch.cuda.current_stream(out.device).synchronize()      # the allocation is torch's, the kernel runs on the library's stream
        A.check(A.lib().mtsb_film_develop(self._h, sensor, C.c_void_p(out.data_ptr()), A.MEM_DEVICE))
        return out

    def film_device_buffer(self, sensor=0):
        """(device pointer, n_floats) of the packed XYZAW film, for NCCL reduction by the caller."""
        p, n = C.c_void_p(), C.c_uint64()
        A.check(A.lib().mtsb_film_device_ptr(self._h, sensor, C.byref(p), C.byref(n)))
        return p.value, n.value

    def render_xyzaw_tensor(self, spp_begin, spp_count, spp_total, seed=0, sensor=0):
        """Clear the film, render sample indices [spp_begin, spp_begin+spp_count) and return the packed XYZAW sums as
        a flat torch CUDA tensor (device-to-device copy; the form `distributed.render_sharded` reduces over NCCL)."""
        import torch
        h, w = self._film_shape(sensor)
        film = torch.empty(h * w * 5, dtype=torch.float32, device=f"cuda:{self.device}")
        self.film_clear(sensor)
        if spp_count > 0:
            self.render_accumulate(sensor, spp_count, seed, spp_begin, spp_total)
        A.check(A.lib().mtsb_film_read(self._h, sensor, C.c_void_p(film.data_ptr()), A.MEM_DEVICE))
        return film

    def render_distributed(self, sensor=0, spp=None, seed=0, group=None):
        """`Integrator.render` over all ranks of the process group: every rank renders its share of the sample
        indices, the film sums are all-reduced (NCCL) and installed as this rank's film; returns linear RGB."""
        import torch
        from . import distributed as D
        spp = int(self.data.spp if spp is None else spp)
        film = D.render_sharded(lambda b, c, t, sd: self.render_xyzaw_tensor(b, c, t, sd, sensor), spp, seed, group)
        torch.cuda.synchronize(self.device)
        self.film_clear(sensor)
        A.check(A.lib().mtsb_film_accumulate(self._h, sensor, C.c_void_p(film.data_ptr()), A.MEM_DEVICE))
        return self.film_develop(sensor)

    def film_put(self, pos, values, sensor=0):
        """``ImageBlock::put(pos, value)`` batched: pos (n,2), values (n,5)."""
        pos = np.ascontiguousarray(pos, np.float32)
        values = np.ascontiguousarray(values, np.float32)
        A.check(A.lib().mtsb_film_put(self._h, sensor, pos.shape[0], C.c_void_p(pos.ctypes.data), C.c_void_p(values.ctypes.data)))

    # ---- parameter updates (traverse()/parameters_changed) ------------------------------------------
    def update_vertices(self, mesh_id, positions):
        self._param_stamps.pop(("vertices", mesh_id), None)
        p, mem, keep = _arg(positions, np.float32, device=self.device)
        A.check(A.lib().mtsb_scene_update_vertices(self._h, mesh_id, p, mem))
        if mem == A.MEM_HOST:
            self.data.meshes[mesh_id].positions = keep.reshape(-1, 3).copy()

    def update_config(self, refit=False, rebuild_every=0):
        """How `update_vertices` brings the BVH up to date: full rebuild (default) or a refit of the last build's topology,
        with a full rebuild on every `rebuild_every`-th update (0 = never)."""
        A.check(A.lib().mtsb_scene_update_config(self._h, int(bool(refit)), int(rebuild_every)))

    def update_texture(self, tex_id, data):
        self._param_stamps.pop(("texture", tex_id), None)
        p, mem, keep = _arg(data, np.float32, device=self.device)
        A.check(A.lib().mtsb_scene_update_texture(self._h, tex_id, p, mem))

    # ---- gradient pass ----------------------------------------------------------------------------
    def grad_config(self, boundary=2, edge_fraction=0.25):
        """Options of the vertex-gradient pass: `boundary` 0 / False = interior term only, 1 = + edge sampling on the film
        (primary visibility), 2 / True = + edge sampling seen from the first path vertex (shadows, occlusion), 3 = + the same
        term at every later path vertex (a shadow seen through a bounce);
        `edge_fraction` = edge samples of each kind per path sample."""
        boundary = 2 if boundary is True else int(boundary)
        A.check(A.lib().mtsb_grad_config(self._h, boundary, float(edge_fraction)))

    def render_grad(self, dL_dimage, sensor=0, spp=None, seed=0, tex_id=-1, mesh_id=-1, spp_begin=0, spp_total=None):
        """Adjoint of `render`: returns (d/d texture `tex_id`, d/d vertex positions of mesh `mesh_id`) of
        sum(dL_dimage * developed image); either is None when its id is negative."""
        spp = self.data.spp if spp is None else int(spp)
        spp_total = spp_begin + spp if spp_total is None else int(spp_total)
        pd, mem, k = _arg(dL_dimage, np.float32, device=self.device)
        gt = gv = None
        if mem == A.MEM_DEVICE:
            import torch
            dev = k.device
            if tex_id >= 0:
                t = self.data.textures[tex_id]
                n = 3 if t["type"] == A.TEX_CONSTANT else t["data"].size
                gt = torch.zeros(n, dtype=torch.float32, device=dev)
            if mesh_id >= 0:
                gv = torch.zeros(self.data.meshes[mesh_id].positions.size, dtype=torch.float32, device=dev)
            pgt = C.c_void_p(gt.data_ptr()) if gt is not None else None
            pgv = C.c_void_p(gv.data_ptr()) if gv is not None else None
            torch.cuda.current_stream(dev).synchronize()      # the zero fills above run on torch's stream, the kernels on the library's
        else:
            if tex_id >= 0:
                t = self.data.textures[tex_id]
                gt = np.zeros(3 if t["type"] == A.TEX_CONSTANT else t["data"].size, np.float32)
            if mesh_id >= 0:
                gv = np.zeros(self.data.meshes[mesh_id].positions.size, np.float32)
            pgt = C.c_void_p(gt.ctypes.data) if gt is not None else None
            pgv = C.c_void_p(gv.ctypes.data) if gv is not None else None
        st = A.Stats()
        A.check(A.lib().mtsb_render_grad(self._h, sensor, spp_begin, spp, spp_total, int(seed), pd, tex_id, pgt, mesh_id, pgv,
                                         mem, C.byref(st)))
        self.last_stats = st.as_dict()
        if gv is not None:
            gv = gv.reshape(-1, 3)
        return gt, gv

    # ---- parity hooks -----------------------------------------------------------------------------
    def sensor_sample_ray(self, pos, sensor=0):
        pos = np.ascontiguousarray(pos, np.float32)
        n = pos.shape[0]
        o, d = np.empty((n, 4), np.float32), np.empty((n, 4), np.float32)
        A.check(A.lib().mtsb_sensor_sample_ray(self._h, sensor, n, C.c_void_p(pos.ctypes.data), C.c_void_p(o.ctypes.data), C.c_void_p(d.ctypes.data)))
        return o, d

    def bsdf_sample(self, bsdf_id, wi, uv, sample1, sample2):
        wi, uv, s1, s2 = (np.ascontiguousarray(x, np.float32) for x in (wi, uv, sample1, sample2))
        n = wi.shape[0]
        out = np.empty((n, 9), np.float32)
        A.check(A.lib().mtsb_bsdf_sample(self._h, bsdf_id, n, *(C.c_void_p(x.ctypes.data) for x in (wi, uv, s1, s2, out))))
        return out

    def bsdf_eval_pdf(self, bsdf_id, wi, uv, wo):
        wi, uv, wo = (np.ascontiguousarray(x, np.float32) for x in (wi, uv, wo))
        n = wi.shape[0]
        ev, pdf = np.empty((n, 3), np.float32), np.empty((n,), np.float32)
        A.check(A.lib().mtsb_bsdf_eval_pdf(self._h, bsdf_id, n, *(C.c_void_p(x.ctypes.data) for x in (wi, uv, wo, ev, pdf))))
        return ev, pdf

    def sample_emitter_direction(self, ref_p, sample2):
        p, s2 = np.ascontiguousarray(ref_p, np.float32), np.ascontiguousarray(sample2, np.float32)
        n = p.shape[0]
        out = np.empty((n, 11), np.float32)
        A.check(A.lib().mtsb_sample_emitter_direction(self._h, n, *(C.c_void_p(x.ctypes.data) for x in (p, s2, out))))
        return out


def pcg32_floats(initstate, initseq, n_draws, device=0):
    s = np.ascontiguousarray(initstate, np.uint64)
    q = np.ascontiguousarray(initseq, np.uint64)
    out = np.empty((s.size, n_draws), np.float32)
    A.check(A.lib().mtsb_pcg32_floats(_context(device), s.size, C.c_void_p(s.ctypes.data), C.c_void_p(q.ctypes.data), n_draws, C.c_void_p(out.ctypes.data)))
    return out


def sampler_seed(n, seed, device=0):
    st, inc = np.empty(n, np.uint64), np.empty(n, np.uint64)
    A.check(A.lib().mtsb_sampler_seed(_context(device), n, int(seed), C.c_void_p(st.ctypes.data), C.c_void_p(inc.ctypes.data)))
    return st, inc
