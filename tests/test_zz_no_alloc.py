"""mtsb_intersect_closest_noalloc (no counterpart in the reference): the allocation-free entry point
must give the hits of mtsb_intersect_closest -- same kernel, so bit-identical -- on consecutive calls that share the fixed
device scratch, must not change the amount of free device memory, and must reject batches beyond its stated maximum.
Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

import helpers as H

pytestmark = pytest.mark.gpu


def _call(lib, A, scene, om, dm, mem_device):
    """-> hits (n,4) float32 numpy, through the C ABI with host or device pointers."""
    n = om.shape[0]
    st = A.Stats()
    if mem_device:
        import torch
        o_t, d_t = torch.from_numpy(om).cuda(), torch.from_numpy(dm).cuda()
        h_t = torch.full((n, 4), -7.0, dtype=torch.float32, device="cuda")
        torch.cuda.synchronize()
        rc = lib.mtsb_intersect_closest_noalloc(scene._h, n, C.c_void_p(o_t.data_ptr()), C.c_void_p(d_t.data_ptr()),
                                                C.c_void_p(h_t.data_ptr()), A.MEM_DEVICE, C.byref(st))
        A.check(rc)
        return h_t.cpu().numpy(), st
    hits = np.full((n, 4), -7.0, np.float32)
    rc = lib.mtsb_intersect_closest_noalloc(scene._h, n, C.c_void_p(om.ctypes.data), C.c_void_p(dm.ctypes.data),
                                            C.c_void_p(hits.ctypes.data), A.MEM_HOST, C.byref(st))
    A.check(rc)
    return hits, st


@pytest.mark.parametrize("mem_device", [False, True])
def test_noalloc_matches_allocating_entry_point_twice(mem_device):
    import torch
    import mitsuba2_b200 as mi
    from mitsuba2_b200 import _abi as A
    mi.set_variant("cuda_b200_rgb")
    lib = A.lib()
    sd = H.sphere_scene(64, 128)
    scene = mi.Scene(sd)
    assert lib.mtsb_noalloc_max_rays() == 1 << 18
    # two different batches in a row: the second one is SMALLER and different, so stale rays / hits / a stale work
    # counter left in the shared scratch by the first call would show
    batches = [H.random_rays(50000, 11, radius=1.5), H.random_rays(20001, 12, radius=1.2)]
    want = [scene.intersect_closest_raw(om, dm).copy() for om, dm in batches]
    _call(lib, A, scene, *batches[0], mem_device)                     # warm-up: module load, symbol resolution
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    got = [_call(lib, A, scene, om, dm, mem_device) for om, dm in batches]
    torch.cuda.synchronize()
    free1 = torch.cuda.mem_get_info()[0]
    for (h, st), w in zip(got, want):
        # integer / same-kernel work: bit-exact, the tolerance of the repo's hit-id tests
        assert np.array_equal(h.view(np.uint32), w.view(np.uint32))
        assert st.closest_rays == h.shape[0] and st.kernel_launches == 1
    assert np.isfinite(want[0][:, 0]).sum() > 1000                    # the batches do hit the mesh
    if not mem_device:                                                # (the device-pointer variant allocates torch tensors in _call)
        assert free1 == free0, f"free device memory changed across allocation-free calls: {free0} -> {free1}"
    # the largest batch the scratch holds, then one ray more
    nmax = int(lib.mtsb_noalloc_max_rays())
    om, dm = H.random_rays(nmax, 13, radius=1.5)
    h, _ = _call(lib, A, scene, om, dm, mem_device)
    assert np.array_equal(h.view(np.uint32), scene.intersect_closest_raw(om, dm).view(np.uint32))
    om2, dm2 = np.concatenate([om, om[:1]]), np.concatenate([dm, dm[:1]])
    hits = np.empty((nmax + 1, 4), np.float32)
    rc = lib.mtsb_intersect_closest_noalloc(scene._h, nmax + 1, C.c_void_p(om2.ctypes.data), C.c_void_p(dm2.ctypes.data),
                                            C.c_void_p(hits.ctypes.data), A.MEM_HOST, None)
    assert rc == -1                                                    # MTSB_E_INVALID
    msg = lib.mtsb_last_error().decode()
    assert "262144" in msg and "exceed" in msg, msg
    # and a call after the rejected one still works
    h, _ = _call(lib, A, scene, *batches[1], mem_device)
    assert np.array_equal(h.view(np.uint32), want[1].view(np.uint32))
    # n == 0 is a no-op
    assert lib.mtsb_intersect_closest_noalloc(scene._h, 0, None, None, None, A.MEM_HOST, None) == 0
