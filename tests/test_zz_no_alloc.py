"""b200pt_intersect_closest_noalloc (include/b200pt.h): the closest-hit traversal with its scratch in a static
__device__ array instead of cudaMalloc'ed buffers - an entry point with no counterpart in the reference.
Checked through the C ABI against b200pt_intersect_closest on the same device
buffers: records identical byte for byte (the bar of the traversal tests), on two calls in a row (the array keeps
the previous call's contents), plus the size limit.  Collected last (file name) so that a failure here hides nothing."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _on_device(torch, a):
    return torch.from_numpy(a.view(np.uint8).reshape(-1).copy()).cuda()


@pytest.mark.parametrize("walk", ["binary", "wide"])
def test_noalloc_entry_matches_intersect_closest(P, monkeypatch, walk):
    """walk = wide: the 8-wide walk, which hands its irregular rays to the binary walk through the scratch list."""
    import torch
    monkeypatch.setenv("B200PT_TRAVERSAL", walk)
    capi = P.capi
    sc = P.scenes.random_soup(30000, seed=123)
    agg = P.B200Aggregate(sc, device=0)
    L = capi.lib()
    batches = [P.scenes.primary_rays(150000, seed=7),
               # second call: fewer rays, axis-aligned ones among them (zero direction components: the rays the wide
               # walk hands to the exact walk through the scratch list), over the first call's leftovers
               P.scenes.adversarial_rays(sc, 40000, seed=8)]
    for rays in batches:
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 7)
        n = rays.shape[0]
        d_rays = _on_device(torch, rays)
        d_ref = torch.zeros(n * capi.HIT_DTYPE.itemsize, dtype=torch.uint8, device="cuda")
        d_new = torch.full((n * capi.HIT_DTYPE.itemsize,), 0xAB, dtype=torch.uint8, device="cuda")
        torch.cuda.synchronize()
        agg.intersect_closest_ptr(d_rays.data_ptr(), n, d_ref.data_ptr(), capi.MEM_DEVICE)
        agg.sync()
        before = torch.cuda.memory_allocated()
        capi.check(L.b200pt_intersect_closest_noalloc(agg._scene, C.c_void_p(d_rays.data_ptr()), n,
                                                      C.c_void_p(d_new.data_ptr())), agg._ctx)
        agg.sync()
        assert torch.cuda.memory_allocated() == before
        ref = d_ref.cpu().numpy().view(capi.HIT_DTYPE)
        new = d_new.cpu().numpy().view(capi.HIT_DTYPE)
        assert (ref["mesh"] != 0xFFFFFFFF).mean() > 0.05
        assert np.array_equal(ref.view("u1"), new.view("u1"))
    # beyond the static scratch: refused with a message, nothing launched
    too_many = capi.NOALLOC_MAX_RAYS + 1
    rc = L.b200pt_intersect_closest_noalloc(agg._scene, C.c_void_p(d_rays.data_ptr()), too_many,
                                            C.c_void_p(d_new.data_ptr()))
    assert rc == capi.ERR_INVALID
    assert b"B200PT_NOALLOC_MAX_RAYS" in L.b200pt_last_error(agg._ctx)
    assert L.b200pt_intersect_closest_noalloc(agg._scene, None, 0, None) == 0
    agg.close()
