"""terra_b200_trace_rays_packed_noalloc (include/terra_b200.h): the closest-hit entry point for callers that cannot
allocate device memory. Its scratch — float4 ray records and the fetch cursor — is a file-scope __device__ array, so
the checks are: same hits as terra_b200_trace_rays_device bit for bit (the bar of tests/test_gpu_parity.py), on two
calls in a row with different batches (the array keeps the first call's contents), no change of the device's free
memory across a call, and a clean refusal above TERRA_B200_NOALLOC_MAX_RAYS.

The file name sorts last on purpose. The test builds its own scene and passes both
alone in a fresh process and after the rest of the suite."""
import ctypes as C

import numpy as np
import pytest

import cases

pytestmark = pytest.mark.gpu

MAX_RAYS = 262144          # TERRA_B200_NOALLOC_MAX_RAYS


def _hits_existing(api, gpu, s, torch, o, d):
    n = o.shape[0]
    o4 = torch.zeros(n, 4, device="cuda"); d4 = torch.zeros(n, 4, device="cuda")
    o4[:, :3] = torch.from_numpy(o).cuda(); d4[:, :3] = torch.from_numpy(d).cuda()
    out = torch.full((n, 4), -7.0, device="cuda")
    torch.cuda.synchronize()
    gpu.terra_b200_trace_rays_device(s.h, n, C.c_void_p(o4.data_ptr()), C.c_void_p(d4.data_ptr()), C.c_void_p(out.data_ptr()), None)
    torch.cuda.synchronize()
    return out.cpu().numpy().view(api.HIT_DTYPE).reshape(n)


def test_no_alloc_entry_point_matches_the_allocating_one_twice(api, scenes, gpu):
    import torch
    s, _ = scenes.sphere_field(gpu, n_spheres=24, n_lat=10, n_lon=12, extent=30.0, seed=3, ggx=False)
    batches = [cases.rays_for(s, api, scenes, 200_000, seed=11), cases.rays_for(s, api, scenes, 70_001, seed=12)]   # second: shorter, odd
    # every device buffer of the test exists before the first call
    dev = []
    for o, d in batches:
        o3 = torch.from_numpy(np.ascontiguousarray(o, np.float32)).cuda().contiguous()
        d3 = torch.from_numpy(np.ascontiguousarray(d, np.float32)).cuda().contiguous()
        dev.append((o3, d3, torch.full((o.shape[0], 4), -7.0, device="cuda")))
    want = [_hits_existing(api, gpu, s, torch, np.ascontiguousarray(o, np.float32), np.ascontiguousarray(d, np.float32)) for o, d in batches]
    torch.cuda.synchronize()
    for k, (o3, d3, out) in enumerate(dev):
        n = o3.shape[0]
        free_before = torch.cuda.mem_get_info()[0]
        rc = gpu.terra_b200_trace_rays_packed_noalloc(s.h, n, C.c_void_p(o3.data_ptr()), C.c_void_p(d3.data_ptr()), C.c_void_p(out.data_ptr()), None)
        free_after = torch.cuda.mem_get_info()[0]
        assert rc == 0
        if k == 1:                                  # call 0 may load the kernels' module lazily
            assert free_after == free_before, "the call changed the device's free memory"
        got = out.cpu().numpy().view(api.HIT_DTYPE).reshape(n)
        assert (want[k]["t"] >= 0).mean() > 0.05, "the batch hits nothing: the case tests nothing"
        for f in ("t", "u", "v"):
            assert np.array_equal(got[f].view(np.uint32), want[k][f].view(np.uint32)), (k, f)
        assert np.array_equal(got["prim"], want[k]["prim"]), k
    # above the limit: refused, nothing written
    n_big = MAX_RAYS + 1
    big = torch.zeros(n_big, 3, device="cuda")
    out = torch.full((n_big, 4), -7.0, device="cuda")
    torch.cuda.synchronize()
    rc = gpu.terra_b200_trace_rays_packed_noalloc(s.h, n_big, C.c_void_p(big.data_ptr()), C.c_void_p(big.data_ptr()), C.c_void_p(out.data_ptr()), None)
    torch.cuda.synchronize()
    assert rc == -1
    assert bool((out == -7.0).all())
    # and the scratch is still good for a third call
    o3, d3, out = dev[0]
    out.fill_(-7.0)
    torch.cuda.synchronize()
    assert gpu.terra_b200_trace_rays_packed_noalloc(s.h, o3.shape[0], C.c_void_p(o3.data_ptr()), C.c_void_p(d3.data_ptr()), C.c_void_p(out.data_ptr()), None) == 0
    got = out.cpu().numpy().view(api.HIT_DTYPE).reshape(-1)
    assert np.array_equal(got["t"].view(np.uint32), want[0]["t"].view(np.uint32))
    s.destroy()
