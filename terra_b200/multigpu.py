"""Multi-GPU rendering: one process per GPU (torch.distributed, NCCL over NVLink/NVSwitch).

The path shards trivially: every (pixel, sample) is independent and the only shared state is the
per-pixel accumulator, which is a sum (reference src/Terra.c:532-572). So the scene is replicated
on every GPU, the work is split either by SAMPLE RANGE (rank r renders samples [base_r, base_r+n_r)
of every pixel; perfect balance; float sums re-associate) or by ROW BANDS (rank r renders bands
r, r+G, ...; the reduce then adds zeros and is bit-exact against one GPU), and the partial
framebuffers are combined with ONE reduce(sum) to rank 0 per frame. No other collective exists
on this path, and none is invented. torch is plumbing here: device buffers, streams, the collective.
"""
from __future__ import annotations

import ctypes as C

import numpy as np


def split_samples(spp: int, world: int, rank: int) -> tuple[int, int]:
    """(first sample index, sample count) of `rank`; counts differ by at most 1 and sum to spp."""
    q, r = divmod(spp, world)
    count = q + (1 if rank < r else 0)
    base = rank * q + min(rank, r)
    return base, count


def split_rows(height: int, world: int, rank: int, band: int = 8) -> list[tuple[int, int]]:
    """Interleaved row bands [(y, h), ...] of `rank`: band b (rows [b*band, (b+1)*band)) goes to b % world."""
    out = []
    for b, y in enumerate(range(0, height, band)):
        if b % world == rank:
            out.append((y, min(band, height - y)))
    return out


def results_as_float(results):
    """TerraRawIntegrationResult[n] {float3 acc; int32 samples} as an (n,4) float tensor whose 4th
    column holds the sample count as a float (exact below 2^24), so one float reduce carries both."""
    import torch
    f = results.view(torch.float32).view(-1, 4).clone()
    f[:, 3] = results.view(torch.int32).view(-1, 4)[:, 3].to(torch.float32)
    return f


def float_as_results(f):
    import torch
    out = f.clone()
    out.view(torch.int32)[:, 3] = f[:, 3].round().to(torch.int32)
    return out


def reduce_partials(partial, dist=None, dst: int = 0):
    """ONE collective: sum the (n,4) float partial framebuffers onto rank `dst` (in place)."""
    if dist is None:
        import torch.distributed as dist
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.reduce(partial, dst=dst, op=dist.ReduceOp.SUM)
    return partial


class DistRenderer:
    """Renders a full frame across the ranks of an initialised process group; rank 0 ends up with the
    reference's framebuffer contents (results[] and tonemapped pixels[])."""

    def __init__(self, lib, scene, cam, width: int, height: int, mode: str = "samples", band: int | None = None):
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        self.lib, self.scene, self.cam = lib, scene, cam
        self.rank = dist.get_rank() if dist.is_initialized() else 0
        self.world = dist.get_world_size() if dist.is_initialized() else 1
        if band is None:
            # about four interleaved bands per rank (balance against per-call launch overhead: a band is
            # one terra_b200_render_device call), a multiple of 4 rows so the 8x4 tiled wave layout applies
            band = max(8, -(-height // (self.world * 4) // 4) * 4)
        self.w, self.h, self.mode, self.band = width, height, mode, band
        self.dev = torch.device("cuda", torch.cuda.current_device())
        self.results = torch.zeros(width * height * 4, dtype=torch.float32, device=self.dev)
        self.pixels = torch.zeros(width * height * 3, dtype=torch.float32, device=self.dev)
        self.host_results = torch.zeros(width * height * 4, dtype=torch.float32).pin_memory()
        self.host_pixels = torch.zeros(width * height * 3, dtype=torch.float32).pin_memory()

    def render(self, spp: int):
        """Adds `spp` samples per pixel (split over the ranks) and reduces. Returns on rank 0 the
        (results, pixels) device tensors; other ranks get (None, None)."""
        torch = self.torch
        stream = torch.cuda.current_stream()
        sp = C.c_void_p(stream.cuda_stream)
        self.results.zero_()
        if self.mode == "samples":
            base, count = split_samples(spp, self.world, self.rank)
            if count:
                self.lib.terra_b200_render_device(C.byref(self.cam), self.scene.h, self.w, self.h, 0, 0, self.w, self.h,
                                                  base, count, C.c_void_p(self.results.data_ptr()), sp)
        else:
            for y, hh in split_rows(self.h, self.world, self.rank, self.band):
                self.lib.terra_b200_render_device(C.byref(self.cam), self.scene.h, self.w, self.h, 0, y, self.w, hh,
                                                  0, spp, C.c_void_p(self.results.data_ptr()), sp)
        if self.world > 1:
            # ONE collective, in place: the accumulators are floats; the int32 `samples` lane of every record is known
            # without communication (the frame started from zero and every pixel received exactly `spp` samples over all
            # ranks), so it is blanked before the sum and written back on rank 0 after it. Two strided fills around
            # the reduce instead of the clone / convert / round / copy chain of round 1 (8 kernels over 33 MB).
            rec = self.results.view(-1, 4)
            rec[:, 3] = 0.0
            reduce_partials(rec, self.dist)
            if self.rank == 0:
                self.results.view(torch.int32).view(-1, 4)[:, 3] = int(spp)
        if self.rank != 0:
            return None, None
        self.lib.terra_b200_resolve_device(self.scene.h, self.w, self.h, C.c_void_p(self.results.data_ptr()),
                                           C.c_void_p(self.pixels.data_ptr()), sp)
        return self.results, self.pixels

    def to_host(self):
        """Device -> pinned host copy of rank 0's framebuffer (what terra_render hands the caller)."""
        self.host_results.copy_(self.results, non_blocking=True)
        self.host_pixels.copy_(self.pixels, non_blocking=True)
        self.torch.cuda.current_stream().synchronize()
        res = self.host_results.numpy().view(np.dtype([("acc", "<f4", (3,)), ("samples", "<i4")])).reshape(self.h, self.w)
        return res, self.host_pixels.numpy().reshape(self.h, self.w, 3)
