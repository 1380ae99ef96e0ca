"""The N>1 host logic on CPU: world_size-2 gloo process group, the same partition + ONE reduce(sum)
that bench.py / DistRenderer run over NCCL. Partial framebuffers are produced by the oracle here
(the checker standing in for the GPU render), so the test also shows the split is correct:
row-band split + reduce is bit-identical to one process; sample split sums the sample counts."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import REPO


def test_split_samples_and_rows():
    from terra_b200.multigpu import split_rows, split_samples
    for spp in (1, 7, 16, 256, 1023):
        for world in (1, 2, 3, 4, 8):
            parts = [split_samples(spp, world, r) for r in range(world)]
            assert sum(c for _, c in parts) == spp
            assert parts[0][0] == 0
            for (b0, c0), (b1, _) in zip(parts, parts[1:]):
                assert b1 == b0 + c0
            assert max(c for _, c in parts) - min(c for _, c in parts) <= 1
    for h in (1, 45, 1080, 2160):
        for world in (1, 2, 4, 8):
            rows = sorted(y for r in range(world) for (y0, hh) in split_rows(h, world, r) for y in range(y0, y0 + hh))
            assert rows == list(range(h))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, mode, out_path):
    sys.path.insert(0, str(REPO))
    sys.path.insert(0, str(REPO / "tests"))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import cases
    from terra_b200 import scenes
    from oracle.binding import harness_api
    api = harness_api()
    from terra_b200.multigpu import float_as_results, reduce_partials, results_as_float, split_rows, split_samples
    prod, oracle = api.TerraLib(), api.OracleLib()
    W = H = 24
    spp = 4
    with cases.no_commit(api):
        s, cam = scenes.cornell(prod, spp=spp, jitter=0.0, integrator=0)
    partial = np.zeros(W * H, api.RESULT_DTYPE)
    if mode == "rows":
        osc = api.OracleScene(oracle, s)
        for y, hh in split_rows(H, world, rank, band=4):
            api.srand(1000 + y)                    # per-band stream so every process draws what one process would
            res, _ = osc.render(cam, W, H, region=(0, y, W, hh))
            partial.reshape(H, W)[y:y + hh] = res[y:y + hh]
    else:
        base, count = split_samples(spp, world, rank)
        s.options().samples_per_pixel = count
        osc = api.OracleScene(oracle, s)
        api.srand(77 + rank)
        res, _ = osc.render(cam, W, H)
        partial[:] = res.reshape(-1)
    t = torch.from_numpy(partial.view(np.float32).copy())
    f = results_as_float(t)
    reduce_partials(f, dist)
    if rank == 0:
        np.save(out_path, float_as_results(f).numpy())
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("mode", ["rows", "samples"])
def test_two_rank_reduce_matches_single_process(tmp_path, mode):
    sys.path.insert(0, str(REPO / "tests"))
    import cases
    from terra_b200 import scenes
    from oracle.binding import harness_api
    api = harness_api()
    out = tmp_path / "reduced.npy"
    mp.spawn(_worker, args=(2, _free_port(), mode, str(out)), nprocs=2, join=True)
    got = np.load(out).view(api.RESULT_DTYPE).reshape(24, 24)
    prod, oracle = api.TerraLib(), api.OracleLib()
    with cases.no_commit(api):
        s, cam = scenes.cornell(prod, spp=4, jitter=0.0, integrator=0)
    osc = api.OracleScene(oracle, s)
    if mode == "rows":
        want = np.zeros((24, 24), api.RESULT_DTYPE)
        for y in range(0, 24, 4):
            api.srand(1000 + y)
            res, _ = osc.render(cam, 24, 24, region=(0, y, 24, 4))
            want[y:y + 4] = res[y:y + 4]
        assert np.array_equal(got["samples"], want["samples"])
        assert np.array_equal(got["acc"].view(np.uint32), want["acc"].view(np.uint32))     # bit-exact
    else:
        assert (got["samples"] == 4).all()
        api.srand(5)
        res, _ = osc.render(cam, 24, 24)
        # different random streams: only the statistics can agree
        a, b = got["acc"].mean(axis=(0, 1)) / 4, res["acc"].mean(axis=(0, 1)) / 4
        assert np.allclose(a, b, rtol=0.25)
