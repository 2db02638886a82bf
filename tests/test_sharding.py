"""CPU, world_size 2 over gloo: the batch-sharding host logic of the multi-GPU path.  The layer has
no cross-image coupling, so a rank's shard must reproduce the corresponding slice of the whole-batch
result exactly (checked here with the CPU oracle standing in for the kernel), the shards must tile
the batch, and the job time is the max over ranks."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from curl_b200 import parallel


def test_shard_bounds_tile_the_batch():
    for n in (0, 1, 7, 8, 64, 65):
        for w in (1, 2, 3, 8):
            spans = [parallel.shard_bounds(n, r, w) for r in range(w)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        parallel.shard_bounds(8, 2, 2)


def _worker(rank, world, port, q):
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, root)
    from oracle import curl_oracle as O
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.set_num_threads(2)
        gen = torch.Generator().manual_seed(5)          # same seed: every rank builds the same batch
        B, H, W = 5, 12, 10
        img = torch.rand(B, 3, H, W, generator=gen)
        mask = torch.rand(B, 1, H, W, generator=gen) > 0.1
        L, R, Hk = (0.1 * torch.randn(B, n, generator=gen) for n in (48, 48, 64))
        gout = torch.randn(B, 3, H, W, generator=gen)
        greg = torch.ones(B)
        whole = O.curl_layer_fwd_bwd(img, mask, L, R, Hk, gout, greg)
        lo, hi = parallel.shard_bounds(B, rank, world)
        s = parallel.shard_batch([img, mask, L, R, Hk, gout, greg], rank, world)
        mine = O.curl_layer_fwd_bwd(*s)
        ok = all(torch.allclose(mine[k], whole[k][lo:hi], rtol=0, atol=2e-6 * max(1.0, whole[k].abs().max().item()))
                 for k in mine)
        # (not bitwise on the CPU oracle: ATen's vectorised pow takes a scalar tail path that depends on
        # the buffer length; the CUDA kernels ARE bitwise shard-invariant, see tests/test_gpu_parity.py)
        # pad shards to equal size for the all_gather helper, then reassemble
        pad = torch.zeros(3, *mine["out"].shape[1:])
        pad[: hi - lo] = mine["out"]
        allp = parallel.gather_shards(pad, world)
        spans = [parallel.shard_bounds(B, r, world) for r in range(world)]
        rebuilt = torch.cat([allp[3 * r: 3 * r + (b - a)] for r, (a, b) in enumerate(spans)])
        ok = ok and torch.allclose(rebuilt, whole["out"], rtol=0, atol=2e-6)
        t = parallel.max_over_ranks(10.0 + rank)
        ok = ok and t == 10.0 + world - 1
        q.put((rank, bool(ok)))
    finally:
        dist.destroy_process_group()


def test_two_rank_sharding_gloo():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=240) for _ in procs)
    for p in procs:
        p.join(timeout=60)
    assert res == [(0, True), (1, True)]
