"""CPU suite, part 3: the multi-GPU host logic without GPUs.

The j-particles are split into contiguous 64-aligned shards (gdy_shard_range, the arithmetic the CUDA
context uses); ranks agree on an NCCL id by broadcast and on timings by a max-reduction.  Here the same
plumbing runs over gloo with world_size 2: the shards tile [0, n) exactly, a rank-0 blob reaches every
rank, and partial sums added over ranks reproduce the full sum (the allreduce contract of
update_acc_jrk; reference's equivalent: src/include/mpi/Hermite4MPI.cpp:202)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from gravidy_b200.hermite4 import Hermite4B200


@pytest.mark.parametrize("n", [1, 63, 64, 65, 97, 1000, 16384, 262144, 1000003])
@pytest.mark.parametrize("world", [1, 2, 3, 4, 8])
def test_shards_tile_the_particle_range(n, world):
    prev = 0
    for rank in range(world):
        b, e = Hermite4B200.shard_range(n, rank, world)
        assert b == prev and e >= b
        assert b % 64 == 0 or b == n
        prev = e
    assert prev == n
    sizes = [np.subtract(*Hermite4B200.shard_range(n, r, world)[::-1]) for r in range(world)]
    if n >= 64 * 4 * world:
        assert max(sizes) - min(sizes) <= 256 + 64   # balanced to a block (the tail shard also loses the padding)


def _worker(rank, world, port, n, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        # (1) the NCCL-id hand-off of bench.py / the host program: 128 bytes from rank 0 to everybody
        blob = torch.zeros(128, dtype=torch.uint8)
        if rank == 0:
            blob.copy_(torch.from_numpy((np.arange(128) * 3 % 251).astype(np.uint8)))
        dist.broadcast(blob, 0)
        # (2) j-sharded partial sums + allreduce == full sum (float64, fixed data)
        rng = np.random.default_rng(7)
        w = rng.standard_normal((n, 6))
        b, e = Hermite4B200.shard_range(n, rank, world)
        part = torch.from_numpy(w[b:e].sum(axis=0))
        dist.all_reduce(part)
        # (3) max-over-ranks timing reduction
        t = torch.tensor([1.0 + rank], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        q.put((rank, bytes(blob.numpy()), part.numpy(), float(t.item()), (b, e)))
    finally:
        dist.destroy_process_group()


def test_world_size_2_gloo_plumbing():
    world, n = 2, 1000
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n, q)) for r in range(world)]
    for p in procs:
        p.start()
    out = sorted(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    expect_blob = bytes((np.arange(128) * 3 % 251).astype(np.uint8))
    full = np.random.default_rng(7).standard_normal((n, 6)).sum(axis=0)
    for rank, blob, part, tmax, (b, e) in out:
        assert blob == expect_blob
        np.testing.assert_allclose(part, full, rtol=1e-12)
        assert tmax == float(world)
    assert np.array_equal(out[0][2], out[1][2])   # bit-identical on every rank
    assert out[0][4][1] == out[1][4][0]
