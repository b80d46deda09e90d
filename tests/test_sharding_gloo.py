"""N > 1 host logic on CPU: world_size-2 gloo runs of the pieces of droid_backends.ba_sharded that do not need a
GPU -- the source-frame partition, the all-reduce of the packed fp64 reduced camera system and the all-gather of the
per-rank depth shards.  The CUDA part of ba_sharded is covered by `bench.py --gpus 2` on the GPU box."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, nbuf, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import droid_backends as B
    ii = torch.from_numpy(np.random.default_rng(3).integers(0, nbuf, size=20 * nbuf))
    bounds = B.shard_bounds(ii, nbuf, world)
    lo, hi = bounds[rank], bounds[rank + 1]
    per = max(bounds[r + 1] - bounds[r] for r in range(world))
    # (1) partial systems: rank r contributes only the source frames of its shard; the sum must equal the whole
    rng = np.random.default_rng(7)
    n6 = 30
    per_frame = rng.normal(size=(nbuf, n6 * n6 + n6))              # what frame k adds to (H - S | v)
    part = torch.from_numpy(per_frame[lo:hi].sum(0))
    dist.all_reduce(part)
    ok_sum = np.allclose(part.numpy(), per_frame.sum(0), rtol=1e-12, atol=1e-12)
    # (2) every rank updates only its own depth maps, then the shards are all-gathered
    disps = torch.zeros(nbuf, 4, 6)
    full = torch.arange(nbuf * 24, dtype=torch.float32).view(nbuf, 4, 6)
    disps[lo:hi] = full[lo:hi]
    B.gather_depth_shards(disps, bounds, rank)
    ok_gather = torch.equal(disps, full)
    q.put((rank, lo, hi, per, bool(ok_sum), bool(ok_gather)))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("nbuf", [9, 16])
def test_sharded_host_logic_world2(lib_built, nbuf):
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, nbuf, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    covered = []
    for rank, lo, hi, per, ok_sum, ok_gather in res:
        assert ok_sum and ok_gather
        covered += list(range(lo, hi))
    assert covered == list(range(nbuf))                            # disjoint, contiguous, complete


def test_shard_range_properties(backends):
    for nbuf in (1, 7, 512, 1000, 1024):
        for world in (1, 2, 4, 8):
            seen = []
            for r in range(world):
                lo, hi, per = backends.shard_range(nbuf, world, r)
                assert 0 <= lo <= hi <= nbuf and hi - lo <= per
                seen += list(range(lo, hi))
            assert seen == list(range(nbuf))


def test_shard_bounds_balance_the_work(backends):
    """Work-balanced contiguous shards: cuts are monotone, cover every frame once, and no shard carries more than
    its share of the cost plus one frame's worth."""
    rng = np.random.default_rng(5)
    nbuf = 300
    deg = rng.integers(4, 32, size=nbuf)
    ii = torch.from_numpy(np.repeat(np.arange(nbuf), deg))
    cost = (deg * (deg + 12.0))
    for world in (1, 2, 4, 8):
        b = backends.shard_bounds(ii, nbuf, world)
        assert len(b) == world + 1 and b[0] == 0 and b[-1] == nbuf and all(b[i] <= b[i + 1] for i in range(world))
        share = cost.sum() / world
        for r in range(world):
            assert cost[b[r]:b[r + 1]].sum() <= share + cost.max() + 1e-9


def test_shard_edge_rows_index_the_ranks_own_rows(backends):
    """A rank that holds only the targets / weights rows of its own edges: edge_row maps every edge of the shard to
    its position among the shard's edges, in edge order; together the shards cover every edge exactly once."""
    rng = np.random.default_rng(2)
    nbuf, ne, world = 40, 500, 4
    ii = torch.from_numpy(rng.integers(0, nbuf, size=ne))
    bounds = backends.shard_bounds(ii, nbuf, world)
    seen = torch.zeros(ne, dtype=torch.int32)
    for r in range(world):
        lo, hi = bounds[r], bounds[r + 1]
        mask = backends.shard_edges(ii, lo, hi)
        rows, count = backends.shard_edge_rows(ii, lo, hi)
        assert rows.dtype == torch.int32 and count == int(mask.sum())
        assert torch.equal(rows[mask].long(), torch.arange(count))          # dense, ascending edge order
        seen += mask.int()
    assert torch.equal(seen, torch.ones(ne, dtype=torch.int32))
