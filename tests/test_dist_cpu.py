"""world_size-2 gloo test of the row-sharded layout (CPU): each rank runs the hot path on its shard of dialogue
turns with no data-path collective; gathered results equal the unsharded run, and the step time reduction is a MAX."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from bist_b200 import synth
from bist_b200.dist import allreduce_grads, gather_rows, max_over_ranks, shard_batch
from oracle import bist_oracle as O

C = dict(seed=11, B=5, T=4, S=3, D=24, d=32, h=4, Lq=3, n_blocks=1)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.set_num_threads(1)
    c = C
    P = synth.make_params(c["seed"], c["d"], c["D"], c["n_blocks"])
    b = synth.make_batch(c["seed"], c["B"], c["T"], c["S"], c["D"], c["Lq"])
    eq = synth.make_encoded_query(c["seed"], c["B"], c["Lq"], c["d"])
    sb, seq, (lo, hi) = shard_batch(b, eq, world, rank)
    with torch.no_grad():
        loc = O.video_path(P, c["h"], sb.fts, seq, sb.query_mask, sb.temporal_mask, c["n_blocks"])
        full = O.video_path(P, c["h"], b.fts, eq, b.query_mask, b.temporal_mask, c["n_blocks"])
    errs = {k: O.norm_err(gather_rows(loc[k], c["B"]), full[k]) for k in ("temporal_ft", "spatial_ft", "spatiotemporal_ft")}
    tmax = max_over_ranks(1.0 + rank)
    if rank == 0:
        q.put((errs, tmax, hi - lo))
    dist.barrier()
    dist.destroy_process_group()


def test_row_sharding_two_ranks_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    errs, tmax, n0 = q.get(timeout=120)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    assert all(e < 1e-6 for e in errs.values()), errs
    assert tmax == 2.0            # max over ranks, not rank 0's own time
    assert n0 == 3                # 5 rows over 2 ranks: 3 + 2


class _BlockFn(torch.autograd.Function):
    """What the layer backward does: one flat gradient block, slices of it returned as the parameter gradients."""

    @staticmethod
    def forward(ctx, x, a, b, scale):
        ctx.scale = scale
        return x * 1.0

    @staticmethod
    def backward(ctx, g):
        from bist_b200.dist import grad_block_ready
        block = torch.arange(10, dtype=torch.float32) * ctx.scale
        grad_block_ready(block)
        return g, block[:4].view(4), block[4:].view(2, 3), None


def _grad_worker(rank, world, port, q):
    from bist_b200.dist import overlap_grad_allreduce
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    res = []
    for overlap in (False, True):
        ps = [torch.nn.Parameter(torch.zeros(4)), torch.nn.Parameter(torch.zeros(2, 3)),
              torch.nn.Parameter(torch.zeros(5)), torch.nn.Parameter(torch.zeros(1))]
        x = torch.ones(3, requires_grad=True)
        loss = _BlockFn.apply(x, ps[0], ps[1], float(rank + 1)).sum() + (ps[2] * (rank + 1)).sum() + ps[3][0] * 10 * (rank + 1)
        if overlap:
            with overlap_grad_allreduce():
                loss.backward()
        else:
            loss.backward()
        # autograd adopts the slices without copying: both gradients still live in the block's storage
        shared = ps[0].grad.untyped_storage().data_ptr() == ps[1].grad.untyped_storage().data_ptr()
        n = allreduce_grads(ps, average=True)
        res.append((n, shared, [p.grad.clone() for p in ps]))
    if rank == 0:
        q.put(res)
    dist.barrier()
    dist.destroy_process_group()


def test_gradient_allreduce_two_ranks_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_grad_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = q.get(timeout=120)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    mean = torch.arange(10, dtype=torch.float32) * 1.5       # mean of 1x and 2x
    for n, shared, grads in res:                             # plain, then with the block's all-reduce started in backward
        assert shared
        assert n == 2                                        # one collective for the block, one for the loose bucket
        assert torch.equal(grads[0], mean[:4]) and torch.equal(grads[1], mean[4:].view(2, 3))
        assert torch.equal(grads[2], torch.full((5,), 1.5)) and torch.equal(grads[3], torch.full((1,), 15.0))
