"""Host-side helpers for the one-process-per-GPU layout.

Dialogue turns (batch rows) are independent on the whole video path — attention never mixes samples and
LayerNorm is per row — so the path shards by rows with NO data-path collective.  The only communication is
bookkeeping (the max-over-ranks step time of bench.py and, in tests, gathering the shards back) and, for the
training step, ONE gradient all-reduce per step (`allreduce_grads`).
"""
import torch
import torch.distributed as dist

from .synth import SynthBatch, shard_rows


def shard_batch(b, encoded_query, world_size, rank):
    """Contiguous, balanced row shard of a batch and of ft['encoded_query'] for `rank`."""
    n = b.fts.shape[0]
    lo, hi = shard_rows(n, world_size, rank)
    sb = SynthBatch()
    for k, v in b.__dict__.items():
        setattr(sb, k, v[lo:hi] if torch.is_tensor(v) and v.dim() > 0 and v.shape[0] == n else v)
    return sb, encoded_query[lo:hi], (lo, hi)


def max_over_ranks(value, device="cpu"):
    """Step time of a multi-rank run = the slowest rank (bench.py)."""
    t = torch.tensor([float(value)], device=device)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def gather_rows(local, n_rows):
    """All-gather row shards produced by shard_batch back into [n_rows, ...] order (tests only)."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return local
    world = dist.get_world_size()
    out = []
    for r in range(world):
        lo, hi = shard_rows(n_rows, world, r)
        buf = torch.empty((hi - lo,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
        if r == dist.get_rank():
            buf.copy_(local)
        dist.broadcast(buf, src=r)
        out.append(buf)
    return torch.cat(out, 0)


def allreduce_grads(params, average=True):
    """Sum (or average) the gradients of `params` over all ranks with as few collectives as possible.

    The backward of a VidEncoderLayer4 writes all of its parameter gradients into one flat fp32 block
    (include/bist_b200.h, "Gradient block") and hands out views, so every layer is ONE all-reduce over the
    views' common base; stand-alone gradients are flattened into one extra bucket.  Returns the number of
    collectives issued.  A no-op (returns 0) without an initialised process group or with world_size 1.
    """
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return 0
    world = dist.get_world_size()
    blocks, loose = {}, []
    for p in params:
        g = p.grad
        if g is None:
            continue
        base = g._base if g._is_view() and g._base is not None and g._base.is_contiguous() else None
        if base is not None:
            blocks.setdefault(id(base), [base, 0])[1] += g.numel()
        else:
            loose.append(g)
    n = 0
    for base, covered in blocks.values():
        if covered == base.numel():          # the block holds nothing but gradients: reduce it in place
            dist.all_reduce(base)
            if average:
                base.div_(world)
            n += 1
        else:                                # partially used base: fall back to its views
            loose.extend(p.grad for p in params if p.grad is not None and p.grad._is_view() and p.grad._base is base)
    if loose:
        flat = torch.cat([g.reshape(-1) for g in loose])
        dist.all_reduce(flat)
        if average:
            flat.div_(world)
        off = 0
        for g in loose:
            g.copy_(flat[off:off + g.numel()].view_as(g))
            off += g.numel()
        n += 1
    return n
