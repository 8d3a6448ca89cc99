"""Host-side helpers for the one-process-per-GPU layout.

Dialogue turns (batch rows) are independent on the whole video path — attention never mixes samples and
LayerNorm is per row — so the path shards by rows with NO data-path collective.  The only communication is
bookkeeping: the max-over-ranks step time of bench.py and, in tests, gathering the shards back.
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
