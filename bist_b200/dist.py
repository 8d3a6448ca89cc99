"""Host-side helpers for the one-process-per-GPU layout.

Dialogue turns (batch rows) are independent on the whole video path — attention never mixes samples and
LayerNorm is per row — so the path shards by rows with NO data-path collective.  The only communication is
bookkeeping (the max-over-ranks step time of bench.py and, in tests, gathering the shards back) and, for the
training step, ONE gradient all-reduce per step (`allreduce_grads`).
"""
import os

import torch
import torch.distributed as dist

from . import arena as _arena
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


def sync_token_counts(batch):
    """Loss normalisers of a row-sharded batch (reference normaliser /root/reference/model/optimize.py:46-50).

    The reference divides the summed loss by the token counts of the WHOLE batch.  With the rows sharded over ranks and the
    gradients AVERAGED by the all-reduce, every rank must divide its local sums by (global count / world size) for the
    averaged gradient to equal the single-GPU gradient on the concatenated batch.  This all-reduces the two counts
    (``ntokens``, ``qntokens``) and keeps ``count / world`` in ``batch.loss_norm`` (a STATIC fp32 [2] device tensor, updated in
    place so that a captured training step sees the new values); ``SimpleLossCompute.loss`` uses it when present.
    Without a process group it holds the local counts (world = 1)."""
    t = torch.stack([batch.ntokens.float(), batch.qntokens.float()])
    world = 1
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        world = dist.get_world_size()
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
    norm = getattr(batch, "loss_norm", None)
    if norm is None or norm.device != t.device:
        batch.loss_norm = norm = torch.empty(2, dtype=torch.float32, device=t.device)
    norm.copy_(t / world)
    return norm


_pending = None          # [(work, storage_ptr, block, need_div, wire)] while overlap_grad_allreduce() is active, else None
_comm_dtype = torch.float32
_wire_cache = {}


def set_grad_comm_dtype(dtype):
    """Element type the gradients cross NVLink in.  ``torch.float32`` (default): the fp32 gradient blocks are reduced in place.
    ``torch.bfloat16``: each block is cast into a persistent bf16 wire buffer, that buffer is all-reduced (half the bytes) and
    the result is written back over the fp32 block — the accumulation inside every rank stays fp32, only the exchanged
    values are rounded (what bf16-mode training does with every activation anyway).  bench.py selects bf16 for bf16 mode."""
    global _comm_dtype
    assert dtype in (torch.float32, torch.bfloat16)
    _comm_dtype = dtype


def _wire(numel, device, slot, dtype=torch.bfloat16):
    "Persistent wire buffer (one per slot = per call site / block, so that captured graphs keep valid addresses)."
    key = (slot, numel, str(device), dtype)
    buf = _wire_cache.get(key)
    if buf is None:
        buf = _wire_cache[key] = torch.empty(numel, dtype=dtype, device=device)
    return buf


# Buckets: the decoder-side modules (attention, feed-forward) hand in their gradient blocks as their backward is enqueued; blocks
# of the SAME stream (autograd runs a node's backward on its forward's stream, and the model's branches live on different streams)
# are packed into one wire buffer and all-reduced as soon as BIST_GRAD_BUCKET_MB of them are there, so the transfer runs beside the
# rest of the backward instead of after it.  0 (the default) disables bucketing: everything left goes in the final launch.
# Measured on 8 B200 (config 3 with audio, bf16 wire): 24 MB buckets 11.86 ms, 48 MB 11.90 ms, none 11.94 ms per step, while the
# end-to-end step (host batches prefetched over PCIe during the backward) went from 14.9 to 16.2 ms — so it stays opt-in.
_BUCKET_BYTES = int(float(os.environ.get("BIST_GRAD_BUCKET_MB", "0")) * (1 << 20))
_buckets = {}


def _flush_bucket(sid):
    blocks = _buckets.pop(sid, None)
    if not blocks:
        return
    op = _reduce_op()
    flat = [b.reshape(-1) for b in blocks]
    sizes = [b.numel() for b in flat]
    wire = _wire(sum(sizes), flat[0].device, ("bucket", len(_pending)), _comm_dtype)
    views = list(torch.split(wire, sizes))
    torch._foreach_copy_(views, flat)
    work = dist.all_reduce(wire, op=op, async_op=True)
    rng = [(b.data_ptr(), b.data_ptr() + b.numel() * b.element_size()) for b in flat]
    _pending.append((work, rng, flat, op == dist.ReduceOp.SUM, views))


class overlap_grad_allreduce:
    """``with overlap_grad_allreduce(): loss.backward()`` — every VidEncoderLayer4 starts the all-reduce of its
    gradient block as soon as its backward kernels are enqueued (NCCL runs it on its own stream, behind the
    kernels that write the block and beside the backward of the earlier layers); ``allreduce_grads`` afterwards
    waits for those and reduces only what is left.  A no-op without a process group or with world_size 1."""

    def __enter__(self):
        global _pending
        if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
            _pending = []
        return self

    def __exit__(self, exc_type, *exc):
        global _pending
        if exc_type is not None and _pending is not None:   # backward raised: finish what was started, disarm
            for work, *_ in _pending:
                work.wait()
            _pending = None
            _buckets.clear()
        return False


def _reduce_op():
    # NCCL averages inside the collective; gloo has no AVG, so sum and divide afterwards.
    return dist.ReduceOp.AVG if dist.get_backend() == "nccl" else dist.ReduceOp.SUM


def grad_block_ready(block, average=True, bucket=False):
    """Called by a backward (encoder.py, modules.py) with its freshly written gradient block.  ``bucket=True``: a small block
    that is collected with its stream's other blocks (see above); else the block is all-reduced on its own right away."""
    if _pending is None:
        return
    if bucket:
        if _BUCKET_BYTES <= 0 or not average or block.dtype != torch.float32 or dist.get_backend() != "nccl":
            return
        sid = torch.cuda.current_stream(block.device).cuda_stream
        lst = _buckets.setdefault(sid, [])
        lst.append(block)
        if sum(b.numel() for b in lst) * 4 >= _BUCKET_BYTES:
            _flush_bucket(sid)
        return
    op = _reduce_op() if average else dist.ReduceOp.SUM
    wire = None
    if _comm_dtype == torch.bfloat16 and block.dtype == torch.float32 and dist.get_backend() == "nccl":
        wire = _wire(block.numel(), block.device, ("block", len(_pending)))   # blocks arrive in the same order every step
        wire.copy_(block.reshape(-1))
    work = dist.all_reduce(block if wire is None else wire, op=op, async_op=True)
    _pending.append((work, [(block.data_ptr(), block.data_ptr() + block.numel() * block.element_size())], [block.reshape(-1)],
                     average and op == dist.ReduceOp.SUM, None if wire is None else [wire]))


def allreduce_grads(params, average=True):
    """Sum (or average) the gradients of `params` over all ranks with as few collectives as possible.

    The backward of a VidEncoderLayer4 writes all of its parameter gradients into one flat fp32 block
    (include/bist_b200.h, "Gradient block") and returns slices of it; autograd adopts those slices as the
    parameters' ``.grad`` without copying (they keep sharing the block's storage, though they are no longer
    autograd views), so every layer is ONE all-reduce over that storage; stand-alone gradients are flattened
    into one extra bucket.  Blocks already started by ``overlap_grad_allreduce`` are only waited for.  Returns
    the number of collectives issued (started ones included).  A no-op (returns 0) without an initialised
    process group or with world_size 1.
    """
    global _pending
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return 0
    world = dist.get_world_size()
    op = _reduce_op() if average else dist.ReduceOp.SUM
    div = average and op == dist.ReduceOp.SUM
    started = []             # address ranges already reduced by overlap_grad_allreduce
    n = 0
    if _pending is not None:
        for sid in list(_buckets):       # what the buckets still hold (the backward is complete: every stream has joined this one)
            _flush_bucket(sid)
    for work, rng, flat, need_div, views in (_pending or []):
        work.wait()
        if views is not None:
            torch._foreach_copy_(flat, views)
        if need_div:
            torch._foreach_div_(flat, world)
        started.extend(rng)
        n += 1
    _pending = None
    _buckets.clear()
    blocks, loose, runs = {}, [], []
    for p in params:
        g = p.grad
        if g is None:
            continue
        st = g.untyped_storage()
        if any(lo <= g.data_ptr() < hi for lo, hi in started):
            continue
        if g.dtype == torch.float32 and g.is_contiguous() and _arena.owns(g):
            runs.append(g)       # carved from the step's gradient arena: merged into contiguous runs below
            continue
        if g.dtype == torch.float32 and g.is_contiguous() and st.nbytes() > g.numel() * 4:
            blocks.setdefault(st.data_ptr(), [st, 0, []])
            blocks[st.data_ptr()][1] += g.numel() * 4
            blocks[st.data_ptr()][2].append(g)
        else:
            loose.append(g)
    flats = []
    if runs:
        # neighbours in the arena (gaps are alignment padding, which is zero on every rank) become ONE flat tensor per run
        runs.sort(key=lambda g: g.data_ptr())
        base = runs[0].untyped_storage()
        lo = hi = None
        for g in runs:
            a, b = g.data_ptr(), g.data_ptr() + g.numel() * 4
            if hi is not None and a - hi <= 1024 and not any(x < a and y > hi for x, y in started):
                hi = max(hi, b)
                continue
            if hi is not None:
                flats.append(torch.empty(0, dtype=torch.float32, device=g.device).set_(base, (lo - base.data_ptr()) // 4, ((hi - lo) // 4,)))
            lo, hi = a, b
        flats.append(torch.empty(0, dtype=torch.float32, device=runs[0].device).set_(base, (lo - base.data_ptr()) // 4, ((hi - lo) // 4,)))
    for st, covered, gs in blocks.values():
        if covered == st.nbytes():           # the storage holds nothing but these gradients: reduce it in place
            flats.append(torch.empty(0, dtype=torch.float32, device=gs[0].device).set_(st, 0, (st.nbytes() // 4,)))
        else:                                # partially used storage: treat its gradients one by one
            loose.extend(gs)
    if dist.get_backend() == "nccl":
        # every block and every stand-alone gradient in place, as ONE grouped NCCL launch (ncclGroupStart/End)
        tensors = flats + [g for g in loose if g.is_contiguous()]
        odd = [g for g in loose if not g.is_contiguous()]
        if tensors and _comm_dtype == torch.bfloat16 and all(t.dtype == torch.float32 for t in tensors):
            # one bf16 wire buffer for everything that is left: cast in (one multi-tensor launch), ONE all-reduce, cast back
            sizes = [t.numel() for t in tensors]
            wire = _wire(sum(sizes), tensors[0].device, ("rest", len(tensors)))
            views = [v.view_as(t) for v, t in zip(torch.split(wire, sizes), tensors)]
            torch._foreach_copy_(views, tensors)
            dist.all_reduce(wire, op=op)
            torch._foreach_copy_(tensors, views)
            n += 1
        elif tensors:
            with dist._coalescing_manager(device=tensors[0].device, async_ops=False):
                for t in tensors:
                    dist.all_reduce(t, op=op)
            n += 1
        for g in odd:
            c = g.contiguous()
            dist.all_reduce(c, op=op)
            g.copy_(c)
            n += 1
        return n
    for flat in flats:
        dist.all_reduce(flat, op=op)
        if div:
            flat.div_(world)
        n += 1
    if len(loose) == 1:
        dist.all_reduce(loose[0], op=op)
        if div:
            loose[0].div_(world)
        n += 1
    elif loose:
        flat = torch.cat([g.reshape(-1).float() for g in loose])
        dist.all_reduce(flat, op=op)
        if div:
            flat.div_(world)
        off = 0
        for g in loose:
            g.copy_(flat[off:off + g.numel()].view_as(g))
            off += g.numel()
        n += 1
    return n
