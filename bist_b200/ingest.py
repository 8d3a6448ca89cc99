"""Feature ingest: what the reference does inside ``collate_fn`` for every batch — load the
per-video ``.npy`` feature arrays, zero-pad them to the longest video of the batch, build the batch tensor and the
temporal mask, copy fp32 to the device (data/dataset.py:146-174, 76-79) — as a cache that is filled once and a collate
that only moves bytes:

  * ``FeatureCache`` keeps every video's ``[T_i, S, D]`` features in ONE page-locked host arena, in bf16 (the compute
    precision of the bf16 mode, rounded to nearest-even exactly as the projection kernel would round the fp32 value)
    or in fp32;
  * ``collate`` writes each video's rows straight from the arena into the zero-padded device batch ``[B, T, S, D]``
    with asynchronous copies (no padded host copy is ever built) and derives ``temporal_mask`` from the clip counts
    (``fts.sum(2).sum(-1) != 0`` of dataset.py:79 for features that are not identically zero on a real clip);
  * with bf16 storage a batch is half the bytes on PCIe and in HBM, and ``VidEncoder8`` reads it with the 2-byte
    projection (``bist_vidproj_fwd_bf16in``).

Layout note: the extractor writes channel-first ``[n_clips, 2048, 4, 4]`` while the loader assumes
channel-last ``[T, H, W, C]``; ``add`` takes either and stores ``[T, S, C]`` (``channels_first=True`` transposes once, at
fill time).
"""
import numpy as np
import torch


class FeatureCache:
    def __init__(self, feature_dim, regions, capacity_clips, dtype=torch.bfloat16, pin=None):
        if dtype not in (torch.bfloat16, torch.float32):
            raise ValueError("FeatureCache stores bf16 or fp32")
        self.D, self.S, self.dtype = int(feature_dim), int(regions), dtype
        pin = torch.cuda.is_available() if pin is None else pin
        self.arena = torch.empty((int(capacity_clips), self.S, self.D), dtype=dtype, pin_memory=bool(pin))
        self.used = 0
        self.index = {}                       # video id -> (first clip, clip count)

    def add(self, vid, feats, channels_first=False):
        """feats: numpy / torch array [T, S, D], [T, H, W, D] or (channels_first) [T, D, H, W]; any float dtype."""
        x = torch.as_tensor(np.asarray(feats) if not torch.is_tensor(feats) else feats)
        if channels_first:
            x = x.reshape(x.shape[0], x.shape[1], -1).transpose(1, 2)          # [T, D, H*W] -> [T, S, D]
        x = x.reshape(x.shape[0], -1, x.shape[-1])                              # dataset.py:151
        if x.shape[1] != self.S or x.shape[2] != self.D:
            raise ValueError("features of %r are %s, cache holds [T, %d, %d]" % (vid, tuple(x.shape), self.S, self.D))
        T = x.shape[0]
        if self.used + T > self.arena.shape[0]:
            raise MemoryError("FeatureCache is full (%d clips)" % self.arena.shape[0])
        self.arena[self.used:self.used + T].copy_(x)                            # the one conversion (RNE to bf16)
        self.index[vid] = (self.used, T)
        self.used += T
        return T

    def __contains__(self, vid):
        return vid in self.index

    def clips(self, vid):
        return self.index[vid][1]

    def host_bytes(self, vids):
        e = 2 if self.dtype == torch.bfloat16 else 4
        return sum(self.index[v][1] for v in vids) * self.S * self.D * e

    def collate(self, vids, device, out=None, T=None, stream=None):
        """Zero-padded batch for the videos ``vids`` on ``device``: (fts [B, T, S, D] in the cache dtype,
        temporal_mask [B, 1, T] bool).  ``out``: a (fts, temporal_mask) pair of preallocated device tensors to fill
        (fixed shapes, e.g. the static inputs of a captured graph); ``T`` defaults to the longest video of the batch."""
        lens = [self.index[v][1] for v in vids]
        B = len(vids)
        T = max(lens) if T is None else int(T)
        if max(lens) > T:
            raise ValueError("a video of %d clips does not fit T=%d" % (max(lens), T))
        if out is None:
            fts = torch.empty((B, T, self.S, self.D), dtype=self.dtype, device=device)
            mask = torch.empty((B, 1, T), dtype=torch.bool, device=device)
        else:
            fts, mask = out
        ctx = torch.cuda.stream(stream) if stream is not None else _null()
        with ctx:
            if fts.is_cuda and self.arena.is_pinned() and fts.is_contiguous():
                self._gather(vids, fts, T)                                     # B copies + fills from one C call
            else:
                for b, v in enumerate(vids):
                    off, n = self.index[v]
                    fts[b, :n].copy_(self.arena[off:off + n], non_blocking=True)
                    if n < T:
                        fts[b, n:].zero_()
            host_mask, done = self._mask_stage(B, T)                                            # pinned ring, guarded by events
            torch.lt(torch.arange(T).unsqueeze(0), torch.tensor(lens).unsqueeze(1), out=host_mask)
            mask.copy_(host_mask.unsqueeze(1), non_blocking=True)
            if done is not None:
                done.record(torch.cuda.current_stream(mask.device))    # the buffer may be rewritten once this copy has run
        return fts, mask

    def _gather(self, vids, fts, T):
        import ctypes
        from . import _lib
        B = len(vids)
        clip = self.S * self.D * self.arena.element_size()
        src = (ctypes.c_void_p * B)()
        dst, nb, poff, pnb = ((ctypes.c_int64 * B)() for _ in range(4))
        base = self.arena.data_ptr()
        for b, v in enumerate(vids):
            off, n = self.index[v]
            src[b] = base + off * clip
            dst[b], nb[b] = b * T * clip, n * clip
            poff[b], pnb[b] = (b * T + n) * clip, (T - n) * clip
        _lib.check(_lib.lib().bist_ingest_gather(src, _lib.ptr(fts), dst, nb, poff, pnb, B, _lib.stream_ptr()),
                   "bist_ingest_gather")

    def _mask_stage(self, B, T):
        """Next pinned staging buffer for a [B, T] mask and the event that guards it.  The H2D copy of a mask is queued behind
        the gather kernel and never blocks the host, so a buffer is only reused after the event recorded behind ITS copy
        has completed (waited for here, on the host) — otherwise batch i could run with batch i+1's mask."""
        ring = self.__dict__.setdefault("_stage", {}).setdefault((B, T), {"bufs": [], "events": [], "next": 0})
        pinned = self.arena.is_pinned()
        if len(ring["bufs"]) < 4:
            ring["bufs"].append(torch.empty((B, T), dtype=torch.bool, pin_memory=pinned))
            ring["events"].append(torch.cuda.Event() if pinned and torch.cuda.is_available() else None)
            i = len(ring["bufs"]) - 1
        else:
            i = ring["next"]
            if ring["events"][i] is not None:
                ring["events"][i].synchronize()
        ring["next"] = (i + 1) % 4
        return ring["bufs"][i], ring["events"][i]


class _null:
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False


def reference_collate(arrays):
    """What data/dataset.py:164-174,79 builds on the host for the same videos (fp32 zero-padded batch + mask); the
    check that ``FeatureCache.collate`` is measured against in tests/."""
    T = max(a.shape[0] for a in arrays)
    S, D = arrays[0].shape[1], arrays[0].shape[2]
    x = np.zeros((len(arrays), T, S, D), dtype=np.float32)
    for b, a in enumerate(arrays):
        x[b, :a.shape[0]] = a
    fts = torch.from_numpy(x)
    return fts, (fts.sum(2).sum(-1) != 0).unsqueeze(-2)
