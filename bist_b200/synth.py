"""Seeded synthetic AVSD-shaped inputs and weights for the BiST video hot path.

Host-side harness code shared by tests/, bench.py, __graft_entry__.py and
oracle/make_golden.py.  Everything is drawn from numpy's legacy ``RandomState`` streams
(frozen by numpy's compatibility policy), so a (seed, shape) pair names the same tensors
in the build container, on the GPU box and in the committed golden fixtures.

Shapes and distributions follow SURVEY.md §8(d):
  * ``fts`` ~ U[0,1) fp32 [B,T,S,D] (ResNeXt post-ReLU features are non-negative), with a
    per-sample valid length T_i ~ U{ceil(T/2)..T} (T_0 = T) and clips t >= T_i zeroed, so
    that ``temporal_mask = fts.sum(2).sum(-1) != 0`` reproduces data/dataset.py:79;
  * token ids ~ U{4..V-1}, pad id 1 on the tail, lengths ~ U{ceil(L/2)..L} (first = L);
  * weights: xavier-uniform matrices (mtn.py:163-165), biases N(0, 0.02), LayerNorm
    gains 1 + N(0, 0.1) and offsets N(0, 0.1) so every affine path is exercised.
"""
import math
from types import SimpleNamespace

import numpy as np
import torch

PAD_ID = 1  # data/dataset.py collate_fn pads with 1


class SynthBatch(SimpleNamespace):
    """Duck-typed stand-in for data/dataset.py:59 ``Batch`` (only the fields the path reads)."""

    def to(self, device, non_blocking=False):
        out = SynthBatch()
        for k, v in self.__dict__.items():
            setattr(out, k, v.to(device, non_blocking=non_blocking) if torch.is_tensor(v) else v)
        return out


def make_batch(seed, B, T, S, D, Lq, vocab=4096, full_clips=False, full_query=False):
    """Seeded batch with the fields of SURVEY.md §8(b): fts, temporal_mask, query, query_mask."""
    rs = np.random.RandomState(seed)
    fts = rs.uniform(0.0, 1.0, size=(B, T, S, D)).astype(np.float32)
    t_len = rs.randint(int(math.ceil(T / 2)), T + 1, size=B)
    t_len[0] = T
    q_len = rs.randint(int(math.ceil(Lq / 2)), Lq + 1, size=B)
    q_len[0] = Lq
    if full_clips:
        t_len[:] = T
    if full_query:
        q_len[:] = Lq
    for i in range(B):
        fts[i, t_len[i]:] = 0.0
    query = rs.randint(4, vocab, size=(B, Lq)).astype(np.int64)
    for i in range(B):
        query[i, q_len[i]:] = PAD_ID
    b = SynthBatch()
    b.fts = torch.from_numpy(fts)
    b.query = torch.from_numpy(query)
    b.query_mask = (b.query != PAD_ID).unsqueeze(-2)                  # dataset.py:66
    b.temporal_mask = (b.fts.sum(2).sum(-1) != 0).unsqueeze(-2)       # dataset.py:79
    b.spatial_mask = (b.fts.sum(1).sum(-1) != 0).unsqueeze(-2)        # dataset.py:78 (unused by the path)
    b.t_len = torch.from_numpy(t_len.astype(np.int64))
    b.q_len = torch.from_numpy(q_len.astype(np.int64))
    return b


def make_encoded_query(seed, B, Lq, d):
    """Stand-in for ft['encoded_query'] (= LayerNorm of scaled embeddings + PE): N(0,1)."""
    rs = np.random.RandomState(seed + 7919)
    return torch.from_numpy(rs.standard_normal(size=(B, Lq, d)).astype(np.float32))


def _xavier(rs, out_f, in_f):
    a = math.sqrt(6.0 / (in_f + out_f))
    return rs.uniform(-a, a, size=(out_f, in_f)).astype(np.float32)


def _bias(rs, n):
    return (0.02 * rs.standard_normal(size=(n,))).astype(np.float32)


def _ln(rs, P, pre, n):
    P[pre + "a_2"] = (1.0 + 0.1 * rs.standard_normal(size=(n,))).astype(np.float32)
    P[pre + "b_2"] = (0.1 * rs.standard_normal(size=(n,))).astype(np.float32)


def _mha(rs, P, pre, d):
    for i in range(4):
        P[f"{pre}linears.{i}.weight"] = _xavier(rs, d, d)
        P[f"{pre}linears.{i}.bias"] = _bias(rs, d)


def _ffn(rs, P, pre, d, d_ff):
    P[pre + "w_1.weight"] = _xavier(rs, d_ff, d)
    P[pre + "w_1.bias"] = _bias(rs, d_ff)
    P[pre + "w_2.weight"] = _xavier(rs, d, d_ff)
    P[pre + "w_2.bias"] = _bias(rs, d)


def layer_param_names(nb_attn, nb_ff):
    """state_dict key suffixes of one VidEncoderLayer4 (SURVEY.md §8(a) parameter list)."""
    names = []
    for a in range(nb_attn):
        for i in range(4):
            names += [f"attn.{a}.linears.{i}.weight", f"attn.{a}.linears.{i}.bias"]
    for f in range(nb_ff):
        names += [f"ff.{f}.w_1.weight", f"ff.{f}.w_1.bias", f"ff.{f}.w_2.weight", f"ff.{f}.w_2.bias"]
    for s in range(nb_attn + nb_ff):
        names += [f"sublayer.{s}.norm.a_2", f"sublayer.{s}.norm.b_2"]
    return names


def make_params(seed, d, D, n_blocks, t2s=True, s2t=True, d_ff=None,
                enc_pre="vid_encoder.", dec_pre="mutlimodal_decoder."):
    """Video-path parameters keyed exactly like the reference ``MTN.state_dict()``.

    Returns dict[str, torch.FloatTensor] (CPU).  With one direction off the layer has
    3 attn / 1 ff / 4 sublayer clones (mtn.py:123-128).
    """
    rs = np.random.RandomState(seed + 104729)
    d_ff = d_ff or 4 * d                                               # mtn.py:70
    nb_attn, nb_ff = (6, 2) if (t2s and s2t) else (3, 1)
    P = {}
    P[enc_pre + "W.weight"] = _xavier(rs, d, D)
    P[enc_pre + "W.bias"] = _bias(rs, d)
    _ln(rs, P, enc_pre + "in_norm.", d)
    for blk in range(n_blocks):
        pre = f"{dec_pre}v_layers.{blk}."
        for a in range(nb_attn):
            _mha(rs, P, f"{pre}attn.{a}.", d)
        for f in range(nb_ff):
            _ffn(rs, P, f"{pre}ff.{f}.", d, d_ff)
        for s in range(nb_attn + nb_ff):
            _ln(rs, P, f"{pre}sublayer.{s}.norm.", d)
    _ln(rs, P, dec_pre + "spatial_out_norm.", d)
    _ln(rs, P, dec_pre + "temporal_out_norm.", d)
    return {k: torch.from_numpy(v) for k, v in P.items()}


def shard_rows(n_rows, world_size, rank):
    """Contiguous, balanced split of ``n_rows`` dialogue turns over ranks.

    Returns (start, stop).  The first ``n_rows % world_size`` ranks get one extra row.
    """
    base, rem = divmod(n_rows, world_size)
    start = rank * base + min(rank, rem)
    return start, start + base + (1 if rank < rem else 0)


# ---------------------------------------------------------------------------------------------- decoder side (§8(f) row 1)
def fill_params(shapes, seed):
    """Seeded values for a {name: shape} table (a ``state_dict`` of the decoder side), drawn in sorted-name order
    so that the reference model in oracle/make_golden.py and the mirror modules in tests/ get the same weights:
    matrices xavier-uniform, LayerNorm gains 1 + N(0, 0.1), offsets N(0, 0.1), other vectors N(0, 0.02)."""
    rs = np.random.RandomState(seed)
    out = {}
    for name in sorted(shapes):
        shp = tuple(shapes[name])
        if len(shp) > 1:
            a = math.sqrt(6.0 / (shp[0] + shp[1]))
            v = rs.uniform(-a, a, size=shp)
        elif name.endswith("a_2"):
            v = 1.0 + 0.1 * rs.standard_normal(shp)
        elif name.endswith("b_2"):
            v = 0.1 * rs.standard_normal(shp)
        else:
            v = 0.02 * rs.standard_normal(shp)
        out[name] = torch.from_numpy(v.astype(np.float32))
    return out


def _pad_mask(rs, B, L):
    n = rs.randint(int(math.ceil(L / 2)), L + 1, size=B)
    n[0] = L
    return torch.from_numpy(np.arange(L)[None, :] < n[:, None]).unsqueeze(1)       # [B,1,L] bool


def make_decoder_inputs(seed, B, Lq, Lh, Lc, Lt, Ta, T, S, d):
    """Seeded inputs of MultimodalDecoder8.forward (decoder.py:107): a batch with the masks of data/dataset.py:59-99
    (trg_mask = pad & causal, make_std_mask :101-104) and the ft dict the text / video encoders would have filled
    (N(0,1) activations).  Returns (batch, ft, encoded_tgt)."""
    rs = np.random.RandomState(seed)
    b = SynthBatch()
    b.query_mask, b.his_mask, b.cap_mask = _pad_mask(rs, B, Lq), _pad_mask(rs, B, Lh), _pad_mask(rs, B, Lc)
    b.audio_mask, b.temporal_mask = _pad_mask(rs, B, Ta), _pad_mask(rs, B, T)
    b.trg_mask = _pad_mask(rs, B, Lt) & torch.tril(torch.ones(Lt, Lt, dtype=torch.bool)).unsqueeze(0)   # [B,Lt,Lt]

    def act(*shape):
        return torch.from_numpy(rs.standard_normal(shape).astype(np.float32))
    ft = {"encoded_query": act(B, Lq, d), "encoded_his": act(B, Lh, d), "encoded_cap": act(B, Lc, d),
          "encoded_audio": act(B, Ta, d), "spatiotemporal_ft": act(B, T, S, d)}
    return b, ft, act(B, Lt, d)


def _tokens(rs, B, L, vocab, unk_rate=0.0):
    ids = rs.randint(4, vocab, size=(B, L)).astype(np.int64)
    if unk_rate > 0:
        ids[rs.uniform(size=(B, L)) < unk_rate] = 0                   # <unk>: masked out of the pointers (mask_unk)
    n = rs.randint(int(math.ceil(L / 2)), L + 1, size=B)
    n[0] = L
    ids[np.arange(L)[None, :] >= n[:, None]] = PAD_ID
    return torch.from_numpy(ids)


def make_full_batch(seed, B, T, S, D, Lq, Lh, Lc, Lt, vocab, Ta=0, Da=0):
    """Seeded batch with every field MTN.forward and SimpleLossCompute read (data/dataset.py:59-104): token ids of the
    query / history / caption / shifted target (pad id 1 on the tails, a few <unk>=0 in the sources), their masks,
    trg_mask = pad & causal, ntokens / qntokens, and the video features of make_batch.  With ``Ta, Da > 0`` the batch
    also carries the audio stream of data/dataset.py:83-87: ``audio_fts`` [B,Ta,Da] (VGGish-like, non-negative, zero
    rows past each sample's length) and ``audio_mask`` = (audio_fts.sum(-1) != 0) [B,1,Ta]."""
    vb = make_batch(seed, B, T, S, D, Lq)
    rs = np.random.RandomState(seed + 7919)
    b = SynthBatch()
    b.fts, b.temporal_mask = vb.fts, vb.temporal_mask
    b.query, b.his, b.cap = _tokens(rs, B, Lq, vocab, 0.1), _tokens(rs, B, Lh, vocab, 0.05), _tokens(rs, B, Lc, vocab, 0.05)
    tgt = _tokens(rs, B, Lt + 1, vocab)
    b.trg, b.trg_y = tgt[:, :-1].contiguous(), tgt[:, 1:].contiguous()
    b.query_mask, b.his_mask, b.cap_mask = ((t != PAD_ID).unsqueeze(-2) for t in (b.query, b.his, b.cap))
    b.trg_mask = (b.trg != PAD_ID).unsqueeze(-2) & torch.tril(torch.ones(Lt, Lt, dtype=torch.bool)).unsqueeze(0)
    b.ntokens, b.qntokens = (b.trg_y != PAD_ID).sum(), (b.query != PAD_ID).sum()
    if Ta > 0 and Da > 0:
        ra = np.random.RandomState(seed + 104729)
        aud = ra.uniform(0.05, 1.0, size=(B, Ta, Da)).astype(np.float32)
        n = ra.randint(int(math.ceil(Ta / 2)), Ta + 1, size=B)
        n[0] = Ta
        aud[np.arange(Ta)[None, :] >= n[:, None]] = 0.0
        b.audio_fts = torch.from_numpy(aud)
        b.audio_mask = (b.audio_fts.sum(-1) != 0).unsqueeze(-2)
    return b


def batch_rows(b, lo, hi):
    """Rows [lo, hi) of a batch as a new batch (dialogue turns are independent)."""
    n = b.query.shape[0] if hasattr(b, "query") else b.fts.shape[0]
    out = SynthBatch()
    for k, v in b.__dict__.items():
        setattr(out, k, v[lo:hi] if torch.is_tensor(v) and v.dim() > 0 and v.shape[0] == n else v)
    return out
