"""Beam search over the library-backed model — the drop-in for
/root/reference/model/decode.py:53-104 ``beam_search_decode`` with the same arguments and the same result
``(n-best [(tokens, score)], best_state)``, plus ``beam_search_decode_batch`` for several dialogue turns at once.

What changes underneath, with identical hypotheses as the outcome:
  * the reference calls ``model.decode`` — the WHOLE ``MultimodalDecoder8.forward`` with its video, caption and audio
    blocks — once per hypothesis per step (≈ beam x max_len times per turn, decode.py:63-66) although those blocks do
    not depend on the target; here they run once per turn (``MultimodalDecoder8.encode_memories``) and every step only
    runs the decoder layers and the generator;
  * all live hypotheses of all turns go through the decoder as ONE batch (the reference is batch 1);
  * the candidate selection keeps the reference's procedure (walk each hypothesis' tokens in descending score, fill
    the beam, then replace the current minimum until a candidate no longer improves it) on the top ``beam + 2`` tokens
    per hypothesis taken on the device — no [V]-sized host transfer or host argsort.
"""
from types import SimpleNamespace

import numpy as np
import torch

_MEM_KEYS = ("encoded_query", "encoded_his", "encoded_cap", "encoded_ft", "cap_ft", "temporal_ft", "spatial_ft",
             "audio_ft", "encoded_audio")
_ROW_FIELDS = ("query", "his", "cap", "query_mask", "his_mask", "cap_mask", "audio_mask")


def _rows(t, owner):
    return t.index_select(0, owner)


def _expand(batch, ft, snaps, owner):
    """Per-hypothesis views of the turn-level batch fields, text memories and per-block memories."""
    b = SimpleNamespace()
    for f in _ROW_FIELDS:
        v = getattr(batch, f, None)
        if torch.is_tensor(v):
            setattr(b, f, _rows(v, owner))
    ft_r = {k: _rows(v, owner) for k, v in ft.items() if k in _MEM_KEYS and torch.is_tensor(v)}
    snaps_r = [{k: _rows(v, owner) for k, v in s.items() if k in _MEM_KEYS and torch.is_tensor(v)} for s in snaps]
    return b, ft_r, snaps_r


def _select(hyps, cands, beam, skip):
    """decode.py:77-96 for one turn: `cands[i]` = (tokens, scores) of hypothesis i in descending score order."""
    new, argmin = [], 0
    for (out, _lp, st), (toks, scores) in zip(hyps, cands):
        for o, new_lp in zip(toks, scores):
            if o in skip:
                continue
            if len(new) == beam:
                if new[argmin][1] < new_lp:
                    new[argmin] = (out + [o], new_lp, st + [int(o)])
                    argmin = min(enumerate(new), key=lambda h: h[1][1])[0]
                else:
                    break
            else:
                new.append((out + [o], new_lp, st + [int(o)]))
                if len(new) == beam:
                    argmin = min(enumerate(new), key=lambda h: h[1][1])[0]
    return new


@torch.no_grad()
def beam_search_decode_batch(model, batch, max_len, start_symbol, unk_symbol, end_symbol, pad_symbol, beam=5, penalty=1.0,
                             nbest=5, min_len=1, train_args=None, dec_eos=False):
    """Beam search for every dialogue turn of ``batch``; returns [(n-best, best_state)] per turn."""
    args = train_args if train_args is not None else model.args
    dec = model.mutlimodal_decoder
    ft = model.encode(batch)                                   # decode.py:54 (text encoders + video projection), once
    snaps = dec.encode_memories(batch, ft)                     # every block's video / caption / audio layers, once
    B = ft['encoded_query'].shape[0]
    dev = ft['encoded_query'].device
    hyps = [[([], np.float32(0.), [int(start_symbol)])] for _ in range(B)]
    comp = [[] for _ in range(B)]
    best = [None] * B
    skip = {unk_symbol} if dec_eos else {unk_symbol, end_symbol}
    k = min(beam + 2, model.generator.vocab_gen.shape[0])
    cache_key, cached = None, None
    for l in range(max_len):
        owner_l = [b for b in range(B) for _ in hyps[b]]
        if not owner_l:
            break
        if cache_key != owner_l:                               # the expansion only changes while the beams fill up
            owner = torch.tensor(owner_l, dtype=torch.long, device=dev)
            cached, cache_key = _expand(batch, ft, snaps, owner), owner_l
        b_r, ft_r, snaps_r = cached
        trg = torch.tensor([st for b in range(B) for (_o, _lp, st) in hyps[b]], dtype=torch.long, device=dev)
        H, L = trg.shape
        b_r.trg = trg
        b_r.trg_mask = torch.tril(torch.ones(L, L, dtype=torch.bool, device=dev)).unsqueeze(0).expand(H, L, L)
        x = model.tgt_embed(trg)                                                   # mtn.py:58
        d_txt = dec.decode_with_memories(b_r, snaps_r, x)
        ft_g = dict(ft_r)
        ft_g['decoded_text'] = d_txt[:, -1:].contiguous()                          # decode.py:67-68
        ft_g['encoded_tgt'] = x[:, -1:].contiguous()
        logp = model.generator(ft_g, b_r, args).reshape(H, -1)                     # decode.py:70
        top_v, top_i = logp.topk(k, dim=-1)
        end_v = logp[:, end_symbol]
        top_v, top_i, end_v = top_v.cpu().numpy(), top_i.cpu().numpy(), end_v.cpu().numpy()   # the step's one sync
        row = 0
        for b in range(B):
            cands = []
            for (out, lp, _st) in hyps[b]:
                if l >= min_len:                                                   # decode.py:73-77
                    new_lp = (end_v[row] + lp) + penalty * (len(out) + 1)
                    comp[b].append((out, new_lp))
                    if best[b] is None or best[b] < new_lp:
                        best[b] = new_lp
                cands.append((top_i[row], top_v[row] + lp))
                row += 1
            hyps[b] = _select(hyps[b], cands, beam, skip)
    results = []
    for b in range(B):
        if comp[b]:
            results.append((sorted(comp[b], key=lambda h: -h[1])[:nbest], best[b]))   # decode.py:99-101
        else:
            results.append(([([], 0)], None))
    return results


def beam_search_decode(model, batch, max_len, start_symbol, unk_symbol, end_symbol, pad_symbol, beam=5, penalty=1.0, nbest=5,
                       min_len=1, train_args=None, dec_eos=False):
    "decode.py:53-104 — one dialogue turn (batch of 1), same return value."
    if batch.query.shape[0] != 1:
        raise ValueError("beam_search_decode takes one dialogue turn (the reference's protocol); "
                         "use beam_search_decode_batch for several")
    return beam_search_decode_batch(model, batch, max_len, start_symbol, unk_symbol, end_symbol, pad_symbol, beam, penalty,
                                    nbest, min_len, train_args, dec_eos)[0]
