"""CPU oracle for the BiST video hot path.  TEST INFRASTRUCTURE ONLY.

This file is a functional (no nn.Module, no state) restatement, on plain torch CPU
tensors, of the reference algorithm for the path named in BASELINE.json:

    VidEncoder8.forward            /root/reference/model/encoder.py:72-93
    VidEncoderLayer4.forward       /root/reference/model/encoder.py:172-201
      .temporal2spatial            /root/reference/model/encoder.py:109-139
      .spatial2temporal            /root/reference/model/encoder.py:141-170
    LayerNorm / SublayerConnection / attention / MultiHeadedAttention /
    PositionwiseFeedForward        /root/reference/model/modules.py:20-113
    output norms                   /root/reference/model/decoder.py:107-129

It deliberately keeps the reference's *formulation* (query block copied S or T times,
K/V projections of every video row, materialised score tensors) so that (a) it is an
independent check of the re-associated CUDA path and (b) its CPU timing is
representative of the reference's own CPU path (bench.py `cpu_baseline`, kind="port").

Who may use it: only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
`--impl reference` legs.  The product package `bist_b200` never imports it.

How it is pinned (the reference ships no tests or golden vectors):
  * tests/golden/*.npz were produced by running the UNMODIFIED reference modules
    (imported from /root/reference in the build container) on seeded inputs; the
    generating script is oracle/make_golden.py; tests/test_oracle.py checks this
    oracle against every fixture (fp32, <=2e-6 normalised error).
  * when /root/reference is present (build container only) tests/test_oracle.py also
    runs the live reference beside the oracle on fresh random shapes.

All parameters are passed as a flat dict using the reference's state_dict key names,
so a reference `state_dict()` can be fed in unchanged.
"""
import math

import torch

__all__ = [
    "layer_norm", "linear", "sdp_attention", "mha", "ffn", "sublayer",
    "vid_encoder8", "temporal2spatial", "spatial2temporal", "vid_encoder_layer4",
    "video_path", "norm_err",
]


def norm_err(a, b):
    """Parity metric of SURVEY.md §8(c): max|a-b| / max|b| (b = reference)."""
    a = a.detach().double().cpu()
    b = b.detach().double().cpu()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


# ----------------------------------------------------------------------------- blocks
def layer_norm(x, a_2, b_2, eps=1e-6):
    """modules.py:28-31 — unbiased std (divide by d-1); eps is added to the std."""
    mu = x.mean(dim=-1, keepdim=True)
    sd = x.std(dim=-1, keepdim=True)          # torch default: Bessel-corrected
    return a_2 * (x - mu) / (sd + eps) + b_2


def linear(x, w, b=None):
    """nn.Linear: x @ w.T + b."""
    y = x.matmul(w.t())
    return y if b is None else y + b


def _apply(drop, site, x):
    """Dropout site hook: `drop(site, x) -> x'` (training-mode tests inject explicit masks); None = eval mode."""
    return x if drop is None else drop(site, x)


def sdp_attention(q, k, v, mask=None, drop=None, site=None):
    """modules.py:54-64.  mask==0 -> score := -1e9; dropout on the probabilities (:62-63)."""
    s = q.matmul(k.transpose(-2, -1)) / math.sqrt(q.shape[-1])
    if mask is not None:
        s = s.masked_fill(mask == 0, -1e9)
    p = _apply(drop, site, torch.softmax(s, dim=-1))
    return p.matmul(v), p


def mha(P, pre, h, query, key, value, mask=None, drop=None):
    """modules.py:81-100.  P[pre+'linears.{0..3}.{weight,bias}'].  Returns (out, attn)."""
    if mask is not None:
        mask = mask.unsqueeze(1)                                   # :85 same mask for all heads
    nb = query.shape[0]
    d_model = P[pre + "linears.0.weight"].shape[0]
    dk = d_model // h

    def split(i, x):
        y = linear(x, P[pre + f"linears.{i}.weight"], P[pre + f"linears.{i}.bias"])
        return y.view(nb, -1, h, dk).transpose(1, 2)               # :89-91

    q, k, v = split(0, query), split(1, key), split(2, value)
    ctx, p = sdp_attention(q, k, v, mask, drop, pre + "p")         # :94
    ctx = ctx.transpose(1, 2).contiguous().view(nb, -1, h * dk)    # :98-99
    return linear(ctx, P[pre + "linears.3.weight"], P[pre + "linears.3.bias"]), p


def ffn(P, pre, x, drop=None):
    """modules.py:112-113: w_2(dropout(relu(w_1 x)))."""
    hid = _apply(drop, pre + "hidden", torch.relu(linear(x, P[pre + "w_1.weight"], P[pre + "w_1.bias"])))
    return linear(hid, P[pre + "w_2.weight"], P[pre + "w_2.bias"])


def sublayer(P, pre, x, fn, drop=None):
    """modules.py:42-44: x + dropout(sublayer(norm(x))); only x is normed."""
    return x + _apply(drop, pre + "out", fn(layer_norm(x, P[pre + "norm.a_2"], P[pre + "norm.b_2"])))


# ----------------------------------------------------------------------------- a1
def vid_encoder8(P, pre, fts, audio_fts=None, no_w_audio=False):
    """encoder.py:72-93.  Returns (spatiotemporal_ft, encoded_audio-or-None).

    vid_position is always None in the reference (mtn.py:108), so :76-79 never runs.
    """
    out = torch.relu(linear(fts, P[pre + "W.weight"], P[pre + "W.bias"]))          # :75
    out = layer_norm(out, P[pre + "in_norm.a_2"], P[pre + "in_norm.b_2"])          # :80
    aud = None
    if audio_fts is not None:
        aud = audio_fts if no_w_audio else torch.relu(
            linear(audio_fts, P[pre + "a_W.weight"], P[pre + "a_W.bias"]))         # :84-87
        aud = layer_norm(aud, P[pre + "a_in_norm.a_2"], P[pre + "a_in_norm.b_2"])  # :90
    return out, aud


# ----------------------------------------------------------------------------- a3 / a4
def temporal2spatial(P, pre, h, x, vft, temporal_mask, ai, fi, si, drop=None):
    """encoder.py:109-139.  x [B,Lq,d]; vft [B,T,S,d]; temporal_mask [B,1,T] bool.

    ai/fi/si = indices of the first attn / ff / sublayer clone used (reference keeps
    them as mutable cursors, encoder.py:122-137).
    """
    B, T, S, d = vft.shape
    Lq = x.shape[1]
    kv = vft.permute(0, 2, 1, 3).reshape(B * S, T, d)                              # :111-112
    xe = x.unsqueeze(1).expand(B, S, Lq, d).reshape(B * S, Lq, d)                  # :114-115
    me = temporal_mask.unsqueeze(1).expand(B, S, 1, T).reshape(B * S, 1, T)        # :117-119
    t_out = sublayer(P, f"{pre}sublayer.{si}.", xe,
                     lambda z: mha(P, f"{pre}attn.{ai}.", h, z, kv, kv, me, drop)[0], drop)    # :121
    kv2 = t_out.reshape(B, S, Lq, d).permute(0, 2, 1, 3).reshape(B * Lq, S, d)     # :125-126
    x1 = x.reshape(B * Lq, 1, d)                                                   # :128
    ts = sublayer(P, f"{pre}sublayer.{si + 1}.", x1,
                  lambda z: mha(P, f"{pre}attn.{ai + 1}.", h, z, kv2, kv2, None, drop)[0], drop)  # :130
    ts = ts.reshape(B, Lq, d)                                                      # :134
    return sublayer(P, f"{pre}sublayer.{si + 2}.", ts,
                    lambda z: ffn(P, f"{pre}ff.{fi}.", z, drop), drop)             # :135


def spatial2temporal(P, pre, h, x, vft, temporal_mask, ai, fi, si, drop=None):
    """encoder.py:141-170 (mirror order: over S per clip without mask, then over T with mask)."""
    B, T, S, d = vft.shape
    Lq = x.shape[1]
    kv = vft.reshape(B * T, S, d)                                                  # :143
    xe = x.unsqueeze(1).expand(B, T, Lq, d).reshape(B * T, Lq, d)                  # :145-146
    s_out = sublayer(P, f"{pre}sublayer.{si}.", xe,
                     lambda z: mha(P, f"{pre}attn.{ai}.", h, z, kv, kv, None, drop)[0], drop)  # :148
    kv2 = s_out.reshape(B, T, Lq, d).permute(0, 2, 1, 3).reshape(B * Lq, T, d)     # :152-153
    x1 = x.reshape(B * Lq, 1, d)                                                   # :155
    me = temporal_mask.unsqueeze(1).expand(B, Lq, 1, T).reshape(B * Lq, 1, T)      # :157-159
    st = sublayer(P, f"{pre}sublayer.{si + 1}.", x1,
                  lambda z: mha(P, f"{pre}attn.{ai + 1}.", h, z, kv2, kv2, me, drop)[0], drop)  # :161
    st = st.reshape(B, Lq, d)                                                      # :165
    return sublayer(P, f"{pre}sublayer.{si + 2}.", st,
                    lambda z: ffn(P, f"{pre}ff.{fi}.", z, drop), drop)             # :166


# ----------------------------------------------------------------------------- a2
def vid_encoder_layer4(P, pre, h, in_t2s, in_s2t, vft, query_mask, temporal_mask,
                       t2s=True, s2t=True, drop=None):
    """encoder.py:172-201 with enc_st_combine == 'none'.  Returns (t2s_out, s2t_out).

    Clone indexing follows the reference cursors: with both directions on, t2s uses
    attn 0-2 / ff 0 / sublayer 0-3 and s2t uses attn 3-5 / ff 1 / sublayer 4-7; with one
    direction off the surviving one uses attn 0-2 / ff 0 / sublayer 0-3.
    """
    ai = fi = si = 0
    o_t2s, o_s2t = in_t2s, in_s2t
    if t2s:
        x = sublayer(P, f"{pre}sublayer.{si}.", in_t2s,
                     lambda z: mha(P, f"{pre}attn.{ai}.", h, z, z, z, query_mask, drop)[0], drop)  # :176
        o_t2s = temporal2spatial(P, pre, h, x, vft, temporal_mask, ai + 1, fi, si + 1, drop)  # :179
        ai, fi, si = ai + 3, fi + 1, si + 4
    if s2t:
        a0, f0, s0 = ai, fi, si
        x = sublayer(P, f"{pre}sublayer.{s0}.", in_s2t,
                     lambda z: mha(P, f"{pre}attn.{a0}.", h, z, z, z, query_mask, drop)[0], drop)  # :184
        o_s2t = spatial2temporal(P, pre, h, x, vft, temporal_mask, a0 + 1, f0, s0 + 1, drop)  # :187
    return o_t2s, o_s2t


# ----------------------------------------------------------------------------- a1 + L x a2 + a10
def video_path(P, h, fts, encoded_query, query_mask, temporal_mask, n_blocks,
               t2s=True, s2t=True,
               enc_pre="vid_encoder.", dec_pre="mutlimodal_decoder.", drop=None):
    """The slice of MTN.forward that is the hot path (mtn.py:49-51, decoder.py:107-129).

    Returns dict(spatiotemporal_ft, t2s, s2t, temporal_ft, spatial_ft) where t2s/s2t are
    the raw streams after the last block and temporal_ft/spatial_ft their output norms
    (decoder.py:126-129; overwritten every block, so only the last block's survive).
    """
    vft, _ = vid_encoder8(P, enc_pre, fts)
    cur_t2s = cur_s2t = encoded_query                                              # decoder.py:109-110
    out = {"spatiotemporal_ft": vft}
    for blk in range(n_blocks):
        cur_t2s, cur_s2t = vid_encoder_layer4(
            P, f"{dec_pre}v_layers.{blk}.", h, cur_t2s, cur_s2t, vft,
            query_mask, temporal_mask, t2s=t2s, s2t=s2t, drop=drop)                # decoder.py:116
    out["t2s"], out["s2t"] = cur_t2s, cur_s2t
    if s2t:
        out["temporal_ft"] = layer_norm(cur_s2t, P[dec_pre + "temporal_out_norm.a_2"],
                                        P[dec_pre + "temporal_out_norm.b_2"])      # decoder.py:127
    if t2s:
        out["spatial_ft"] = layer_norm(cur_t2s, P[dec_pre + "spatial_out_norm.a_2"],
                                       P[dec_pre + "spatial_out_norm.b_2"])        # decoder.py:129
    return out


# ----------------------------------------------------------------------------- §8(f) row 1: decoder-side layers
# Functional restatement of CapEncoderLayer / AudioEncoderLayer (encoder.py:203-235), MultimodalDecoderLayer12
# (decoder.py:11-60) and MultimodalDecoder8 (decoder.py:62-186) for enc_st_combine == 'none' (the configuration of
# train_configs.py:40 / exec.sh; the other modes overwrite ft['spatiotemporal_ft'] and only run with one block,
# SURVEY.md §8 a2).  `cfg` carries the args fields the reference reads: nb_venc_blocks, nb_cenc_blocks,
# nb_aenc_blocks, enc_vc_combine, include_caption, dec_st_combine, t2s, s2t.  `masks`: trg, his, query, cap, audio,
# temporal.  Pinned by tests/golden/decoder_*.npz (outputs of the unmodified reference, oracle/make_golden.py).
def tgifqa_head(ft, query_mask, w, b):
    """Answer head of BASELINE config 4.  NOT a restatement of reference code: the TGIF-QA branch is not in the tree
    (/root/reference/README.md:105-106), so this is the build's own spec and parity for it is UNPINNED:
    masked mean over the query positions of LN(s2t) + LN(t2s) (= temporal_ft + spatial_ft, model/decoder.py:126-129),
    then Linear(d_model, n_answers)."""
    m = query_mask.reshape(query_mask.shape[0], -1, 1).to(ft["temporal_ft"].dtype)
    pooled = ((ft["temporal_ft"] + ft["spatial_ft"]) * m).sum(1) / m.sum(1).clamp_min(1.0)
    return linear(pooled, w, b)


def memory_encoder_layer(P, pre, h, x, mem, self_mask, mem_mask, drop=None):
    """CapEncoderLayer.forward / AudioEncoderLayer.forward: self attention, attention over the memory, FFN."""
    x = sublayer(P, pre + "sublayer.0.", x, lambda z: mha(P, pre + "attn.0.", h, z, z, z, self_mask, drop)[0], drop)
    x = sublayer(P, pre + "sublayer.1.", x, lambda z: mha(P, pre + "attn.1.", h, z, mem, mem, mem_mask, drop)[0], drop)
    return sublayer(P, pre + "sublayer.2.", x, lambda z: ffn(P, pre + "ff.", z, drop), drop)


def decoder_memories(cfg):
    """The attention memories MultimodalDecoderLayer12 walks after self / history / query attention, as
    (ft key, mask name) in order, and whether the two video streams are attended in parallel (decoder.py:27-57)."""
    v, c, a = cfg["nb_venc_blocks"] > 0, cfg["nb_cenc_blocks"] > 0, cfg.get("nb_aenc_blocks", 0) > 0
    if v and c and cfg.get("enc_vc_combine", "none") != "none":
        return [("encoded_ft", "query")], False                                    # :28-30
    mems = []
    if cfg["include_caption"] != "none":                                           # :32-37
        mems.append(("cap_ft", "query") if c else ("encoded_cap", "cap"))
    parallel = False
    if v:                                                                          # :38-54 (enc_st_combine == 'none')
        if cfg["dec_st_combine"] == "seq":
            if cfg["s2t"]:
                mems.append(("temporal_ft", "query"))
            if cfg["t2s"]:
                mems.append(("spatial_ft", "query"))
        else:
            mems += [("temporal_ft", "query"), ("spatial_ft", "query")]
            parallel = True
    if a:                                                                          # :55-57
        mems.append(("audio_ft", "query"))
    return mems, parallel


def decoder_layer12(P, pre, h, cfg, x, ft, masks, drop=None):
    """MultimodalDecoderLayer12.forward (decoder.py:20-60)."""
    def att(i, z, mem, mask):
        return sublayer(P, f"{pre}sublayer.{i}.", z,
                        lambda y: mha(P, f"{pre}attn.{i}.", h, y, y if mem is None else mem, y if mem is None else mem,
                                      mask, drop)[0], drop)
    x = att(0, x, None, masks["trg"])                                              # :21
    x = att(1, x, ft["encoded_his"], masks["his"])                                 # :22
    x = att(2, x, ft["encoded_query"], masks["query"])                             # :23
    mems, parallel = decoder_memories(cfg)
    i = 3
    for key, mname in mems:
        if parallel and key == "temporal_ft":                                      # :47-52 both from the same input
            t_x = att(i, x, ft["temporal_ft"], masks["query"])
            s_x = att(i + 1, x, ft["spatial_ft"], masks["query"])
            x = t_x + s_x
            i += 2
        elif parallel and key == "spatial_ft":
            continue
        else:
            x = att(i, x, ft[key], masks[mname])
            i += 1
    return sublayer(P, f"{pre}sublayer.{i}.", x, lambda z: ffn(P, pre + "ff.", z, drop), drop)   # :58


def vc_combine(P, pre, cfg, ft):
    """encoded_ft of decoder.py:137-167 (enc_st_combine == 'none'): sum, or softmax gate over the streams."""
    mode = cfg.get("enc_vc_combine", "none")
    a = cfg.get("nb_aenc_blocks", 0) > 0
    if mode == "sum":
        return ft["temporal_ft"] + ft["spatial_ft"] + ft["cap_ft"]                 # :141
    temp = [ft["encoded_query"], ft["cap_ft"]]                                     # :148
    if cfg["t2s"]:
        temp.append(ft["spatial_ft"])
    if cfg["s2t"]:
        temp.append(ft["temporal_ft"])
    if a:
        temp.append(ft["audio_ft"])
    score = torch.softmax(linear(torch.cat(temp, dim=-1), P[pre + "vc_combine_W.weight"], P[pre + "vc_combine_W.bias"]), dim=-1)
    if cfg["t2s"] and cfg["s2t"]:
        streams = ["temporal_ft", "spatial_ft", "cap_ft"]                          # :157-160
    elif not cfg["t2s"]:
        streams = ["temporal_ft", "cap_ft"]
    else:
        streams = ["spatial_ft", "cap_ft"]
    out = sum(score[:, :, i:i + 1] * ft[k] for i, k in enumerate(streams))
    if a:
        out = out + score[:, :, 3:4] * ft["audio_ft"]                              # :167 (index 3, as the reference)
    return out


def multimodal_decoder8(P, pre, h, cfg, ft, x, masks, n_blocks, drop=None):
    """MultimodalDecoder8.forward (decoder.py:107-186), enc_st_combine == 'none'.  ft must hold encoded_query,
    encoded_his, (encoded_cap), (spatiotemporal_ft), (encoded_audio); returns ft with decoded_text added."""
    ft = dict(ft)
    eq = ft["encoded_query"]
    cur = {"t2s": eq, "s2t": eq, "audio": eq, "cap": eq}                           # :108-112
    v, c, a = cfg["nb_venc_blocks"] > 0, cfg["nb_cenc_blocks"] > 0, cfg.get("nb_aenc_blocks", 0) > 0
    for blk in range(n_blocks):
        if v:                                                                      # :115-129
            cur["t2s"], cur["s2t"] = vid_encoder_layer4(
                P, f"{pre}v_layers.{blk}.", h, cur["t2s"], cur["s2t"], ft["spatiotemporal_ft"], masks["query"],
                masks["temporal"], t2s=bool(cfg["t2s"]), s2t=bool(cfg["s2t"]), drop=drop)
            if cfg["s2t"]:
                ft["temporal_ft"] = layer_norm(cur["s2t"], P[pre + "temporal_out_norm.a_2"], P[pre + "temporal_out_norm.b_2"])
            if cfg["t2s"]:
                ft["spatial_ft"] = layer_norm(cur["t2s"], P[pre + "spatial_out_norm.a_2"], P[pre + "spatial_out_norm.b_2"])
        if c:                                                                      # :130-132
            cur["cap"] = memory_encoder_layer(P, f"{pre}c_layers.{blk}.", h, cur["cap"], ft["encoded_cap"],
                                              masks["query"], masks["cap"], drop)
            ft["cap_ft"] = layer_norm(cur["cap"], P[pre + "cap_out_norm.a_2"], P[pre + "cap_out_norm.b_2"])
        if a:                                                                      # :133-135
            cur["audio"] = memory_encoder_layer(P, f"{pre}a_layers.{blk}.", h, cur["audio"], ft["encoded_audio"],
                                                masks["query"], masks["audio"], drop)
            ft["audio_ft"] = layer_norm(cur["audio"], P[pre + "a_out_norm.a_2"], P[pre + "a_out_norm.b_2"])
        if v and c and cfg.get("enc_vc_combine", "none") in ("sum", "dyn"):        # :137-167
            ft["encoded_ft"] = vc_combine(P, pre, cfg, ft)
        x = decoder_layer12(P, f"{pre}layers.{blk}.", h, cfg, x, ft, masks, drop)  # :182
    ft["decoded_text"] = layer_norm(x, P[pre + "norm.a_2"], P[pre + "norm.b_2"])   # :185
    return ft


# ----------------------------------------------------------------------------- §8(f) row 2: text side, generators, loss
# Embeddings / PositionalEncoding (modules.py:115-141), the LayerNorm-only text Encoder (encoder.py:11-40),
# Generator / PointerGenerator / MultiPointerGenerator (generator.py), LabelSmoothing (label_smoothing.py),
# SimpleLossCompute (optimize.py:36-94) and MTN.forward (mtn.py:31-61).  Pinned by tests/golden/model_*.npz.
def positional_table(max_len, d):
    """modules.py:127-136 (the registered buffer ``pe`` without its batch dimension)."""
    pe = torch.zeros(max_len, d)
    position = torch.arange(0., max_len).unsqueeze(1)
    div_term = torch.exp(torch.arange(0., d, 2) * -(math.log(10000.0) / d))
    pe[:, 0::2] = torch.sin(position * div_term)
    pe[:, 1::2] = torch.cos(position * div_term)
    return pe


def embed(P, pre, ids, drop=None, site=None):
    """query_embed = Sequential(Embeddings, PositionalEncoding) (mtn.py:78-80): lut(x) * sqrt(d) + pe, then dropout."""
    lut = P[pre + "0.lut.weight"]
    d = lut.shape[1]
    x = lut[ids] * math.sqrt(d) + positional_table(ids.shape[1], d).to(lut.device).unsqueeze(0)
    return _apply(drop, site, x)


def pointer_probs(P, pre, x, encoded_text, text_mask):
    """pointer_attn(x, text, text, mask).attn of a 1-head MultiHeadedAttention (generator.py:62-63): [B, Lt, Ls]."""
    q = linear(x, P[pre + "linears.0.weight"], P[pre + "linears.0.bias"])
    k = linear(encoded_text, P[pre + "linears.1.weight"], P[pre + "linears.1.bias"])
    s = q.matmul(k.transpose(-2, -1)) / math.sqrt(q.shape[-1])
    return torch.softmax(s.masked_fill(text_mask == 0, -1e9), dim=-1)


def _pointer_source(name, b, ft):
    return {"query": (b.query, ft["encoded_query"], b.query_mask), "his": (b.his, ft["encoded_his"], b.his_mask),
            "cap": (b.cap, ft["encoded_cap"], b.cap_mask)}[name]


def generator_logp(P, pre, ft, b, ptr_ft, mask_unk=1):
    """log p_out of PointerGenerator (one source, sigmoid switch, generator.py:29-75) or MultiPointerGenerator
    (several sources, softmax switch, generator.py:77-127) with the tied embedding ``vocab_gen``."""
    x = ft["decoded_text"]
    p_vocab = torch.softmax(x.matmul(P[pre + "vocab_gen"].t()), dim=-1)
    names = ptr_ft.split(",")
    multi = len(names) > 1
    ptrs, vecs = [], []
    for i, name in enumerate(names):
        text, enc, mask = _pointer_source(name, b, ft)
        if mask_unk:
            mask = mask & (text != 0).unsqueeze(-2)
        ap = pointer_probs(P, pre + (f"pointer_attn.{i}." if multi else "pointer_attn."), x, enc, mask)
        idx = text.unsqueeze(1).expand_as(ap)
        ptrs.append(torch.zeros_like(p_vocab).scatter_add(2, idx, ap))
        vecs.append(ap.matmul(enc))
    W, bias = P[pre + "pointer_gen_W.weight"], P[pre + "pointer_gen_W.bias"]
    if multi:
        sw = torch.softmax(linear(torch.cat([x, ft["encoded_tgt"]] + vecs, -1), W, bias), dim=-1)
        p_out = sum(sw[:, :, i:i + 1] * ptrs[i] for i in range(len(names))) + sw[:, :, -1:] * p_vocab
    else:
        s = torch.sigmoid(linear(torch.cat([x, vecs[0], ft["encoded_tgt"]], -1), W, bias))
        p_out = (1 - s) * ptrs[0] + s * p_vocab
    return torch.log(p_out)


def label_smoothing_kl(logp, target, pad, smoothing):
    """LabelSmoothing.forward (label_smoothing.py:20-30): summed KL(true_dist || exp(logp)); logp [R, V]."""
    V = logp.shape[1]
    t = torch.full_like(logp, smoothing / (V - 2))
    t.scatter_(1, target.unsqueeze(1), 1.0 - smoothing)
    t[:, pad] = 0
    t[target == pad] = 0
    return torch.where(t > 0, t * (t.clamp_min(1e-38).log() - logp), torch.zeros_like(t)).sum()


def loss_compute(P, args, ft, b, pad, smoothing=0.1):
    """SimpleLossCompute.__call__ (optimize.py:46-82) for enc_st_combine == 'none'.  Returns (loss, parts)."""
    V = P["generator.vocab_gen"].shape[0]
    out = generator_logp(P, "generator.", ft, b, args["ptr_ft"], args.get("mask_unk", 1))
    parts = {"out": label_smoothing_kl(out.reshape(-1, V), b.trg_y.reshape(-1), pad, smoothing) / b.ntokens.float()}
    if args.get("auto_encoder", 0):
        keys = (["cap_ft"] if args["nb_cenc_blocks"] > 0 else []) + (["audio_ft"] if args.get("nb_aenc_blocks", 0) > 0 else [])
        if args["nb_venc_blocks"] > 0:
            keys += (["temporal_ft"] if args["s2t"] else []) + (["spatial_ft"] if args["t2s"] else [])
        for k in keys:
            lp = torch.log_softmax(ft[k].matmul(P["ae_generator.proj"].t()), dim=-1)        # generator.py:22-25
            parts[k + "_ae"] = label_smoothing_kl(lp.reshape(-1, V), b.query.reshape(-1), pad, smoothing) / b.qntokens.float()
    return sum(parts.values()), parts


def mtn_encode(P, b, drop=None):
    """MTN.encode (mtn.py:36-52): embeddings -> text-encoder norms -> video projection; target-independent."""
    ft = {}
    eq, ec, eh = (embed(P, "query_embed.", ids, drop, f"query_embed.{n}") for n, ids in (("query", b.query), ("cap", b.cap), ("his", b.his)))
    for i, (key, x) in enumerate((("encoded_query", eq), ("encoded_cap", ec), ("encoded_his", eh))):   # encoder.py:19-40
        ft[key] = layer_norm(x, P[f"text_encoder.norm.{i}.a_2"], P[f"text_encoder.norm.{i}.b_2"])
    audio = getattr(b, "audio_fts", None) if "vid_encoder.a_W.weight" in P else None      # a_N > 0, encoder.py:83
    ft["spatiotemporal_ft"], enc_audio = vid_encoder8(P, "vid_encoder.", b.fts, audio)
    if enc_audio is not None:
        ft["encoded_audio"] = enc_audio
    return ft


def mtn_decode(P, args, b, ft, h, n_blocks, drop=None):
    """MTN.decode (mtn.py:54-62): target embedding -> multimodal decoder (which re-runs its video / caption blocks)."""
    ft = dict(ft)
    ft["encoded_tgt"] = embed(P, "tgt_embed.", b.trg, drop, "tgt_embed.trg")
    masks = dict(trg=b.trg_mask, his=b.his_mask, query=b.query_mask, cap=b.cap_mask, audio=getattr(b, "audio_mask", None),
                 temporal=b.temporal_mask)
    return multimodal_decoder8(P, "mutlimodal_decoder.", h, args, ft, ft["encoded_tgt"], masks, n_blocks, drop)


def mtn_forward(P, args, b, h, n_blocks, drop=None):
    """MTN.forward (mtn.py:31-34): encode, then decode."""
    return mtn_decode(P, args, b, mtn_encode(P, b, drop), h, n_blocks, drop)


# ----------------------------------------------------------------------------- §8(f) row 3: beam search
def beam_search(P, args, b, h, n_blocks, max_len, start_symbol, unk_symbol, end_symbol, beam=5, penalty=1.0, nbest=5,
                min_len=1, dec_eos=False):
    """beam_search_decode (decode.py:53-104) for ONE dialogue turn, in the reference's order of operations: one
    encode, then per step and per live hypothesis one full decode (video / caption blocks included) + generator on the
    last position, the completed-hypothesis list scored with the end symbol, and the fill-then-replace-the-minimum
    candidate selection over the descending argsort of the [V] scores.  Returns (n-best [(tokens, score)], best_state).
    Pinned by tests/golden/beam_*.npz (the unmodified reference's hypotheses)."""
    import copy
    import numpy as np
    ft = mtn_encode(P, b)
    hyps, comp, best = [([], 0., [int(start_symbol)])], [], None
    skip = (unk_symbol,) if dec_eos else (unk_symbol, end_symbol)
    for l in range(max_len):
        new, argmin = [], 0
        for out, lp, st in hyps:
            bb = copy.copy(b)
            bb.trg = torch.tensor([st], dtype=torch.long)
            bb.trg_mask = torch.tril(torch.ones(1, len(st), len(st), dtype=torch.bool))      # subsequent_mask
            f = mtn_decode(P, args, bb, ft, h, n_blocks)
            f["decoded_text"], f["encoded_tgt"] = f["decoded_text"][:, -1:], f["encoded_tgt"][:, -1:]
            logp = generator_logp(P, "generator.", f, bb, args["ptr_ft"], args.get("mask_unk", 1))
            lp_vec = np.squeeze(logp.detach().numpy() + lp)
            if l >= min_len:
                new_lp = lp_vec[end_symbol] + penalty * (len(out) + 1)
                comp.append((out, new_lp))
                if best is None or best < new_lp:
                    best = new_lp
            for o in np.argsort(lp_vec)[::-1]:
                if o in skip:
                    continue
                new_lp = lp_vec[o]
                if len(new) == beam:
                    if new[argmin][1] < new_lp:
                        new[argmin] = (out + [o], new_lp, st + [int(o)])
                        argmin = min(enumerate(new), key=lambda e: e[1][1])[0]
                    else:
                        break
                else:
                    new.append((out + [o], new_lp, st + [int(o)]))
                    if len(new) == beam:
                        argmin = min(enumerate(new), key=lambda e: e[1][1])[0]
        hyps = new
    if comp:
        return sorted(comp, key=lambda e: -e[1])[:nbest], best
    return [([], 0)], None
