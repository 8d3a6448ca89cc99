#!/usr/bin/env python
"""Generate tests/golden/*.npz by running the UNMODIFIED reference (TEST INFRASTRUCTURE).

Run in the build container only (needs /root/reference):

    python oracle/make_golden.py            # rewrites tests/golden/

For every case below the reference model is built with ``model.mtn.make_model``
(/root/reference/model/mtn.py:63-167), its video-path parameters are overwritten with the
seeded values of ``bist_b200.synth.make_params`` (checked to be a key-for-key, shape-for-shape
subset of the reference ``state_dict``), and the hot path is driven exactly as
``MultimodalDecoder8.forward`` drives it (/root/reference/model/decoder.py:107-129).  Only the
case description (seeds + shapes) and the reference OUTPUTS are stored; inputs and weights
are regenerated from the seeds by the tests, which keeps the fixtures small.

A second fixture (modules.npz) pins the generic blocks of model/modules.py
(LayerNorm, MultiHeadedAttention incl. ``.attn``, PositionwiseFeedForward,
SublayerConnection) on their own.
"""
import argparse
import json
import os
import sys
import types
import warnings

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
REF = os.environ.get("BIST_REFERENCE", "/root/reference")

from bist_b200 import synth  # noqa: E402

CASES = {
    # name: dict(seed, B, T, S, D, d, h, Lq, n_blocks, t2s, s2t [, full_clips])
    "both_small":  dict(seed=11, B=3, T=5, S=4, D=96, d=64, h=4, Lq=6, n_blocks=2, t2s=1, s2t=1),
    "t2s_only":    dict(seed=12, B=2, T=4, S=3, D=80, d=64, h=4, Lq=5, n_blocks=2, t2s=1, s2t=0),
    "s2t_only":    dict(seed=13, B=2, T=4, S=3, D=80, d=64, h=4, Lq=5, n_blocks=2, t2s=0, s2t=1),
    "lq1_ragged":  dict(seed=14, B=2, T=3, S=7, D=40, d=32, h=2, Lq=1, n_blocks=1, t2s=1, s2t=1),
    "exec_sh_d128": dict(seed=15, B=2, T=6, S=16, D=256, d=128, h=8, Lq=5, n_blocks=1, t2s=1, s2t=1),
    "long_video":  dict(seed=16, B=1, T=33, S=49, D=64, d=64, h=8, Lq=3, n_blocks=1, t2s=1, s2t=1),
    "cfg1_d512":   dict(seed=17, B=4, T=8, S=16, D=2048, d=512, h=8, Lq=16, n_blocks=3, t2s=1, s2t=1),
}


def import_reference():
    """Import /root/reference unmodified; stub the absent, unused ``nltk``."""
    if not os.path.isdir(REF):
        raise SystemExit(f"reference tree not found at {REF}")
    nltk = types.ModuleType("nltk")
    nltk_util = types.ModuleType("nltk.util")
    nltk_util.ngrams = lambda *a, **k: []
    nltk.util = nltk_util
    sys.modules.setdefault("nltk", nltk)
    sys.modules.setdefault("nltk.util", nltk_util)
    if REF not in sys.path:
        sys.path.insert(0, REF)
    warnings.filterwarnings("ignore")
    import model.mtn as mtn            # noqa
    import model.modules as modules    # noqa
    return mtn, modules


def ref_args(c):
    return argparse.Namespace(
        nb_blocks=c["n_blocks"], nb_venc_blocks=c["n_blocks"], nb_cenc_blocks=0, nb_aenc_blocks=0,
        d_model=c["d"], att_h=c["h"], dropout=0.1, ptr_gen=1, ptr_ft="query", auto_encoder=0,
        t2s=c["t2s"], s2t=c["s2t"], enc_st_combine="none", enc_vc_combine="none",
        dec_st_combine="seq", include_caption="none", mask_unk=1)


def run_reference_case(mtn, c):
    torch.manual_seed(0)
    model = mtn.make_model(64, 64, ref_args(c), ft_sizes=[c["D"]]).eval()
    P = synth.make_params(c["seed"], c["d"], c["D"], c["n_blocks"], bool(c["t2s"]), bool(c["s2t"]))
    sd = model.state_dict()
    for k, v in P.items():                       # pins the parameter names and shapes (§8(b))
        assert k in sd, f"{k} is not a reference state_dict key"
        assert tuple(sd[k].shape) == tuple(v.shape), (k, sd[k].shape, v.shape)
    vid_keys = [k for k in sd if k.startswith("vid_encoder.") or ".v_layers." in k
                or "spatial_out_norm" in k or "temporal_out_norm" in k]
    assert sorted(vid_keys) == sorted(P.keys()), "synth.make_params must cover the whole video path"
    model.load_state_dict(P, strict=False)

    b = synth.make_batch(c["seed"], c["B"], c["T"], c["S"], c["D"], c["Lq"])
    eq = synth.make_encoded_query(c["seed"], c["B"], c["Lq"], c["d"])
    out = {}
    with torch.no_grad():
        ft = {"encoded_query": eq}
        ft = model.vid_encoder(b, ft)                                 # mtn.py:49-51
        dec = model.mutlimodal_decoder
        in_ft = {"t2s": eq, "s2t": eq}                                # decoder.py:109-110
        for blk in range(c["n_blocks"]):
            in_ft = dec.v_layers[blk](in_ft, ft, b)                   # decoder.py:116
            if c["t2s"]:
                out[f"t2s_{blk}"] = in_ft["t2s"].numpy().copy()
            if c["s2t"]:
                out[f"s2t_{blk}"] = in_ft["s2t"].numpy().copy()
        if c["s2t"]:
            out["temporal_ft"] = dec.temporal_out_norm(in_ft["s2t"]).numpy()   # decoder.py:127
        if c["t2s"]:
            out["spatial_ft"] = dec.spatial_out_norm(in_ft["t2s"]).numpy()     # decoder.py:129
    vft = ft["spatiotemporal_ft"].numpy()
    if vft.size > 200_000:                       # keep the fixture small: strided sample
        out["vft_sample"] = vft[:, ::3, ::5, :].copy()
        out["vft_stride"] = np.array([3, 5])
    else:
        out["vft_sample"] = vft
        out["vft_stride"] = np.array([1, 1])
    return out


def run_modules_fixture(modules):
    """Generic blocks of model/modules.py on their own."""
    rs = np.random.RandomState(99)
    d, h, Bp, Lq, Lk, dff = 48, 4, 3, 5, 7, 96
    out = {"meta": np.array(json.dumps(dict(d=d, h=h, Bp=Bp, Lq=Lq, Lk=Lk, dff=dff)))}
    x = torch.from_numpy(rs.standard_normal((Bp, Lq, d)).astype(np.float32))
    kv = torch.from_numpy(rs.standard_normal((Bp, Lk, d)).astype(np.float32))
    mask_k = torch.from_numpy(rs.uniform(size=(Bp, 1, Lk)) > 0.3)
    mask_k[:, :, 0] = True
    mask_qk = torch.from_numpy(rs.uniform(size=(Bp, Lq, Lk)) > 0.3)
    mask_qk[:, :, 0] = True
    mask_qk[1, 2, :] = False                      # a fully masked row -> uniform probabilities
    out.update(x=x.numpy(), kv=kv.numpy(), mask_k=mask_k.numpy(), mask_qk=mask_qk.numpy())

    ln = modules.LayerNorm(d)
    att = modules.MultiHeadedAttention(h, d).eval()
    ff = modules.PositionwiseFeedForward(d, dff).eval()
    sc = modules.SublayerConnection(d, 0.1).eval()
    with torch.no_grad():
        for name, m in (("ln", ln), ("att", att), ("ff", ff), ("sc", sc)):
            for k, p in m.state_dict().items():
                v = rs.standard_normal(tuple(p.shape)).astype(np.float32) * (0.2 if p.dim() > 1 else 0.1)
                if k.endswith("a_2"):
                    v = 1.0 + v
                p.copy_(torch.from_numpy(v))
                out[f"{name}.{k}"] = v
        out["ln_out"] = ln(x).numpy()
        out["att_out_k"] = att(x, kv, kv, mask_k).numpy()
        out["att_p_k"] = att.attn.numpy().copy()
        out["att_out_qk"] = att(x, kv, kv, mask_qk).numpy()
        out["att_p_qk"] = att.attn.numpy().copy()
        out["att_out_nomask"] = att(x, kv, kv, None).numpy()
        out["ff_out"] = ff(x).numpy()
        out["sc_out"] = sc(x, ff).numpy()
    return out


# ---- decoder side: MultimodalDecoder8 with its caption / audio / decoder layers ----------
DEC_CASES = {
    # exec.sh-like AVSD setting: caption encoder + softmax gate over (temporal, spatial, caption)
    "avsd_dyn":     dict(seed=31, B=2, Lq=5, Lh=9, Lc=7, Lt=6, Ta=4, T=4, S=3, d=64, h=4, n_blocks=2, nb_cenc=2, nb_aenc=0,
                         vc="dyn", cap="summary", dec_st="seq", t2s=1, s2t=1),
    "audio_dyn":    dict(seed=32, B=2, Lq=4, Lh=6, Lc=5, Lt=5, Ta=6, T=3, S=4, d=64, h=4, n_blocks=2, nb_cenc=2, nb_aenc=2,
                         vc="dyn", cap="summary", dec_st="seq", t2s=1, s2t=1),
    "vc_sum":       dict(seed=33, B=2, Lq=4, Lh=6, Lc=5, Lt=5, Ta=3, T=3, S=4, d=32, h=2, n_blocks=1, nb_cenc=1, nb_aenc=0,
                         vc="sum", cap="summary", dec_st="seq", t2s=1, s2t=1),
    "nocap_seq":    dict(seed=34, B=3, Lq=5, Lh=7, Lc=4, Lt=6, Ta=3, T=4, S=3, d=64, h=8, n_blocks=2, nb_cenc=0, nb_aenc=0,
                         vc="none", cap="none", dec_st="seq", t2s=1, s2t=1),
    # (include_caption != 'none' with nb_cenc_blocks == 0 and video blocks indexes one attention past what
    #  mtn.py:139-150 allocates: the reference itself raises IndexError, so there is no such case)
    "nocap_par":    dict(seed=35, B=2, Lq=5, Lh=7, Lc=6, Lt=4, Ta=3, T=4, S=3, d=64, h=4, n_blocks=1, nb_cenc=0, nb_aenc=0,
                         vc="none", cap="none", dec_st="par", t2s=1, s2t=1),
    "cenc_aenc":    dict(seed=36, B=2, Lq=5, Lh=7, Lc=6, Lt=4, Ta=5, T=4, S=3, d=64, h=4, n_blocks=1, nb_cenc=1, nb_aenc=1,
                         vc="none", cap="summary", dec_st="seq", t2s=1, s2t=1),
    "s2t_only_dyn": dict(seed=37, B=2, Lq=5, Lh=7, Lc=6, Lt=4, Ta=5, T=4, S=3, d=64, h=4, n_blocks=1, nb_cenc=1, nb_aenc=0,
                         vc="dyn", cap="summary", dec_st="seq", t2s=0, s2t=1),
}


def dec_args(c):
    return argparse.Namespace(
        nb_blocks=c["n_blocks"], nb_venc_blocks=c["n_blocks"], nb_cenc_blocks=c["nb_cenc"], nb_aenc_blocks=c["nb_aenc"],
        d_model=c["d"], att_h=c["h"], dropout=0.1, ptr_gen=1, ptr_ft="query", auto_encoder=0,
        t2s=c["t2s"], s2t=c["s2t"], enc_st_combine="none", enc_vc_combine=c["vc"],
        dec_st_combine=c["dec_st"], include_caption=c["cap"], mask_unk=1)


def run_decoder_case(mtn, c):
    torch.manual_seed(0)
    model = mtn.make_model(64, 64, dec_args(c), ft_sizes=[48, 16]).eval()
    dec = model.mutlimodal_decoder
    shapes = {k: tuple(v.shape) for k, v in dec.state_dict().items()}
    dec.load_state_dict(synth.fill_params(shapes, c["seed"]), strict=True)
    b, ft, x = synth.make_decoder_inputs(c["seed"], c["B"], c["Lq"], c["Lh"], c["Lc"], c["Lt"], c["Ta"], c["T"], c["S"], c["d"])
    with torch.no_grad():
        out_ft = dec(b, dict(ft), x)                                   # decoder.py:107-186, as mtn.py:61 calls it
    out = {k: out_ft[k].numpy().copy() for k in ("decoded_text", "temporal_ft", "spatial_ft", "cap_ft", "audio_ft",
                                                  "encoded_ft") if k in out_ft}
    out["param_shapes"] = np.array(json.dumps(sorted([k, list(v)] for k, v in shapes.items())))
    return out


# ---- full model: MTN.forward + SimpleLossCompute of the unmodified reference --------------
MODEL_CASES = {
    # exec.sh-like: caption encoder, softmax gate, two pointer sources, auto-encoder heads
    "avsd_full":   dict(seed=41, B=3, T=4, S=3, D=40, d=64, h=4, Lq=6, Lh=9, Lc=7, Lt=5, V=97, n_blocks=2, nb_cenc=2,
                        vc="dyn", cap="summary", ptr_ft="query,cap", ae=1, t2s=1, s2t=1),
    "single_ptr":  dict(seed=42, B=2, T=3, S=4, D=32, d=32, h=2, Lq=5, Lh=6, Lc=4, Lt=4, V=50, n_blocks=1, nb_cenc=0,
                        vc="none", cap="none", ptr_ft="query", ae=1, t2s=1, s2t=1),
    "his_ptr_noae": dict(seed=43, B=2, T=3, S=4, D=32, d=32, h=4, Lq=5, Lh=8, Lc=4, Lt=6, V=64, n_blocks=1, nb_cenc=1,
                         vc="sum", cap="summary", ptr_ft="his,query", ae=0, t2s=1, s2t=1),
    # BASELINE config 3 names the audio encoder: ft_sizes=[D, Da] builds VidEncoder8.a_W / a_in_norm (mtn.py:112-113,
    # encoder.py:83-91, with args.noW_venc=0 injected), AudioEncoderLayer blocks, a fourth gate column
    # (decoder.py:154,167) and the audio auto-encoder head (optimize.py:61-65)
    "avsd_audio":  dict(seed=44, B=3, T=4, S=3, D=40, d=64, h=4, Lq=6, Lh=8, Lc=7, Lt=5, V=83, n_blocks=2, nb_cenc=2,
                        nb_aenc=2, Ta=6, Da=12, vc="dyn", cap="summary", ptr_ft="query,cap", ae=1, t2s=1, s2t=1),
    "audio_seq":   dict(seed=45, B=2, T=3, S=4, D=32, d=32, h=2, Lq=5, Lh=6, Lc=4, Lt=4, V=50, n_blocks=1, nb_cenc=0,
                        nb_aenc=1, Ta=5, Da=10, vc="none", cap="none", ptr_ft="query", ae=1, t2s=1, s2t=1),
}
PAD, SMOOTHING = 1, 0.1          # train.py:93 (vocab['<blank>'] is the collate pad id, data/dataset.py)


def model_args(c):
    return argparse.Namespace(
        nb_blocks=c["n_blocks"], nb_venc_blocks=c["n_blocks"], nb_cenc_blocks=c["nb_cenc"], nb_aenc_blocks=c.get("nb_aenc", 0),
        d_model=c["d"], att_h=c["h"], dropout=0.1, ptr_gen=1, ptr_ft=c["ptr_ft"], auto_encoder=c["ae"],
        t2s=c["t2s"], s2t=c["s2t"], enc_st_combine="none", enc_vc_combine=c["vc"],
        dec_st_combine="seq", include_caption=c["cap"], mask_unk=1, noW_venc=0)


def model_ft_sizes(c):
    return [c["D"]] + ([c["Da"]] if c.get("nb_aenc", 0) > 0 else [])


def model_batch(c):
    return synth.make_full_batch(c["seed"], c["B"], c["T"], c["S"], c["D"], c["Lq"], c["Lh"], c["Lc"], c["Lt"], c["V"],
                                 Ta=c.get("Ta", 0), Da=c.get("Da", 0))


TIED = ("tgt_embed.0.lut.weight", "generator.vocab_gen", "ae_generator.proj")


def model_params(shapes, seed):
    """fill_params over the whole state_dict; the tied copies of the embedding take the query_embed values and the
    positional tables (buffers) are left to the model."""
    P = synth.fill_params({k: v for k, v in shapes.items() if not k.endswith(".pe")}, seed)
    for k in TIED:
        if k in P:
            P[k] = P["query_embed.0.lut.weight"]
    return P


def run_model_case(mtn, c):
    import model.label_smoothing as ls
    import model.optimize as opt
    torch.Tensor.cuda = lambda self, *a, **k: self        # generator.py:66,113 hard-code .cuda() (harness shim, §8(c))
    torch.manual_seed(0)
    args = model_args(c)
    model = mtn.make_model(c["V"], c["V"], args, ft_sizes=model_ft_sizes(c)).eval()
    shapes = {k: tuple(v.shape) for k, v in model.state_dict().items()}
    model.load_state_dict(model_params(shapes, c["seed"]), strict=False)
    b = model_batch(c)
    crit = ls.LabelSmoothing(size=c["V"], padding_idx=PAD, smoothing=SMOOTHING)
    with torch.no_grad():
        ft = model.forward(b)
        logp = model.generator(ft, b, args)
        lc = opt.SimpleLossCompute(model.generator, model.ae_generator, crit, opt=None, args=args)
        losses = lc(ft, b)
        # total normalised loss exactly as optimize.py:48-82 sums it
        total = crit(logp.contiguous().view(-1, c["V"]), b.trg_y.contiguous().view(-1)) / b.ntokens.float()
        if c["ae"]:
            for k in (["cap_ft"] if c["nb_cenc"] > 0 else []) + (["audio_ft"] if c.get("nb_aenc", 0) > 0 else []) + ["temporal_ft", "spatial_ft"]:
                ae = model.ae_generator(ft, b, args, k)
                total = total + crit(ae.contiguous().view(-1, c["V"]), b.query.contiguous().view(-1)) / b.qntokens.float()
    out = {k: ft[k].numpy().copy() for k in ("encoded_query", "encoded_his", "encoded_cap", "encoded_tgt", "decoded_text")}
    for k in ("encoded_audio", "audio_ft"):
        if k in ft:
            out[k] = ft[k].numpy().copy()
    out["logp"] = logp.numpy().copy()
    out["loss_out_sum"] = np.array(float(losses["out"]))
    out["loss_total"] = np.array(float(total))
    out["param_shapes"] = np.array(json.dumps(sorted([k, list(v)] for k, v in shapes.items())))
    return out


# ---- beam search: decode.py:53-104 of the unmodified reference, one turn at a time ----------
BEAM = dict(beam=3, max_len=6, nbest=3, min_len=1, penalty=1.0, start=2, unk=0, end=3, pad=1)
BEAM_CASES = ("avsd_full", "single_ptr")


def run_beam_case(mtn, c):
    import model.decode as dec
    torch.Tensor.cuda = lambda self, *a, **k: self
    torch.manual_seed(0)
    args = model_args(c)
    model = mtn.make_model(c["V"], c["V"], args, ft_sizes=model_ft_sizes(c)).eval()
    shapes = {k: tuple(v.shape) for k, v in model.state_dict().items()}
    model.load_state_dict(model_params(shapes, c["seed"]), strict=False)
    b = model_batch(c)
    out = {}
    g = BEAM
    with torch.no_grad():
        for i in range(c["B"]):
            bi = synth.batch_rows(b, i, i + 1)
            hyps, best = dec.beam_search_decode(model, bi, g["max_len"], g["start"], g["unk"], g["end"], g["pad"], beam=g["beam"],
                                                penalty=g["penalty"], nbest=g["nbest"], min_len=g["min_len"], train_args=args)
            toks = np.full((g["nbest"], g["max_len"]), -1, dtype=np.int64)
            for j, (seq, _s) in enumerate(hyps):
                toks[j, :len(seq)] = [int(t) for t in seq]
            out[f"tokens_{i}"] = toks
            out[f"scores_{i}"] = np.array([float(s) for _seq, s in hyps], dtype=np.float64)
            out[f"best_{i}"] = np.array(float(best))
    out["beam"] = np.array(json.dumps(g))
    return out


def main():
    mtn, modules = import_reference()
    gold = os.path.join(ROOT, "tests", "golden")
    os.makedirs(gold, exist_ok=True)
    only = set(sys.argv[1:])      # e.g. `make_golden.py model_avsd_audio`: (re)generate the named model fixtures only
    if only:
        for name, c in MODEL_CASES.items():
            if f"model_{name}" in only:
                out = run_model_case(mtn, c)
                out["meta"] = np.array(json.dumps(c))
                path = os.path.join(gold, f"model_{name}.npz")
                np.savez_compressed(path, **out)
                print(f"{path}: {os.path.getsize(path) / 1024:.1f} KiB  loss_total={float(out['loss_total']):.6f}")
        return
    for name, c in CASES.items():
        out = run_reference_case(mtn, c)
        out["meta"] = np.array(json.dumps(c))
        path = os.path.join(gold, f"path_{name}.npz")
        np.savez_compressed(path, **out)
        print(f"{path}: {os.path.getsize(path) / 1024:.1f} KiB")
    for name, c in DEC_CASES.items():
        out = run_decoder_case(mtn, c)
        out["meta"] = np.array(json.dumps(c))
        path = os.path.join(gold, f"decoder_{name}.npz")
        np.savez_compressed(path, **out)
        print(f"{path}: {os.path.getsize(path) / 1024:.1f} KiB", sorted(k for k in out if k not in ("meta", "param_shapes")))
    for name, c in MODEL_CASES.items():
        out = run_model_case(mtn, c)
        out["meta"] = np.array(json.dumps(c))
        path = os.path.join(gold, f"model_{name}.npz")
        np.savez_compressed(path, **out)
        print(f"{path}: {os.path.getsize(path) / 1024:.1f} KiB  loss_total={float(out['loss_total']):.6f}")
    for name in BEAM_CASES:
        c = MODEL_CASES[name]
        out = run_beam_case(mtn, c)
        out["meta"] = np.array(json.dumps(c))
        path = os.path.join(gold, f"beam_{name}.npz")
        np.savez_compressed(path, **out)
        print(f"{path}: {os.path.getsize(path) / 1024:.1f} KiB", out["tokens_0"].tolist(), out["scores_0"].tolist())
    path = os.path.join(gold, "modules.npz")
    np.savez_compressed(path, **run_modules_fixture(modules))
    print(f"{path}: {os.path.getsize(path) / 1024:.1f} KiB")


if __name__ == "__main__":
    main()
