"""GPU parity at BASELINE.json's FULL sizes (run with -m gpu on the B200 box).  The CUDA path runs the whole batch; the CPU
oracle (reference formulation, pinned by the goldens from the unmodified reference: tests/test_oracle*.py) runs a subset of
rows — dialogue turns are independent, so the oracle on rows R must equal rows R of the full-batch result.  Different
shapes select different kernels (fused stage-1 tiling, GEMM tile templates, stage-2 variants), so the small-shape goldens do
not cover these.  Tolerances are the north_star's: max|a-b|/max|b| <= 1e-4 fp32, <= 2e-2 bf16; greedy ids identical in fp32."""
import pytest
import torch

from bist_b200 import LabelSmoothing, SimpleLossCompute, VideoPath, synth
from bist_b200.core import FULL_SHAPES, PAD, SMOOTHING, full_model
from oracle import bist_oracle as O

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
TOL = {"fp32": 1e-4, "bf16": 2e-2}


def _video_case(c, precision, rows, seed=1):
    P = synth.make_params(0, c["d"], c["D"], c["n_blocks"])
    m = VideoPath(c["d"], c["h"], c["n_blocks"], c["D"], precision=precision).to(DEV)
    m.load_state_dict(P, strict=True)
    m.requires_grad_(False)
    b = synth.make_batch(seed, c["B"], c["T"], c["S"], c["D"], c["Lq"])
    eq = synth.make_encoded_query(seed, c["B"], c["Lq"], c["d"])
    with torch.no_grad():
        got = m(b.to(DEV), eq.to(DEV))
        ref = O.video_path(P, c["h"], b.fts[rows], eq[rows], b.query_mask[rows], b.temporal_mask[rows], c["n_blocks"])
    return got, ref, b


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_config2_full_size_vs_oracle(precision):
    "Config 2: B=64, T=40, S=16, D=2048, d=512, 3 blocks; oracle on four rows, three of them ragged."
    c = dict(B=64, T=40, S=16, D=2048, d=512, h=8, Lq=16, n_blocks=3)
    rows = [0, 5, 17, 63]
    got, ref, b = _video_case(c, precision, rows)
    assert (~b.temporal_mask[rows]).any() and (~b.query_mask[rows]).any()      # padded clips and padded query positions are present
    for k in ("spatiotemporal_ft", "temporal_ft", "spatial_ft"):
        e = O.norm_err(got[k][rows].float().cpu(), ref[k])
        assert e < TOL[precision], (precision, k, e)


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_config5_long_video_vs_oracle(precision):
    "Config 5 shape: S=49 (7x7), T=128, d=512, Lq=16 (B=4 here; bench.py runs B=8): one region per t2s tile, two clips per s2t tile."
    c = dict(B=4, T=128, S=49, D=2048, d=512, h=8, Lq=16, n_blocks=3)
    rows = [0, 3]
    got, ref, _ = _video_case(c, precision, rows, seed=5)
    for k in ("spatiotemporal_ft", "temporal_ft", "spatial_ft"):
        e = O.norm_err(got[k][rows].float().cpu(), ref[k])
        assert e < TOL[precision], (precision, k, e)


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_config4_tgifqa_shape_vs_oracle(precision):
    """Config 4 shape: B=128, T=64, S=16.  The encoder is pinned by the main-branch oracle; the answer head is this repo's own
    spec (the TGIF-QA branch is not in the reference tree, README.md:105-106) and is checked against its restatement."""
    from bist_b200.tgifqa import TGIF_SHAPES, TgifQaModel
    c = TGIF_SHAPES
    rows = [0, 3, 77, 127]
    torch.manual_seed(0)
    m = TgifQaModel(c["d"], c["h"], c["n_blocks"], c["D"], c["n_answers"], precision=precision).to(DEV)
    P = synth.make_params(0, c["d"], c["D"], c["n_blocks"])
    m.video.load_state_dict(P, strict=True)
    m.requires_grad_(False)
    b = synth.make_batch(7, c["B"], c["T"], c["S"], c["D"], c["Lq"])
    eq = synth.make_encoded_query(7, c["B"], c["Lq"], c["d"])
    with torch.no_grad():
        got = m(b.to(DEV), eq.to(DEV))
        ref = O.video_path(P, c["h"], b.fts[rows], eq[rows], b.query_mask[rows], b.temporal_mask[rows], c["n_blocks"])
        ref_logits = O.tgifqa_head(ref, b.query_mask[rows], m.answer.weight.float().cpu(), m.answer.bias.float().cpu())
    for k in ("temporal_ft", "spatial_ft"):
        e = O.norm_err(got[k][rows].float().cpu(), ref[k])
        assert e < TOL[precision], (precision, k, e)
    e = O.norm_err(got["answer_logits"][rows].float().cpu(), ref_logits)
    assert e < TOL[precision], (precision, "answer_logits", e)
    if precision == "fp32":
        assert torch.equal(got["answer_logits"][rows].argmax(-1).cpu(), ref_logits.argmax(-1))


@pytest.mark.parametrize("audio", [False, True], ids=["no_audio", "audio"])
def test_config3_full_model_vs_oracle(audio):
    """Config 3: the full model at B=64, V=4096, d=512, 3/3/3(/3) blocks in fp32: decoded text, generator log-probabilities,
    greedy ids and the un-normalised loss sums of a row subset against the oracle on those rows."""
    c = FULL_SHAPES
    torch.manual_seed(0)
    m, lc = full_model("fp32", audio=audio)
    m = m.to(DEV).eval()
    b = synth.make_full_batch(1, c["B"], c["T"], c["S"], c["D"], c["Lq"], c["Lh"], c["Lc"], c["Lt"], c["V"],
                              Ta=c["Ta"] if audio else 0, Da=c["Da"] if audio else 0)
    rows = (9, 12)                                                     # rows 9..11 as a sub-batch for the oracle
    sub = synth.batch_rows(b, *rows)
    sub.ntokens, sub.qntokens = (sub.trg_y != PAD).sum(), (sub.query != PAD).sum()
    P = {k: v.detach().float().cpu() for k, v in m.state_dict().items() if not k.endswith(".pe")}
    a = m.args
    args = dict(nb_venc_blocks=a.nb_venc_blocks, nb_cenc_blocks=a.nb_cenc_blocks, nb_aenc_blocks=a.nb_aenc_blocks if audio else 0,
                enc_vc_combine=a.enc_vc_combine, include_caption=a.include_caption, dec_st_combine=a.dec_st_combine,
                t2s=a.t2s, s2t=a.s2t, enc_st_combine="none", ptr_ft=a.ptr_ft, auto_encoder=a.auto_encoder, mask_unk=1)
    with torch.no_grad():
        bd = b.to(DEV)
        ft = m(bd)
        logp = m.generator(ft, bd, m.args)
        rft = O.mtn_forward(P, args, sub, a.att_h, a.nb_blocks)
        ref_total, ref_parts = O.loss_compute(P, args, rft, sub, PAD, SMOOTHING)
        # the same rows through the CUDA path as their own batch: the loss normalisers are per batch (optimize.py:46-50)
        sd = sub.to(DEV)
        sft = m(sd)
        total, parts = lc.loss(sft, sd)
    lo, hi = rows
    for k in ("decoded_text", "temporal_ft", "spatial_ft"):
        e = O.norm_err(ft[k][lo:hi].float().cpu(), rft[k])
        assert e < 1e-4, (k, e)
        assert O.norm_err(sft[k].float().cpu(), rft[k]) < 1e-4, k
    assert abs(float(total) - float(ref_total)) < 1e-4 * abs(float(ref_total)), (float(total), float(ref_total))
    # generator log-probabilities and greedy ids of those rows (taken from the FULL-batch run)
    slogp = m.generator(sft, sd, m.args)
    assert O.norm_err(logp[lo:hi].cpu(), slogp.cpu()) < 1e-5
    ref_lp = O.generator_logp(P, "generator.", rft, sub, a.ptr_ft, 1)
    assert O.norm_err(slogp.cpu(), ref_lp) < 1e-4
    assert torch.equal(slogp.argmax(-1).cpu(), ref_lp.argmax(-1))


def test_config5_beam_search_turn_vs_oracle():
    "One dialogue turn of the config-5 shape through beam search (beam 3, max_len 4 to bound the oracle's time): ids identical, scores 1e-4."
    from bist_b200.decode import beam_search_decode
    shape = dict(B=1, T=128, S=49)
    c = dict(FULL_SHAPES, **shape)
    torch.manual_seed(0)
    m, _ = full_model("fp32", shape)
    m = m.to(DEV).eval()
    b = synth.make_full_batch(5, 1, c["T"], c["S"], c["D"], c["Lq"], c["Lh"], c["Lc"], c["Lt"], c["V"])
    P = {k: v.detach().float().cpu() for k, v in m.state_dict().items() if not k.endswith(".pe")}
    a = m.args
    args = dict(nb_venc_blocks=a.nb_venc_blocks, nb_cenc_blocks=a.nb_cenc_blocks, nb_aenc_blocks=0,
                enc_vc_combine=a.enc_vc_combine, include_caption=a.include_caption, dec_st_combine=a.dec_st_combine,
                t2s=a.t2s, s2t=a.s2t, enc_st_combine="none", ptr_ft=a.ptr_ft, auto_encoder=a.auto_encoder, mask_unk=1)
    with torch.no_grad():
        ref_h, ref_best = O.beam_search(P, args, b, a.att_h, a.nb_blocks, 4, 2, 0, 3, beam=3, penalty=1.0, nbest=3, min_len=1)
        got_h, got_best = beam_search_decode(m, b.to(DEV), 4, 2, 0, 3, 1, beam=3, penalty=1.0, nbest=3, min_len=1, train_args=m.args)
    assert [[int(t) for t in s] for s, _ in got_h] == [[int(t) for t in s] for s, _ in ref_h]
    for (_, gs), (_, rs) in zip(got_h, ref_h):
        assert abs(float(gs) - float(rs)) < 1e-4 * max(1.0, abs(float(rs)))
