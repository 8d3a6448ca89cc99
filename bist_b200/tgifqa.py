"""BASELINE config 4: the TGIF-QA video-QA variant — B=128 turns, T=64 clips, S=16 regions, answer-classification head.

The TGIF-QA branch is NOT in the reference tree (/root/reference/README.md:105-106), so there is no reference head to match:
this head is this repo's OWN SPEC, as SURVEY.md §8(c) prescribes — masked mean over the Lq query positions of
``LN(s2t) + LN(t2s)`` (= ft['temporal_ft'] + ft['spatial_ft'] of model/decoder.py:126-129) followed by
``Linear(d_model, n_answers)``.  The ENCODER under it (VidEncoder8 + VidEncoderLayer4 blocks + output norms) is the
main-branch one and is pinned by the main-branch oracle at this shape (tests/test_gpu_sizes.py).  Multi-GPU: batch-sharded
replicas, no collective (inference)."""
import torch
import torch.nn as nn

from . import _lib
from .encoder import VideoPath
from .modules import _BF16Cache, linear

TGIF_SHAPES = dict(B=128, T=64, S=16, D=2048, d=512, h=8, Lq=16, n_answers=1000, n_blocks=3)


class TgifQaModel(nn.Module):
    def __init__(self, d_model=512, h=8, n_blocks=3, ft_size=2048, n_answers=1000, precision="fp32"):
        super().__init__()
        self.video = VideoPath(d_model, h, n_blocks, ft_size, precision=precision)
        self.answer = nn.Linear(d_model, n_answers)
        self._cache = _BF16Cache()
        self.eval()

    def forward(self, b, encoded_query):
        ft = self.video(b, encoded_query)
        m = b.query_mask.reshape(b.query_mask.shape[0], -1, 1).to(ft["temporal_ft"].dtype)       # [B, Lq, 1]
        pooled = ((ft["temporal_ft"] + ft["spatial_ft"]) * m).sum(1) / m.sum(1).clamp_min(1.0)    # masked mean over Lq
        ft["answer_logits"] = linear(pooled.contiguous(), self.answer, self._cache)
        return ft


def bench_leg(args, env):
    "bench.py object `tgifqa`: turns/s of the config-4 forward per N GPUs (replicas, no collective)."
    from . import synth
    dev, rank, world, timed = env["dev"], env["rank"], env["world"], env["timed"]
    c = TGIF_SHAPES
    torch.manual_seed(0)
    m = TgifQaModel(c["d"], c["h"], c["n_blocks"], c["D"], c["n_answers"], precision=args.precision).to(dev)
    m.video.load_state_dict(synth.make_params(0, c["d"], c["D"], c["n_blocks"]), strict=True)
    b = synth.make_batch(7 + rank, c["B"], c["T"], c["S"], c["D"], c["Lq"]).to(dev)
    eq = synth.make_encoded_query(7 + rank, c["B"], c["Lq"], c["d"]).to(dev)
    with torch.no_grad():
        for _ in range(3):
            m(b, eq)
        g = torch.cuda.CUDAGraph()
        s = torch.cuda.Stream(device=dev)
        s.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(s):
            m(b, eq)
        torch.cuda.current_stream(dev).wait_stream(s)
        with torch.cuda.graph(g):
            out = m(b, eq)
        for _ in range(3):
            g.replay()
        n = max(3, min(args.steps, 10))
        ms = timed(g.replay, n) / n
    assert torch.isfinite(out["answer_logits"].float()).all()
    N = c["B"] * c["T"] * c["S"]
    e = 2 if args.precision == "bf16" else 4
    alg = N * c["D"] * 4 + N * c["d"] * e * (1 + c["n_blocks"])        # fts read + vft write + one vft read per block
    peaks = env["peaks"]
    return {"metric": "tgifqa_turns_per_sec", "value": world * c["B"] / (ms * 1e-3), "unit": "turns/s", "ms_per_step": ms,
            "workload": "config 4: B=128 T=64 S=16 encoder + answer head (own spec: masked mean of LN(s2t)+LN(t2s) -> Linear(d, 1000)), replicas",
            "encoder_algorithmic_gbs": alg / (ms * 1e-3) / 1e9, "encoder_frac_of_hbm_peak": alg / (ms * 1e-3) / 1e9 / peaks["hbm"]}
