"""Feature ingest.  CPU: FeatureCache.collate against the reference's collate semantics
(data/dataset.py:164-174 zero-pad to the longest video, :79 temporal mask), ragged inputs, both storage dtypes and both
array layouts.  GPU: a bf16 feature batch through the 2-byte projection against the fp32 batch through the fused kernel,
and the whole video path on cached features against the oracle."""
import numpy as np
import pytest
import torch

from bist_b200.ingest import FeatureCache, reference_collate


def _videos(seed, lens, S, D):
    rs = np.random.RandomState(seed)
    return [rs.uniform(0, 1, size=(n, S, D)).astype(np.float32) for n in lens]


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_collate_matches_reference_collate(dtype):
    lens, S, D = [5, 2, 7, 1], 4, 24
    vids = _videos(3, lens, S, D)
    cache = FeatureCache(D, S, capacity_clips=sum(lens), dtype=dtype, pin=False)
    for i, v in enumerate(vids):
        assert cache.add("v%d" % i, v) == lens[i]
    order = ["v2", "v0", "v3", "v1"]
    fts, mask = cache.collate(order, "cpu")
    ref, ref_mask = reference_collate([vids[2], vids[0], vids[3], vids[1]])
    assert fts.shape == ref.shape and fts.dtype == dtype
    assert torch.equal(fts.float(), ref.to(dtype).float())          # fp32: exact; bf16: the round-to-nearest-even value
    assert torch.equal(mask, ref_mask)
    fts2, mask2 = cache.collate(["v1", "v3"], "cpu", T=4)           # fixed T (static graph inputs): padded beyond the batch max
    assert fts2.shape == (2, 4, S, D) and not mask2[0, 0, 2:].any() and float(fts2[0, 2:].abs().sum()) == 0.0
    assert cache.host_bytes(order) == sum(lens) * S * D * (4 if dtype == torch.float32 else 2)


def test_layouts_and_capacity():
    T, H, Wd, C = 3, 2, 2, 8
    rs = np.random.RandomState(1)
    x = rs.uniform(size=(T, H, Wd, C)).astype(np.float32)           # channel-last, as the loader assumes (dataset.py:151)
    cache = FeatureCache(C, H * Wd, capacity_clips=2 * T, dtype=torch.float32, pin=False)
    cache.add("last", x)
    cache.add("first", np.transpose(x, (0, 3, 1, 2)), channels_first=True)   # what the extractor writes (resnext.py:136)
    a, _ = cache.collate(["last"], "cpu")
    b, _ = cache.collate(["first"], "cpu")
    assert torch.equal(a, b) and torch.equal(a[0], torch.from_numpy(x.reshape(T, H * Wd, C)))
    with pytest.raises(MemoryError):
        cache.add("more", x)
    with pytest.raises(ValueError):
        FeatureCache(C, 3, 4, pin=False).add("bad", x)


@pytest.mark.gpu
def test_bf16_features_through_the_two_byte_projection():
    from bist_b200 import VideoPath, synth
    from oracle import bist_oracle as O
    c = dict(seed=611, B=3, T=6, S=4, D=128, d=64, h=4, Lq=5, n_blocks=2)
    lens = [6, 3, 4]
    vids = _videos(c["seed"], lens, c["S"], c["D"])
    cache = FeatureCache(c["D"], c["S"], capacity_clips=sum(lens), dtype=torch.bfloat16)
    for i, v in enumerate(vids):
        cache.add(i, v)
    fts16, tmask = cache.collate([0, 1, 2], "cuda:0")
    ref_fts, ref_mask = reference_collate(vids)
    assert torch.equal(tmask.cpu(), ref_mask)
    P = synth.make_params(c["seed"], c["d"], c["D"], c["n_blocks"])
    m = VideoPath(c["d"], c["h"], c["n_blocks"], c["D"], precision="bf16").to("cuda:0")
    m.load_state_dict(P, strict=True)
    b = synth.make_batch(c["seed"], c["B"], c["T"], c["S"], c["D"], c["Lq"])
    eq = synth.make_encoded_query(c["seed"], c["B"], c["Lq"], c["d"])
    bd = b.to("cuda:0")
    with torch.no_grad():
        bd.fts, bd.temporal_mask = fts16, tmask
        out16 = m(bd, eq.to("cuda:0"))
        bd.fts = ref_fts.to("cuda:0")                                # the reference's fp32 batch through the fused kernel
        out32 = m(bd, eq.to("cuda:0"))
    ref = O.video_path(P, c["h"], ref_fts, eq, b.query_mask, ref_mask, c["n_blocks"])
    for k in ("spatiotemporal_ft", "temporal_ft", "spatial_ft"):
        assert O.norm_err(out16[k].float().cpu(), out32[k].float().cpu()) < 1e-2, k     # same math, different kernels
        assert O.norm_err(out16[k].float().cpu(), ref[k]) < 2e-2, k


@pytest.mark.gpu
def test_back_to_back_collates_keep_their_own_masks():
    """Eight ragged batches collated back to back WITHOUT synchronising in between, behind a long-running kernel on the same
    stream (so the mask copies are still queued when the next collate rewrites a staging buffer): every batch keeps its own
    temporal mask (the pinned staging ring is guarded by CUDA events)."""
    import numpy as np
    import torch
    from bist_b200.ingest import FeatureCache
    rs = np.random.RandomState(3)
    S, D, T, B, n_vid = 4, 64, 12, 6, 20
    lens = rs.randint(3, T + 1, size=n_vid)
    cache = FeatureCache(D, S, int(lens.sum()), dtype=torch.float32)
    for v in range(n_vid):
        cache.add(v, rs.uniform(0.1, 1, size=(int(lens[v]), S, D)).astype(np.float32))
    dev = torch.device("cuda:0")
    picks = [[int(v) for v in rs.permutation(n_vid)[:B]] for _ in range(8)]
    torch.cuda.synchronize()
    torch.cuda._sleep(200_000_000)          # ~0.1 s of GPU spin: everything below queues up behind it
    outs = [cache.collate(vids, dev, T=T) for vids in picks]
    torch.cuda.synchronize()
    for vids, (fts, mask) in zip(picks, outs):
        want = torch.tensor([[t < lens[v] for t in range(T)] for v in vids])
        assert torch.equal(mask[:, 0].cpu(), want)
        assert torch.equal((fts.float().sum((2, 3)) != 0).cpu(), want)
