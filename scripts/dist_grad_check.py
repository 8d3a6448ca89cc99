"""N-GPU check (torchrun, NCCL): the row-sharded training step's gradients, averaged by the all-reduce, equal the gradients of
the single-GPU step on the CONCATENATED batch (reference normaliser /root/reference/model/optimize.py:46-50).

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 scripts/dist_grad_check.py

Every rank builds the same model (config-3 shapes at a reduced width so that the check is quick, audio stream on, fp32,
eval mode = dropout off) and the same global batch of 16 * world dialogue turns; rank r trains on rows [16r, 16r+16) with
dist.sync_token_counts + dist.allreduce_grads, eagerly AND through the captured graph with the all-reduce inside
(GraphedTrainStep(allreduce=True)); rank 0 also runs the whole batch alone.  Prints max normalised gradient error; exit 1 above 1e-4
(--wire-bf16: the gradients are exchanged as bf16, tolerance 1e-2)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from bist_b200 import synth
from bist_b200.core import PAD, full_args, full_model
from bist_b200.dist import allreduce_grads, overlap_grad_allreduce, set_grad_comm_dtype, sync_token_counts
from bist_b200.graph import GraphedTrainStep


def nerr(a, b):
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    wire_bf16 = "--wire-bf16" in sys.argv      # gradients cross NVLink as bf16 (dist.set_grad_comm_dtype): rounding of the exchanged values only
    set_grad_comm_dtype(torch.bfloat16 if wire_bf16 else torch.float32)
    tol = 1e-2 if wire_bf16 else 1e-4
    shapes = dict(B=16 * world, T=12, S=16, D=256, d=128, h=2, Lq=9, Lh=20, Lc=11, Lt=7, V=501, Ta=12, Da=128)
    args = full_args(shapes["d"], shapes["h"], n_blocks=2, dropout=0.0, audio=True)
    torch.manual_seed(0)
    model, lc = full_model("fp32", shapes, args=args, audio=True)
    model = model.to(dev).eval()      # no dropout anywhere (the attention dropout is fixed at 0.1 in train(), mtn.py:77); gradients still flow
    params = list(model.parameters())
    gb = synth.make_full_batch(3, shapes["B"], shapes["T"], shapes["S"], shapes["D"], shapes["Lq"], shapes["Lh"], shapes["Lc"],
                               shapes["Lt"], shapes["V"], Ta=shapes["Ta"], Da=shapes["Da"])
    lo, hi = 16 * rank, 16 * rank + 16
    sb = synth.batch_rows(gb, lo, hi)
    sb.ntokens, sb.qntokens = (sb.trg_y != PAD).sum(), (sb.query != PAD).sum()
    sb = sb.to(dev)

    def grads():
        return [p.grad.detach().clone() if p.grad is not None else None for p in params]

    # ---- sharded, eager --------------------------------------------------------------------------------------------------
    sync_token_counts(sb)
    for p in params:
        p.grad = None
    with overlap_grad_allreduce():
        lc.loss(model(sb), sb)[0].backward()
    ncoll = allreduce_grads(params)
    torch.cuda.synchronize()
    g_eager = grads()
    # ---- sharded, captured graph with the all-reduce inside ------------------------------------------------------------------
    for p in params:
        p.grad = None
    step = GraphedTrainStep(lambda: lc.loss(model(sb), sb)[0], params, warmup=2, allreduce=True)
    for _ in range(2):
        step()
    torch.cuda.synchronize()
    g_graph = grads()
    coll_graph = step.collectives
    step.close()
    del step
    # ---- the whole batch on one GPU (every rank computes it: same result, no communication) -----------------------------------
    for p in params:
        p.grad = None
    wb = gb.to(dev)
    lc.loss(model(wb), wb)[0].backward()
    torch.cuda.synchronize()
    g_ref = grads()
    scale = max(float(g.abs().max()) for g in g_ref if g is not None)
    worst_e = worst_g = 0.0
    for (n, _), a, b, r in zip(model.named_parameters(), g_eager, g_graph, g_ref):
        if r is None:
            continue
        if float(r.abs().max()) < 1e-6 * scale:          # zero by softmax shift invariance (key biases): rounding noise only
            continue
        worst_e, worst_g = max(worst_e, nerr(a, r)), max(worst_g, nerr(b, r))
    t = torch.tensor([worst_e, worst_g], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ok = float(t.max()) < tol
    if rank == 0:
        print("dist_grad_check: world %d, wire %s, %d collectives eager / %d in graph; max normalised gradient error vs the single-GPU step on the "
              "concatenated batch: eager %.3e, graph %.3e (tolerance %.0e) -> %s"
              % (world, "bf16" if wire_bf16 else "fp32", ncoll, coll_graph, float(t[0]), float(t[1]), tol, "OK" if ok else "FAIL"), flush=True)
    torch.cuda.synchronize()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
