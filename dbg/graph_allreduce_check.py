"""2+ GPUs: gradients left by GraphedTrainStep(allreduce=True) (NCCL captured inside the training graph, stream branches on)
equal the mean over ranks of the eager local gradients.  torchrun --nproc-per-node 2 dbg/graph_allreduce_check.py"""
import os, sys, torch, torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local); dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
from bist_b200.core import full_model, full_batch
from bist_b200.graph import GraphedTrainStep
shapes = dict(B=4, T=5, S=4, D=64, d=64, h=4, Lq=6, Lh=8, Lc=7, Lt=5, V=128, Ta=6, Da=16)
for precision in ("fp32", "bf16"):
    m, lc = full_model(precision, shapes, audio=True)
    m = m.to(dev).eval()                                   # no dropout: the gradients are deterministic
    for p in m.parameters():
        dist.broadcast(p.data, 0)
    b = full_batch(10 + rank, dev, shapes, audio=True)
    loss = lc.loss(m(b), b)[0]
    loss.backward()
    params = [p for p in m.parameters() if p.grad is not None]
    mean_g = []
    for p in params:
        s = p.grad.clone(); dist.all_reduce(s); mean_g.append(s / world)
    del loss
    for p in m.parameters():
        p.grad = None
    print(f"[rank {rank}] {precision}: eager mean gradients done, capturing", flush=True)
    step = GraphedTrainStep(lambda: lc.loss(m(b), b)[0], list(m.parameters()), warmup=2, allreduce=True)
    print(f"[rank {rank}] {precision}: captured, {step.collectives} collectives", flush=True)
    worst = 0.0
    for _ in range(3):
        step()
        torch.cuda.synchronize()
        for p, s in zip(params, mean_g):
            worst = max(worst, float((p.grad - s).abs().max() / s.abs().max().clamp_min(1e-20)))
    step.close()
    del step                                               # graphs holding NCCL kernels must go before the process group
    print(f"[rank {rank}] {precision}: worst {worst:.3e}", flush=True)
    if rank == 0:
        print(precision, "collectives per replay", step.collectives, "params", len(params), "worst rel err vs mean of ranks", worst)
import gc; gc.collect(); torch.cuda.synchronize()
os._exit(0)                                                # destroy_process_group() after NCCL graph capture hangs at exit
