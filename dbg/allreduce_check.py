import os, sys, torch, torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local); dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
from bist_b200.core import full_model, full_batch
from bist_b200.dist import allreduce_grads
shapes = dict(B=4, T=5, S=4, D=64, d=64, h=4, Lq=6, Lh=8, Lc=7, Lt=5, V=128)
m, lc = full_model("fp32", shapes)
m = m.to(dev).eval()
for p in m.parameters():
    dist.broadcast(p.data, 0)
b = full_batch(10 + rank, dev, shapes)
lc.loss(m(b), b)[0].backward()
params = [p for p in m.parameters() if p.grad is not None]
local_g = [p.grad.clone() for p in params]
n = allreduce_grads(list(m.parameters()))
worst = 0.0
for p, g in zip(params, local_g):
    s = g.clone(); dist.all_reduce(s); s /= world
    worst = max(worst, float((p.grad - s).abs().max() / s.abs().max().clamp_min(1e-20)))
if rank == 0:
    print("collectives", n, "params", len(params), "worst rel err vs mean of ranks", worst)
dist.destroy_process_group()
