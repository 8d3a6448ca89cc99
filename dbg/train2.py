import os, sys, time, torch, torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local); dev = torch.device("cuda", local)
if world > 1: dist.init_process_group("nccl", device_id=dev)
from bist_b200 import VideoPath, synth
from bist_b200.dist import allreduce_grads, overlap_grad_allreduce
import bench
c = bench.CFG
tm = VideoPath(c["d"], c["h"], c["n_blocks"], c["D"], dropout=0.1, precision="bf16").to(dev)
tm.load_state_dict(synth.make_params(0, c["d"], c["D"], c["n_blocks"]), strict=True); tm.train()
opt = torch.optim.Adam(tm.parameters(), lr=1e-4, fused=True)
db = synth.make_batch(1 + rank, c["B"], c["T"], c["S"], c["D"], c["Lq"]).to(dev)
deq = synth.make_encoded_query(1 + rank, c["B"], c["Lq"], c["d"]).to(dev)
tgt = torch.randn(c["B"], c["Lq"], c["d"], device=dev)
plist = list(tm.parameters())
def ev(): e = torch.cuda.Event(enable_timing=True); e.record(); return e
acc = [0.0] * 4; host = [0.0] * 4
for it in range(13):
    torch.cuda.synchronize(); 
    if world > 1: dist.barrier()
    torch.cuda.synchronize()
    t = [time.perf_counter()]; e = [ev()]
    opt.zero_grad(set_to_none=True)
    out = tm(db, deq)
    loss = ((out["temporal_ft"] - tgt) ** 2).mean() + ((out["spatial_ft"] - tgt) ** 2).mean()
    e.append(ev()); t.append(time.perf_counter())
    with overlap_grad_allreduce(): loss.backward()
    e.append(ev()); t.append(time.perf_counter())
    n = allreduce_grads(plist); e.append(ev()); t.append(time.perf_counter())
    opt.step(); e.append(ev()); t.append(time.perf_counter())
    torch.cuda.synchronize()
    if it >= 3:
        for i in range(4): acc[i] += e[i].elapsed_time(e[i + 1]) / 10; host[i] += (t[i + 1] - t[i]) * 100
if rank == 0:
    print("collectives", n, "views", sum(1 for p in plist if p.grad is not None and p.grad._is_view()), "of", len(plist))
    print("gpu ms  fwd %.3f bwd %.3f allreduce %.3f adam %.3f  total %.3f" % (*acc, sum(acc)))
    print("host ms fwd %.3f bwd %.3f allreduce %.3f adam %.3f  total %.3f" % (*host, sum(host)))
if world > 1: dist.destroy_process_group()
