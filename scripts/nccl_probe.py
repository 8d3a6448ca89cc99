"""All-reduce bandwidth of the box's NCCL path at the gradient-block sizes of the training step.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 scripts/nccl_probe.py

Prints, from rank 0, one line per size: MB, ms per all-reduce, algorithmic GB/s (bytes / time) and bus GB/s
(2(n-1)/n x that).  Used to tell the cost of the gradient exchange apart from the cost of the backward kernels.
"""
import os

import torch
import torch.distributed as dist


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    if rank == 0:
        print("p2p 0->1:", torch.cuda.can_device_access_peer(0, 1) if torch.cuda.device_count() > 1 else None, flush=True)
    for mb in (1, 4, 16, 42, 130, 256):
        x = torch.ones(mb * (1 << 20) // 4, device=dev)
        for _ in range(5):
            dist.all_reduce(x)
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n = 20
        e0.record()
        for _ in range(n):
            dist.all_reduce(x)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / n
        if rank == 0:
            gbs = mb * (1 << 20) / (ms * 1e-3) / 1e9
            print("%4d MB  %.3f ms  alg %.1f GB/s  bus %.1f GB/s" % (mb, ms, gbs, gbs * 2 * (world - 1) / world), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
