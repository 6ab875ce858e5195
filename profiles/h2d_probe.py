"""How much pinned host -> device bandwidth does the box deliver when every GPU copies at once?  (bench.py's e2e leg
moves one 37 MB text batch per GPU per step.)  torchrun --nproc-per-node G profiles/h2d_probe.py"""
import json
import os

import torch
import torch.distributed as dist

rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
dev = torch.device("cuda", torch.cuda.current_device())
dist.init_process_group("nccl", device_id=dev)
n = 37_290_008
host = [torch.empty(n, dtype=torch.uint8).pin_memory() for _ in range(4)]
dst = [torch.empty(n, dtype=torch.uint8, device=dev) for _ in range(2)]
tok = torch.zeros(1, device=dev)


def run(active, reps=60):
    dist.all_reduce(tok)
    torch.cuda.synchronize()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    if active:
        for i in range(reps):
            dst[i % 2].copy_(host[i % 4], non_blocking=True)
    t1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([t0.elapsed_time(t1)], device=dev)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return reps * n / (float(ms) * 1e-3) / 1e9


run(True, 10)
alone = run(rank == 0)
together = run(True)
if rank == 0:
    print(json.dumps({"gpus": world, "bytes_per_copy": n, "one_gpu_copying_GBps": round(alone, 1),
                      "all_gpus_copying_GBps_per_gpu": round(together, 1),
                      "all_gpus_copying_GBps_total": round(together * world, 1)}))
dist.barrier()
dist.destroy_process_group()
