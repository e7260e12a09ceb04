"""Concurrent host->device copies on N ranks (torchrun): is the e2e input bound by the shared PCIe uplinks?
Every rank copies its 154 MB of pinned fp32 images (the bench's per-step input) to its own GPU, all ranks at
once (barrier, CUDA events, max over ranks).  One JSON line on rank 0: GB/s per rank and in total."""
import json
import os

import torch
import torch.distributed as dist

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
res = {}
for name, n, dt in (("fp32_154MB", 256 * 3 * 224 * 224, torch.float32), ("uint8_38MB", 256 * 3 * 224 * 224, torch.uint8)):
    src = torch.empty(n, dtype=dt).pin_memory()
    dst = torch.empty(n, dtype=dt, device=dev)
    for _ in range(3):
        dst.copy_(src, non_blocking=True)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 20
    e0.record()
    for _ in range(reps):
        dst.copy_(src, non_blocking=True)
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1) / reps], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms = float(ms.item())
    gbs = n * src.element_size() / ms / 1e6
    res[name] = {"ms_per_copy_max_over_ranks": round(ms, 3), "GBps_per_rank": round(gbs, 1), "GBps_total": round(gbs * world, 1)}
if rank == 0:
    print(json.dumps({"h2d_concurrent": res, "ranks": world}), flush=True)
if world > 1:
    dist.destroy_process_group()
