"""aggregate device->host copy rate of N ranks copying a C5 batch's colour frames (629 MB) into pinned memory at the same
time: the floor of the e2e leg at N GPUs, with nothing of the renderer involved (torch only)"""
import os, time, json, torch
import torch.distributed as dist
rank = int(os.environ.get("RANK", 0)); world = int(os.environ.get("WORLD_SIZE", 1)); local = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
n = 4096 * 320 * 240
d = torch.zeros(n, dtype=torch.int16, device="cuda")
h = torch.empty(n, dtype=torch.int16).pin_memory()
for _ in range(2):
    h.copy_(d, non_blocking=True)
torch.cuda.synchronize()
if world > 1:
    dist.barrier()
t0 = time.perf_counter()
for _ in range(10):
    h.copy_(d, non_blocking=True)
torch.cuda.synchronize()
dt = (time.perf_counter() - t0) / 10
t = torch.tensor([dt], dtype=torch.float64, device="cuda")
if world > 1:
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
if rank == 0:
    print(json.dumps({"n_gpus": world, "ms_per_copy_max_over_ranks": t.item() * 1e3, "per_gpu_GBps": n * 2 / t.item() / 1e9, "aggregate_GBps": world * n * 2 / t.item() / 1e9}))
if world > 1:
    dist.destroy_process_group()
