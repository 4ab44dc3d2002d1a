"""Times a flat all-reduce of the gradient volume of one train step (torchrun; diagnostics only)."""
import os, sys, torch, torch.distributed as dist
rank = int(os.environ["RANK"]); lr = int(os.environ["LOCAL_RANK"]); world = int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(lr); dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
for mb in (1, 8, 24, 64):
    n = mb * 1024 * 1024 // 4
    x = torch.ones(n, device=dev)
    for _ in range(5): dist.all_reduce(x)
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20): dist.all_reduce(x)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 20
    if rank == 0:
        print(f"world {world} allreduce {mb} MiB: {ms*1e3:.1f} us, busbw {2*(world-1)/world*mb/1024/ (ms/1e3):.1f} GiB/s", flush=True)
# many small tensors, grouped
ts = [torch.ones(s, device=dev) for s in (300000, 300000, 4500000, 100000, 300000, 400000, 1200000, 1536, 16384, 226055)]
from torch.distributed.distributed_c10d import _coalescing_manager
for _ in range(3):
    with _coalescing_manager(device=dev):
        for t in ts: dist.all_reduce(t)
torch.cuda.synchronize(); dist.barrier()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20):
    with _coalescing_manager(device=dev):
        for t in ts: dist.all_reduce(t)
e1.record(); torch.cuda.synchronize()
if rank == 0: print(f"grouped 10 tensors ({sum(t.numel() for t in ts)*4/2**20:.1f} MiB): {e0.elapsed_time(e1)/20*1e3:.1f} us")
dist.destroy_process_group()
