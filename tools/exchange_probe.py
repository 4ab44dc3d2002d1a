"""Times gp_allreduce_sum_f32 (csrc/exchange.cu) against NCCL over a range of sizes / CTA counts (torchrun; diagnostics).

    python -m torch.distributed.run --nproc-per-node N tools/exchange_probe.py [MiB ...]
"""
import ctypes
import os
import sys

import torch
import torch.distributed as dist
import torch.distributed._symmetric_memory as symm

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gaussianprediction_b200 import _lib  # noqa: E402

rank, lr, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
L = _lib.lib()
sizes = [int(a) for a in sys.argv[1:]] or [8, 64, 512]
MAXB = 592
flag_bytes = int(L.gp_allreduce_flag_bytes(MAXB))
for mb in sizes:
    n = mb * 1024 * 1024 // 4
    buf = symm.empty(n + flag_bytes // 4, dtype=torch.float32, device=dev)
    buf.zero_()
    hdl = symm.rendezvous(buf, dist.group.WORLD)
    torch.cuda.synchronize()
    dist.barrier()
    peers = (ctypes.c_void_p * world)(*[int(p) for p in hdl.buffer_ptrs])
    mc = int(hdl.multicast_ptr or 0)
    x = torch.ones(n, device=dev)
    for _ in range(3):
        dist.all_reduce(x)
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        dist.all_reduce(x)
    e1.record(); torch.cuda.synchronize()
    t_nccl = e0.elapsed_time(e1) / 10
    if rank == 0:
        print(f"world {world} {mb} MiB: NCCL {t_nccl*1e3:.0f} us ({mb/1024/(t_nccl/1e3):.0f} GiB/s alg), multicast={'yes' if mc else 'no'}", flush=True)
    for use_mc in ([True, False] if mc else [False]):
        for blocks in (24, 48, 96, 148, 296, 592):
            # correctness once per config
            buf[:n].fill_(float(rank + 1))
            torch.cuda.synchronize(); dist.barrier()
            args = (rank, world, mc if use_mc else None, peers, 0, n, n * 4, blocks, torch.cuda.current_stream().cuda_stream)
            _lib.check(L.gp_allreduce_sum_f32(*args), "ar")
            torch.cuda.synchronize(); dist.barrier()
            want = float(world * (world + 1) // 2)
            ok = bool((buf[:n] == want).all())
            buf[:n].zero_()
            torch.cuda.synchronize(); dist.barrier()
            e0.record()
            for _ in range(10):
                _lib.check(L.gp_allreduce_sum_f32(*args), "ar")
            e1.record(); torch.cuda.synchronize()
            t = e0.elapsed_time(e1) / 10
            tt = torch.tensor([t], device=dev); dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            if rank == 0:
                print(f"   {'multimem' if use_mc else 'peer ld/st'} blocks {blocks:4d}: {float(tt)*1e3:.0f} us ({mb/1024/(float(tt)/1e3):.0f} GiB/s alg) ok={ok}", flush=True)
    del buf, hdl
dist.barrier()
dist.destroy_process_group()
