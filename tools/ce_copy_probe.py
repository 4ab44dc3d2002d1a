"""Copy-engine peer copies over NVLink (torchrun, N >= 2; diagnostics): bandwidth of pulling a peer's symmetric buffer into
local memory with cudaMemcpyAsync, alone and while this library's neighbour search runs on the SMs.

    python -m torch.distributed.run --nproc-per-node 2 tools/ce_copy_probe.py [MiB]
"""
import os
import sys

import torch
import torch.distributed as dist
import torch.distributed._symmetric_memory as symm

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gaussianprediction_b200 import synthetic as syn  # noqa: E402
from gaussianprediction_b200.deform import knn_keypoints, spatial_order  # noqa: E402

rank, lr, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
mb = int(sys.argv[1]) if len(sys.argv) > 1 else 544
n = mb * 1024 * 1024 // 4
buf = symm.empty(n, dtype=torch.float32, device=dev)
buf.fill_(float(rank + 1))
hdl = symm.rendezvous(buf, dist.group.WORLD)
torch.cuda.synchronize()
dist.barrier()
peer = (rank + 1) % world
src = hdl.get_buffer(peer, (n,), torch.float32)
dst = torch.empty(n, device=dev)
side = torch.cuda.Stream(device=dev)


def timed(fn, reps=5):
    fn()
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def copy_main():
    dst.copy_(src, non_blocking=True)


t_copy = timed(copy_main)
ok = bool((dst == float(peer + 1)).all())

# the neighbour search of the C5 job (2 M x 512, 35-D) alone, and with the copy in flight on a side stream
N, Kp = 2_000_000, 512
gs = syn.make_gaussians(N, seed=1)
g = torch.Generator().manual_seed(0)
kp = gs.xyz[torch.randperm(N, generator=g)[:Kp]].contiguous().to(dev)
kf = gs.motion_feature[:Kp].contiguous().to(dev)
xyz, mf = gs.xyz.to(dev), gs.motion_feature.to(dev)
order = spatial_order(xyz)
hint = None
_, hint, _ = knn_keypoints(xyz, mf, kp, kf, 6, "hybird", 0.5, idx32_only=True, order=order)


def knn():
    knn_keypoints(xyz, mf, kp, kf, 6, "hybird", 0.5, idx32_only=True, order=order, hint=hint)


t_knn = timed(knn)


def both():
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        dst.copy_(src, non_blocking=True)
    knn()
    torch.cuda.current_stream().wait_stream(side)


t_both = timed(both)
if rank == 0:
    print(f"world {world}: peer copy of {mb} MiB by cudaMemcpyAsync: {t_copy:.3f} ms = {mb / 1024 / (t_copy / 1e3):.0f} GiB/s "
          f"(correct: {ok}); neighbour search alone {t_knn:.3f} ms; search + copy together {t_both:.3f} ms")
dist.barrier()
dist.destroy_process_group()
