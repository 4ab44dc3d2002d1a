"""Timeline of two pipelined steps (torchrun, N >= 2): when do the stage graphs and the exchange kernels run?
GPU times from events on the main / exchange streams, host times from perf_counter (diagnostics only)."""
import os, sys, time
import torch, torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from gaussianprediction_b200 import pipeline
rank, lr, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(lr); dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
name = sys.argv[1] if len(sys.argv) > 1 else bench.HEADLINE
wl = dict(bench.WORKLOADS[name], name=name)
r = bench.OurRunner(wl, dev, 16, world, 8, exchange="nvls")
st = bench.workload_stats(r)
r.capacity = int(st["max_R"] * 1.25) + 65536
r.step(0)
r.enable_graph()
for s in range(4):
    r.step_graph(s)
torch.cuda.synchronize(); dist.barrier()
pipeline.TRACE = []
t0 = torch.cuda.Event(enable_timing=True); t0.record(); h0 = time.perf_counter()
for s in range(4, 6):
    pipeline._mark(f"--- step {s}")
    r.step_graph(s)
r.batch.flush()
pipeline._mark("flushed")
torch.cuda.synchronize()
if rank == 0:
    print(r.exchange_kind)
    for label, e, h in pipeline.TRACE:
        print(f"gpu {t0.elapsed_time(e):9.3f} ms  host {(h - h0) * 1e3:9.3f} ms  {label}")
dist.barrier(); dist.destroy_process_group()
