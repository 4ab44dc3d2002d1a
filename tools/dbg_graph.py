import sys, traceback
sys.path.insert(0, '/root/repo')
import torch, bench
wl = bench.WORKLOADS["tiny"]
dev = torch.device("cuda", 0)
r = bench.OurRunner(wl, dev, 4, 1)
st = bench.workload_stats(r)
r.capacity = int(st["max_R"] * 1.25) + 65536
for i in range(3):
    r.step(i)
torch.cuda.synchronize()
try:
    r.enable_graph()
    print("graph ok")
except Exception:
    traceback.print_exc()
