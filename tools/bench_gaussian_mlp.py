"""Per-Gaussian Deformable_Field ([N,104] -> 4 x 256 -> 8) forward + backward: tensor-core path vs the torch modules
(cuBLAS fp32), CUDA-event times.  python tools/bench_gaussian_mlp.py [N ...]"""
import os, sys, json
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gaussianprediction_b200.deform import DeformableField, gaussian_mlp

def timeit(fn, reps=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps

for N in [int(a) for a in sys.argv[1:]] or [100_000, 1_000_000]:
    torch.manual_seed(0)
    net = DeformableField(104, 8, 4, 256).cuda()
    x = torch.randn(N, 104, device="cuda", requires_grad=True)
    cot = torch.randn(N, 8, device="cuda")
    def ours_f():
        with torch.no_grad(): gaussian_mlp(net, x)
    def torch_f():
        with torch.no_grad(): net(x)
    def ours_fb():
        for p in net.parameters(): p.grad = None
        x.grad = None
        (gaussian_mlp(net, x) * cot).sum().backward()
    def torch_fb():
        for p in net.parameters(): p.grad = None
        x.grad = None
        (net(x) * cot).sum().backward()
    flop = 2.0 * N * (104 * 256 + 3 * 256 * 256 + 256 * 8)
    r = dict(N=N, ours_fwd_ms=timeit(ours_f), torch_fwd_ms=timeit(torch_f), ours_fwd_bwd_ms=timeit(ours_fb), torch_fwd_bwd_ms=timeit(torch_fb))
    r["ours_fwd_tflops_fp32_equivalent"] = flop / r["ours_fwd_ms"] / 1e9
    r["ours_fwd_tflops_tf32_issued"] = 3 * flop / r["ours_fwd_ms"] / 1e9
    print(json.dumps({k: (round(v, 4) if isinstance(v, float) else v) for k, v in r.items()}))
