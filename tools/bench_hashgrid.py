"""Hash-grid blend-weight network (16 levels x 4 features, T = 2^19, MLP 64 x 2 -> 12): forward / forward + backward ms."""
import os, sys, json
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gaussianprediction_b200 import synthetic as syn
from gaussianprediction_b200.deform import spatial_order
from gaussianprediction_b200.hashgrid import HashGridWeights, _HashEncode

def timeit(fn, reps=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps

for N in [int(a) for a in sys.argv[1:]] or [100_000, 1_000_000, 2_000_000]:
    torch.manual_seed(0)
    net = HashGridWeights(12).cuda()
    xyz = ((syn.make_gaussians(N, seed=1).xyz + 1.3) / 2.6).cuda()       # the cloud scaled into [0,1]^3
    cot = torch.randn(N, 12, device="cuda")
    r = dict(N=N)
    for name, order in (("index_order", None), ("morton_order", spatial_order(xyz))):
        net.order = order
        def enc_f():
            with torch.no_grad(): _HashEncode.apply(xyz, net.table, net.cfg, order)
        def fwd():
            with torch.no_grad(): net(xyz)
        def fb():
            for p in net.parameters(): p.grad = None
            (net(xyz) * cot).sum().backward()
        r[name] = dict(encode_ms=round(timeit(enc_f), 4), forward_ms=round(timeit(fwd), 4), fwd_bwd_ms=round(timeit(fb), 4))
    print(json.dumps(r))

# per-kernel split of one forward + backward at the last N (library kernels only; torch.mm weight gradients are the rest)
import ctypes
from gaussianprediction_b200 import _lib
L = _lib.lib()
for name, order in (("index_order", None), ("morton_order", spatial_order(xyz))):
    net.order = order
    L.gp_profile_enable(1)
    for _ in range(3):
        for p in net.parameters(): p.grad = None
        (net(xyz) * cot).sum().backward()
    torch.cuda.synchronize()
    buf = ctypes.create_string_buffer(1 << 14)
    L.gp_profile_report(buf, len(buf)); L.gp_profile_enable(0)
    print(name, {l.split()[0]: round(float(l.split()[2]) / 3, 3) for l in buf.value.decode().splitlines()})
