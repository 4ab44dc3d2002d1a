"""Diagnostics: device time of the torch-side ops of one train iteration (keypoint MLP etc.) replayed as a graph."""
import sys, torch
sys.path.insert(0, ".")
from gaussianprediction_b200.deform import DeformableField, positional_encoding
dev = "cuda"
Kp = 512
net = DeformableField(104, 7, 4, 256).to(dev)
kp = torch.randn(Kp, 3, device=dev, requires_grad=True); kf = torch.randn(Kp, 32, device=dev, requires_grad=True)
t = torch.tensor([0.3], device=dev)
g_dx = torch.randn(Kp, 3, device=dev); g_dq = torch.randn(Kp, 4, device=dev)
params = [kp, kf] + list(net.parameters())
def body():
    for p in params: p.grad = None
    te = positional_encoding(t, 6).unsqueeze(0)
    noise = torch.randn_like(kp) * 0.0
    xe = positional_encoding(kp + noise, 10)
    tt = te.repeat((Kp, 1))
    d = net(torch.cat([kf, xe, tt], dim=-1))
    dx, dq = d[..., 0:3], torch.nn.functional.normalize(d[..., 3:7])
    torch.autograd.backward([dx, dq], [g_dx, g_dq])
s = torch.cuda.Stream(); s.wait_stream(torch.cuda.current_stream())
with torch.cuda.stream(s):
    for _ in range(3): body()
torch.cuda.current_stream().wait_stream(s); torch.cuda.synchronize()
for p in params: p.grad = None
g = torch.cuda.CUDAGraph()
with torch.cuda.graph(g): body()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(50): g.replay()
e1.record(); torch.cuda.synchronize()
print("keypoint MLP (PE, cat, 5 Linear, normalize) fwd+bwd as a graph: %.4f ms" % (e0.elapsed_time(e1) / 50))
from gaussianprediction_b200.deform import keypoint_mlp
def body_fused():
    for p in params: p.grad = None
    noise = torch.randn_like(kp) * 0.0
    dx, dq, _ = keypoint_mlp(net, kp + noise, kf, t, 10, 6, normalize_q=True)
    torch.autograd.backward([dx, dq], [g_dx, g_dq])
s.wait_stream(torch.cuda.current_stream())
with torch.cuda.stream(s):
    for _ in range(3): body_fused()
torch.cuda.current_stream().wait_stream(s); torch.cuda.synchronize()
for p in params: p.grad = None
gf = torch.cuda.CUDAGraph()
with torch.cuda.graph(gf): body_fused()
torch.cuda.synchronize()
e0.record()
for _ in range(50): gf.replay()
e1.record(); torch.cuda.synchronize()
print("same through gp_kpmlp_* as a graph: %.4f ms" % (e0.elapsed_time(e1) / 50))
# image-sized torch ops of the step: SH cat + probe loss
N = 100000
dc = torch.randn(N, 1, 3, device=dev, requires_grad=True); rest = torch.randn(N, 15, 3, device=dev, requires_grad=True)
img = torch.randn(3, 800, 800, device=dev, requires_grad=True); G = torch.randn(3, 800, 800, device=dev)
gsh = torch.randn(N, 16, 3, device=dev)
def body2():
    dc.grad = None; rest.grad = None; img.grad = None
    sh = torch.cat((dc, rest), dim=1)
    sh.backward(gsh)
    (img * G).sum().backward()
s.wait_stream(torch.cuda.current_stream())
with torch.cuda.stream(s):
    for _ in range(3): body2()
torch.cuda.current_stream().wait_stream(s); torch.cuda.synchronize()
dc.grad = None; rest.grad = None; img.grad = None
g2 = torch.cuda.CUDAGraph()
with torch.cuda.graph(g2): body2()
torch.cuda.synchronize()
e0.record()
for _ in range(50): g2.replay()
e1.record(); torch.cuda.synchronize()
print("SH cat fwd+bwd + probe loss fwd+bwd as a graph: %.4f ms" % (e0.elapsed_time(e1) / 50))
