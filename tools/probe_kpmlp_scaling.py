"""Diagnostics: device time of the fused keypoint MLP kernels vs the number of keypoints (blocks = Kp/4)."""
import sys, torch
sys.path.insert(0, ".")
from gaussianprediction_b200 import _lib
from gaussianprediction_b200.deform import DeformableField, keypoint_mlp
dev = "cuda"
L = _lib.lib()
net = DeformableField(104, 7, 4, 256).to(dev)
t = torch.tensor([0.3], device=dev)
for Kp in (64, 128, 256, 512, 1024, 2048):
    kp = torch.randn(Kp, 3, device=dev, requires_grad=True); kf = torch.randn(Kp, 32, device=dev, requires_grad=True)
    g = torch.randn(Kp, 7, device=dev)
    def body():
        for p in list(net.parameters()) + [kp, kf]: p.grad = None
        torch.autograd.backward(keypoint_mlp(net, kp, kf, t, 10, 6, normalize_q=True)[:2], [g[:, :3].contiguous(), g[:, 3:].contiguous()])
    for _ in range(3): body()
    torch.cuda.synchronize()
    L.gp_profile_enable(1)
    for _ in range(10): body()
    torch.cuda.synchronize()
    buf = (__import__("ctypes").c_char * 4096)()
    n = L.gp_profile_report(buf, 4096)
    L.gp_profile_enable(0)
    print(Kp, buf.value.decode().replace("\n", " | ")[:300])
