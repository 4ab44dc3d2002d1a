"""GPU tests of the hash-grid blend-weight network (csrc/hashgrid.cu + csrc/linear_tc.cu) against the CPU restatement of
the published algorithm (oracle/hashgrid_oracle.py; parity unpinned: tinycudann is not in the reference tree).
Tolerances: encoding 1e-6 absolute of its scale, logits 1e-4 relative, gradients 1e-3 relative."""
import numpy as np
import pytest
import torch

from helpers import rel_err

pytestmark = pytest.mark.gpu


def _net(seed=0, **kw):
    from gaussianprediction_b200.hashgrid import HashGridWeights
    torch.manual_seed(seed)
    net = HashGridWeights(**kw)
    with torch.no_grad():
        net.table.uniform_(-0.5, 0.5)           # larger than the training init so that errors are visible
    return net


@pytest.mark.parametrize("log2_t,N", [(19, 4000), (12, 3000)])       # 12: every level above the third is hashed
def test_hashgrid_network_vs_oracle(gplib, log2_t, N):
    from gaussianprediction_b200.deform import spatial_order
    from oracle import hashgrid_oracle as ho
    net = _net(log2_hashmap_size=log2_t)
    g = torch.Generator().manual_seed(3)
    xyz = torch.rand(N, 3, generator=g) * 1.6 - 0.3            # also outside [0,1] and negative (cells wrap like tcnn's cast)
    cot = torch.randn(N, 12, generator=g)
    kw = dict(n_levels=net.cfg[0], base_resolution=net.cfg[1], per_level_scale=net.cfg[2], log2_hashmap_size=net.cfg[3])
    table = net.table.detach().clone().requires_grad_(True)
    ws = [w.detach().clone().requires_grad_(True) for w in net.weights]
    want_enc = ho.encode(xyz, table, **kw)
    want = ho.network(xyz, table, *ws, **kw)
    (want * cot).sum().backward()
    net = net.cuda()
    x = xyz.cuda()
    from gaussianprediction_b200.hashgrid import _HashEncode
    for order in (None, spatial_order(x)):
        net.order = order
        for p in net.parameters():
            p.grad = None
        enc = _HashEncode.apply(x, net.table, net.cfg, order)
        e_enc = float((enc.cpu() - want_enc.detach()).abs().max())
        print("encoding max abs err", e_enc)
        assert e_enc <= 2e-6 * max(1.0, float(want_enc.abs().max()))
        out = net(x)
        assert rel_err(out, want.detach()) <= 1e-4
        (out * cot.cuda()).sum().backward()
        assert rel_err(net.table.grad, table.grad) <= 1e-3
        for a, b in zip(net.weights, ws):
            assert rel_err(a.grad, b.grad) <= 1e-3
    # frozen logits (evaluation): the network is evaluated once
    frozen = net.freeze(x)
    assert net(x) is frozen


def test_hashgrid_as_weights_model_in_the_keypoint_stage(gplib):
    """The network in place of the logits table: forward + backward through the whole deformation."""
    from gaussianprediction_b200 import synthetic as syn
    from gaussianprediction_b200.deform import KeypointDeformation, default_args
    from gaussianprediction_b200.hashgrid import HashGridWeights
    gs = syn.make_gaussians(5000, seed=9)
    kp, kf = syn.make_keypoints(gs, Kp=64, seed=10)
    torch.manual_seed(1)
    wm = HashGridWeights(n_out=12)
    model = KeypointDeformation(default_args(feature_amplify=0.5, norm_rotation=True), gs.xyz, gs.rotation, gs.scaling, gs.opacity,
                                gs.motion_feature, kp, kf, weights_model=wm).cuda()
    xyz, q, s, o = model(torch.tensor([0.3], device="cuda"), 50_000)
    (xyz.sum() + q.square().sum()).backward()
    assert torch.isfinite(wm.table.grad).all() and float(wm.table.grad.abs().sum()) > 0
    assert all(torch.isfinite(w.grad).all() and float(w.grad.abs().sum()) > 0 for w in wm.weights)
