"""GPU parity of the step after the backward: gp_adam_step / gp_densification_stats against
torch.optim.Adam (the reference's optimizer, scene/gaussian_model.py:452) and the reference's index expressions."""
import numpy as np
import pytest
import torch

from helpers import rel_err

pytestmark = pytest.mark.gpu


def test_fused_adam_vs_torch_optim_adam(gplib):
    from gaussianprediction_b200.optim import FusedAdam
    torch.manual_seed(1)
    shapes = [(10007, 3), (10007, 1, 3), (10007, 15, 3), (10007, 1), (10007, 4), (256, 104), (256,), (7, 256), (7,), (5,)]
    lrs = [1.6e-4, 2.5e-3, 2.5e-3 / 20, 5e-2, 1e-3, 8e-4, 8e-4, 8e-4, 8e-4, 1e-3]
    ref = [torch.nn.Parameter(torch.randn(s, device="cuda")) for s in shapes]
    mine = [torch.nn.Parameter(p.detach().clone()) for p in ref]
    groups = lambda ps: [{"params": [p], "lr": lr, "name": "g%d" % i} for i, (p, lr) in enumerate(zip(ps, lrs))]
    o_ref = torch.optim.Adam(groups(ref), lr=0.0, eps=1e-15)
    o_mine = FusedAdam(groups(mine), lr=0.0, eps=1e-15)
    for step in range(1, 8):
        for i, (a, b) in enumerate(zip(ref, mine)):
            if i == 9 and step % 2 == 0:
                a.grad = b.grad = None                    # a parameter without gradient this step is skipped (own step count)
                continue
            g = torch.randn_like(a) * (10.0 ** (-(step % 4)))
            a.grad, b.grad = g.clone(), g.clone()
        if step == 4:                                     # the reference rewrites the group learning rates every iteration
            for og, mg in zip(o_ref.param_groups, o_mine.param_groups):
                og["lr"] = mg["lr"] = og["lr"] * 0.5
        o_ref.step()
        o_mine.step()
        o_ref.zero_grad(set_to_none=True)
        o_mine.zero_grad(set_to_none=True)
    for a, b in zip(ref, mine):
        assert float((a - b).abs().max()) <= 2e-6 * max(1.0, float(a.abs().max()))
        assert rel_err(o_mine.state[b]["exp_avg"], o_ref.state[a]["exp_avg"]) <= 1e-5
        assert rel_err(o_mine.state[b]["exp_avg_sq"], o_ref.state[a]["exp_avg_sq"]) <= 1e-5
        assert int(o_mine.state[b]["step"]) == int(o_ref.state[a]["step"])
    assert all(p.grad is None for p in mine)
    cpu_p = torch.nn.Parameter(torch.randn(3))
    cpu_p.grad = torch.randn(3)
    with pytest.raises(TypeError):
        FusedAdam([cpu_p], lr=1e-3).step()               # no CPU path


def test_adam_and_stats_vs_cpu_oracle(gplib):
    from gaussianprediction_b200.optim import FusedAdam, add_densification_stats
    from oracle import optim_oracle as oo
    g = torch.Generator().manual_seed(2)
    p0 = torch.randn(1003, 3, generator=g)
    p = torch.nn.Parameter(p0.clone().cuda())
    opt = FusedAdam([{"params": [p], "lr": 3e-3}], lr=0.0, eps=1e-15)
    o = dict(p=p0.numpy().copy(), m=np.zeros((1003, 3), np.float32), v=np.zeros((1003, 3), np.float32))
    for step in range(1, 4):
        gr = torch.randn(1003, 3, generator=g) * 1e-2
        p.grad = gr.cuda()
        opt.step()
        oo.adam_step(o["p"], gr.numpy(), o["m"], o["v"], step, 3e-3)
    assert np.abs(p.detach().cpu().numpy() - o["p"]).max() <= 2e-6
    # densification statistics
    P = 5000
    radii = ((torch.rand(P, generator=g) * 30).int() * (torch.rand(P, generator=g) > 0.3).int()).int()
    vgrad = torch.randn(P, 3, generator=g) * 1e-3
    st = [torch.rand(P, generator=g) * 20, torch.rand(P, 1, generator=g), torch.rand(P, 1, generator=g).round(),
          torch.rand(P, 1, generator=g) * 1e-3]
    want = [t.numpy().copy() for t in st]
    oo.densification_stats(vgrad.numpy(), radii.numpy(), *want)
    dev = [t.cuda() for t in st]
    add_densification_stats(vgrad.cuda(), radii.cuda(), dev[1], dev[2], xyz_gradient_accum_max=dev[3], max_radii2D=dev[0])
    for a, b in zip(dev, want):
        assert np.allclose(a.cpu().numpy(), b, rtol=1e-6, atol=0)
    with pytest.raises(TypeError):
        add_densification_stats(vgrad, radii, st[1], st[2])
