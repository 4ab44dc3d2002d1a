"""Host-side multi-rank logic on CPU: gloo backend, world_size 2 (the GPU path uses the same code over NCCL)."""
import os

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from gaussianprediction_b200.parallel import FlatGradAllReduce, reduce_densification_stats, shard_samples
    torch.manual_seed(0)                                    # replicated parameters
    params = [torch.nn.Parameter(torch.randn(5, 3)), torch.nn.Parameter(torch.randn(7)), torch.nn.Parameter(torch.randn(2, 2))]
    n_samples = 8
    mine = shard_samples(n_samples, rank, world)
    # per-sample "loss": deterministic function of the sample index; losses of a step are summed (train.py:113-133)
    loss = sum(((s + 1) * (params[0] ** 2).sum() + (s + 2) * params[1].sum()) for s in mine)
    loss.backward()                                         # params[2] gets no gradient on any rank
    red = FlatGradAllReduce(params)
    red.reduce()
    # the grouped variant falls back to the flat path on gloo and must agree with it
    from gaussianprediction_b200.parallel import GroupedGradAllReduce
    params2 = [torch.nn.Parameter(p.detach().clone()) for p in params]
    loss2 = sum(((s + 1) * (params2[0] ** 2).sum() + (s + 2) * params2[1].sum()) for s in mine)
    loss2.backward()
    GroupedGradAllReduce(params2).reduce()
    assert torch.allclose(params2[0].grad, params[0].grad) and torch.allclose(params2[1].grad, params[1].grad)
    radii = torch.tensor([rank + 1, 5 - rank, 0], dtype=torch.int32)
    vis = torch.tensor([rank == 0, rank == 1, False])
    radii, vis = reduce_densification_stats(radii, vis)
    if rank == 0:
        torch.save(dict(g0=params[0].grad.clone(), g1=params[1].grad.clone(), g2=params[2].grad.clone(), mine=mine,
                        radii=radii, vis=vis, same_storage=params[0].grad.data_ptr() == red.views[0].data_ptr(),
                        nbytes=red.nbytes), out)
    dist.barrier()
    dist.destroy_process_group()


def test_sample_sharding_and_flat_allreduce_gloo(tmp_path):
    from gaussianprediction_b200.parallel import shard_samples
    assert shard_samples(8, 1, 2) == [1, 3, 5, 7] and shard_samples(5, 3, 4) == [3] and shard_samples(3, 3, 4) == []
    allidx = sorted(sum((shard_samples(10, r, 4) for r in range(4)), []))
    assert allidx == list(range(10))                        # every sample rendered exactly once
    out = str(tmp_path / "r0.pt")
    port = 29500 + (os.getpid() % 2000)
    mp.spawn(_worker, args=(2, port, out), nprocs=2, join=True)
    r = torch.load(out)
    torch.manual_seed(0)
    p0, p1 = torch.randn(5, 3), torch.randn(7)
    w0 = sum(s + 1 for s in range(8))
    w1 = sum(s + 2 for s in range(8))
    assert torch.allclose(r["g0"], 2 * w0 * p0) and torch.allclose(r["g1"], torch.full((7,), float(w1)))
    assert float(r["g2"].abs().sum()) == 0 and r["mine"] == [0, 2, 4, 6] and r["same_storage"] and r["nbytes"] == 28 * 4   # 15 -> 16, 7 -> 8, 4: views are 16-byte aligned
    assert r["radii"].tolist() == [2, 5, 0] and r["vis"].tolist() == [True, True, False]


def _bucket_worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from gaussianprediction_b200._sinks import grad_sinks, out_like, sink_for
    from gaussianprediction_b200.parallel import FlatGradAllReduce
    torch.manual_seed(0)
    a, b, c = (torch.nn.Parameter(torch.randn(4, 3)), torch.nn.Parameter(torch.randn(6)), torch.nn.Parameter(torch.randn(2, 5)))
    red = FlatGradAllReduce([a, b, c], first=[c])           # c's gradient is "ready early": it leads the buffer
    early, late = red.buckets()
    layout = (red.params[0] is c, red.bucket_ranges[0][1], early.numel(), late.numel())
    # a producer that looks its destination up (what the autograd Functions do in their forward) writes in place
    with grad_sinks(zip(red.params, red.views)):
        dst = sink_for(c)
        hit = dst is not None and dst.data_ptr() == red.views[0].data_ptr() and sink_for(torch.zeros(3)) is None
        g = out_like(dst, (2, 5), dtype=torch.float32)
    g.copy_(torch.full((2, 5), float(rank + 1)))
    c.grad = g
    a.grad = torch.full((4, 3), 10.0 * (rank + 1))          # arrives in a tensor of its own: packed by copy; b: no gradient
    moved_early = red.pack(0, red.bucket_ranges[0][1])
    w1 = dist.all_reduce(early, async_op=True)              # the early bucket goes first ...
    moved_late = red.pack(red.bucket_ranges[0][1], None)
    w2 = dist.all_reduce(late, async_op=True)               # ... the rest follows (pipeline.TrainStep._replay)
    w1.wait(); w2.wait()
    if rank == 0:
        torch.save(dict(layout=layout, hit=hit, moved=(moved_early, moved_late), c=c.grad.clone(), a=a.grad.clone(),
                        b=b.grad.clone(), in_flat=all(p.grad.data_ptr() == v.data_ptr() for p, v in zip(red.params, red.views))),
                   out)
    dist.barrier()
    dist.destroy_process_group()


def test_bucketed_exchange_and_grad_sinks_gloo(tmp_path):
    out = str(tmp_path / "b0.pt")
    port = 31500 + (os.getpid() % 2000)
    mp.spawn(_bucket_worker, args=(2, port, out), nprocs=2, join=True)
    r = torch.load(out)
    assert r["layout"] == (True, 1, 12, 20)        # 10 -> 12 | 12 + 6 -> 20: every view starts 16-byte aligned
    assert r["hit"] and r["hit"] and r["moved"] == (0, 2) and r["in_flat"]
    assert torch.equal(r["c"], torch.full((2, 5), 3.0)) and torch.equal(r["a"], torch.full((4, 3), 30.0))
    assert float(r["b"].abs().sum()) == 0


def test_flat_views_are_16_byte_aligned_for_any_gaussian_count():
    """The backward kernels store 128-bit vectors straight into the reducer's views (pipeline.TrainStep); after
    densification the Gaussian count is rarely a multiple of 4 (N = 6001 made `_rotation` start mid-vector)."""
    from gaussianprediction_b200.parallel import FlatGradAllReduce
    N = 6001
    ps = [torch.nn.Parameter(torch.zeros(*shape)) for shape in ((N, 1, 3), (N, 15, 3), (N, 3), (N, 4), (N, 3), (N, 1),
                                                                (N, 12), (37, 3), (37, 32), (7,), (7, 256))]
    red = FlatGradAllReduce(ps, first=ps[:2], second=ps[2:7])
    base = red.flat.data_ptr()
    assert all((v.data_ptr() - base) % 16 == 0 for v in red.views)
    assert all(v.shape == p.shape for v, p in zip(red.views, red.params))
    r = red.bucket_ranges
    assert r[0][2] == 0 and r[-1][3] == red.flat.numel() and all(a[3] == b[2] for a, b in zip(r, r[1:]))
    assert all(e0 % 4 == 0 for _, _, e0, _ in r)
    # views do not overlap
    ends = sorted(((v.data_ptr() - base) // 4, (v.data_ptr() - base) // 4 + v.numel()) for v in red.views)
    assert all(a[1] <= b[0] for a, b in zip(ends, ends[1:]))


def test_accumulating_sinks_are_a_separate_nested_map():
    """``grad_sinks(pairs, accumulate=...)``: plain and accumulating destinations are looked up separately (only kernels
    that can add in place ask for the second kind), nested scopes extend the maps and restore them on exit."""
    from gaussianprediction_b200._sinks import acc_sink_for, grad_sinks, sink_for
    a, b, c = torch.zeros(8, 3), torch.zeros(5), torch.zeros(4, 4)
    da, db, sb = torch.zeros(8, 3), torch.zeros(5), torch.zeros(5)
    assert sink_for(a) is None and acc_sink_for(a) is None
    with grad_sinks([(a, da), (b, db)], accumulate=[(b, sb)]):
        assert sink_for(a) is da and sink_for(b) is db and sink_for(c) is None
        assert acc_sink_for(b) is sb and acc_sink_for(a) is None
        assert sink_for(a.view(-1)) is da                       # matched by storage, not by tensor identity
        assert sink_for(torch.zeros(0)) is None and acc_sink_for(None) is None
        dc = torch.zeros(4, 4)
        with grad_sinks([], accumulate=[(c, dc)]):
            assert acc_sink_for(c) is dc and acc_sink_for(b) is sb and sink_for(a) is da
        assert acc_sink_for(c) is None
        with pytest.raises(AssertionError):
            with grad_sinks([], accumulate=[(a, torch.zeros(3, 8))]):   # shape mismatch is refused
                pass
        assert acc_sink_for(b) is sb                            # ... and leaves the maps as they were
    assert sink_for(a) is None and acc_sink_for(b) is None
