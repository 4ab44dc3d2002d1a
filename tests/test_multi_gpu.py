"""Two-rank GPU test of the data-parallel step (skipped on a one-GPU box): ``bench.py --gpus 2`` on the tiny workload under
torchrun, with this library's exchange kernel and with NCCL.  bench.py itself asserts, outside the timed region, that
the exchanged gradient buffer equals the sum of the per-rank buffers formed independently (all_gather + add) and that
the radii equal the element-wise max over the ranks (``exchange_check``); the test reads that verdict."""
import json
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("exchange", ["nvls", "nccl"])
def test_two_rank_step_sums_gradients(gplib, exchange):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    port = 29600 + os.getpid() % 300
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(ROOT, "bench.py"), "--gpus", "2", "--workload", "tiny", "--steps", "3",
           "--warmup", "3", "--no-profile", "--exchange", exchange]
    out = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    line = json.loads([l for l in out.stdout.splitlines() if l.startswith("{")][-1])
    chk = line["exchange_check"]
    assert chk["ok"] and chk["radii_equal"] and max(chk["max_rel_err_per_bucket"]) <= 1e-4, chk
    assert line["n_gpus"] == 2 and line["config"]["samples_per_step"] == 8


def test_two_rank_step_uses_the_copy_engines_by_default(gplib):
    """At two ranks the step's exchange is the copy-engine path (parallel.NvlsFlatGradAllReduce.ce_pull); the check above
    ran through it -- this one reads which exchange the line names, and that GP_EXCHANGE_CE=0 still selects the kernel."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    kinds = {}
    for ce in ("1", "0"):
        port = 29300 + (os.getpid() + int(ce)) % 250
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
               "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "bench.py"), "--gpus", "2", "--workload", "tiny",
               "--steps", "3", "--warmup", "3", "--no-profile", "--exchange", "nvls"]
        out = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=600, env=dict(os.environ, GP_EXCHANGE_CE=ce))
        assert out.returncode == 0, out.stderr[-2000:]
        line = json.loads([l for l in out.stdout.splitlines() if l.startswith("{")][-1])
        assert line["exchange_check"]["ok"], line["exchange_check"]
        kinds[ce] = line["exchange_check"]["kind"]
    assert "copy engines" in kinds["1"] and "peer loads/stores" in kinds["0"], kinds


def test_exchange_pieces_on_one_gpu(gplib):
    """The kernels of the copy-engine path with a world of one rank: the flag barrier meets itself (epochs grow by one per
    call on its slot), the copy copies, the in-place add and max do what they say."""
    import ctypes
    from gaussianprediction_b200 import _lib
    L = gplib
    dev = torch.device("cuda:0")
    flags = torch.zeros(int(L.gp_allreduce_flag_bytes(4)) // 4, dtype=torch.int32, device=dev)
    peers = (ctypes.c_void_p * 1)(flags.data_ptr())
    s = _lib.stream_ptr()
    for _ in range(3):
        _lib.check(L.gp_peer_barrier_launch(0, 1, peers, 0, 2, s), "barrier")
    torch.cuda.synchronize()
    words = flags[:4 * 17].view(4, 17)              # [slot][arrival word of each of 16 peers | the slot's epoch]
    assert int(words[2, 0]) == 3 and int(words[2, 16]) == 3 and int(words[0].abs().sum()) == 0
    g = torch.Generator().manual_seed(1)
    a, b = torch.randn(4096 + 64, generator=g).to(dev), torch.randn(4096 + 64, generator=g).to(dev)
    c = torch.empty_like(b)
    _lib.check(L.gp_copy_async(c.data_ptr(), b.data_ptr(), b.numel() * 4, s), "copy")
    want = a + b
    _lib.check(L.gp_add_inplace_f32(a.data_ptr(), c.data_ptr(), a.numel(), s), "add")
    torch.cuda.synchronize()
    assert torch.equal(c, b) and torch.equal(a, want)
    x = torch.randint(-5, 50, (1003,), generator=g, dtype=torch.int32).to(dev)
    y = torch.randint(-5, 50, (1003,), generator=g, dtype=torch.int32).to(dev)
    wantm = torch.maximum(x, y)
    _lib.check(L.gp_max_inplace_s32(x.data_ptr(), y.data_ptr(), x.numel(), s), "max")
    torch.cuda.synchronize()
    assert torch.equal(x, wantm)
    assert L.gp_add_inplace_f32(a.data_ptr() + 4, c.data_ptr(), 8, s) != 0      # misaligned: refused
