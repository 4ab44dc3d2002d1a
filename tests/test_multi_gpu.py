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
