#!/usr/bin/env python
"""Headline benchmark of the per-frame render hot path: train iterations (keypoint deformation forward -> rasterize
forward -> backward through both) of the reference's ``--batch`` loop (train.py:113-133).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload NAME] [--mode train|extrapolate]

ONE job at every N (so that per-N values compare): BASELINE.json configs[4] -- 2 M Gaussians, 800x800, 512 keypoints,
nearest_num 6, SH degree 3, **8 (camera, time) samples per step split over the N ranks** (strong scaling): rank r
renders samples r, r+N, ... of the step, gradients are summed over the samples of a rank in place and over the ranks
with ONE exchange per step (SUM of gradients, MAX of radii = OR of visibility), started bucket by bucket while the
last sample's backward is still running.  One JSON line on stdout (rank 0):

  value        ms per train iteration, whole job = step time / 8, inputs resident in HBM, max over ranks
  e2e          same through the public API with every sample's camera + time + target image coming from pinned host
               memory and the step's loss read back
  secondary    (N = 1) ms per iteration on the other BASELINE.json configs, one sample per step: c4_1m_800x800 (the
               1 M-Gaussian scene north_star states its 2x target on), c2 (100 k), c3 (500 k, 536x960)
  roofline     the dominant kernel of the step (CUDA events around every launch in a second pass, gp_profile_enable)
               against MEASURED_PEAKS.json / the measured FP32 peak; ``rooflines`` holds the same for every kernel
  cpu_baseline the deformation stage (the only part of the path with a CPU implementation) on BASELINE.json
               configs[0] (20 k x 512, forward), oracle/ on the host cores: 1 thread and all cores
  --impl reference   the reference's own CUDA path (oracle/_ref/*.so + its torch ops, oracle/ref_pipeline.py), same job
  --mode extrapolate configs[3]: 1 M Gaussians, 200 timesteps x 50 views, frames sharded over the ranks, frames/s
"""
import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402

METRIC = "train-iter ms (deform+raster fwd/bwd)"
WORKLOADS = {
    "c2_dnerf_100k_800x800": dict(N=100_000, W=800, H=800, Kp=512, K=6, shape="dnerf"),
    "c3_hyper_500k_536x960": dict(N=500_000, W=536, H=960, Kp=512, K=6, shape="hyper"),
    "c4_1m_800x800": dict(N=1_000_000, W=800, H=800, Kp=512, K=6, shape="dnerf"),
    "c5_2m_800x800": dict(N=2_000_000, W=800, H=800, Kp=512, K=6, shape="dnerf"),
    "tiny": dict(N=5_000, W=160, H=112, Kp=64, K=6, shape="dnerf"),
}
HEADLINE = "c5_2m_800x800"
SECONDARY = ("c4_1m_800x800", "c2_dnerf_100k_800x800", "c3_hyper_500k_536x960")
SAMPLES_PER_STEP = 8
ITER = 50_000          # > second_stage_iteration + xyz_noise_iteration: keypoint stage, noise factor 0


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=HEADLINE, choices=list(WORKLOADS))
    ap.add_argument("--samples-per-step", type=int, default=SAMPLES_PER_STEP)
    ap.add_argument("--mode", default="train", choices=["train", "extrapolate"])
    ap.add_argument("--frames", type=int, default=200 * 50, help="extrapolate: timesteps x views")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-profile", action="store_true")
    ap.add_argument("--no-secondary", action="store_true")
    ap.add_argument("--ncams", type=int, default=16)
    ap.add_argument("--exchange", default=os.environ.get("GP_EXCHANGE", "auto"), choices=["auto", "nccl", "nvls"])
    return ap.parse_args()


# ------------------------------------------------------------------------------------------------ scene

def build_inputs(wl, ncams, seed=2024):
    from gaussianprediction_b200 import synthetic as syn
    gs = syn.make_gaussians(wl["N"], seed=seed, shape=wl["shape"])
    kp, kf = syn.make_keypoints(gs, Kp=wl["Kp"], seed=seed + 1)
    logits = syn.make_logits(wl["N"], wl["K"], seed=seed + 2)
    if wl["shape"] == "hyper":
        g = torch.Generator().manual_seed(seed + 3)
        jit = (torch.rand(ncams, 2, generator=g) - 0.5) * 0.6
        cams = [syn.hyper_camera(wl["W"], wl["H"], focal=700.0 * wl["W"] / 536.0, jitter=(float(a), float(b))) for a, b in jit]
    else:
        cams = syn.camera_ring(ncams, wl["W"], wl["H"], seed=seed + 3)
    times = torch.linspace(0.05, 0.95, ncams)
    g = torch.Generator().manual_seed(seed + 4)
    grads = [torch.randn(3, wl["H"], wl["W"], generator=g) * 1e-2 for _ in range(4)]   # dL/dcolor
    return gs, kp, kf, logits, cams, times, grads


def model_args():
    from gaussianprediction_b200.deform import default_args
    return default_args(feature_amplify=0.5, norm_rotation=True)     # D-NeRF scripts: feature_amplify 0.5, norm_rotation


def pin_cam(cam):
    from types import SimpleNamespace
    c = SimpleNamespace(**vars(cam))
    for k in ("world_view_transform", "full_proj_transform", "camera_center"):
        setattr(c, k, getattr(cam, k).pin_memory())
    return c


def cam_to(cam, dev, non_blocking=False):
    from types import SimpleNamespace
    c = SimpleNamespace(**vars(cam))
    for k in ("world_view_transform", "full_proj_transform", "camera_center"):
        setattr(c, k, getattr(cam, k).to(dev, non_blocking=non_blocking))
    return c


# ------------------------------------------------------------------------------------------------ clocks sampler

class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.rows, self.proc, self.idx = [], None, gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.idx}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm = [float(r[1]) for r in self.rows if len(r) >= 9 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if len(r) >= 9 and r[2].replace(".", "").isdigit()]
        reasons = set()
        for r in self.rows:
            if len(r) >= 9:
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------------------ our arm

class OurRunner:
    """This package's arm: ``S`` samples per step, this rank renders samples rank, rank + world, ... of each step."""

    def __init__(self, wl, dev, ncams, world, samples_per_step=1, exchange="auto"):
        from gaussianprediction_b200.deform import KeypointDeformation
        from gaussianprediction_b200.renderer import SceneModel
        self.wl, self.dev, self.world, self.S = wl, dev, world, int(samples_per_step)
        assert self.S % world == 0, "samples per step must divide evenly over the ranks"
        self.local = self.S // world
        self.exchange = exchange
        self.rank = int(os.environ.get("RANK", "0")) if world > 1 else 0
        gs, kp, kf, logits, cams, times, grads = build_inputs(wl, ncams)
        torch.manual_seed(2024)
        deform = KeypointDeformation(model_args(), gs.xyz, gs.rotation, gs.scaling, gs.opacity, gs.motion_feature, kp, kf,
                                     logits=logits)
        self.pc = SceneModel(deform, gs.features_dc, gs.features_rest).to(dev)
        self.params = [p for p in self.pc.parameters()]
        self.cams_host = [pin_cam(c) for c in cams]
        self.cams_dev = [cam_to(c, dev) for c in cams]
        self.times_host = times.pin_memory()
        self.times_dev = [times[i:i + 1].to(dev) for i in range(len(times))]
        self.grads_host = [g.pin_memory() for g in grads]
        self.grads_dev = [g.to(dev) for g in grads]
        self.bg = torch.zeros(3, device=dev)
        self.capacity = None
        self.batch = None
        self.h2d_bytes = self.S * (grads[0].numel() * 4 + (16 + 16 + 3 + 1) * 4)     # whole job, per step
        self.d2h_bytes = 4 * world

    # sample g of the global sequence: step s, slot k of this rank -> g = s * S + rank + k * world (parallel.shard_samples)
    def _ids(self, s):
        return [s * self.S + self.rank + k * self.world for k in range(self.local)]

    def sample(self, g, host=False):
        k = g % len(self.cams_dev)
        if host:
            return self.cams_host[k], self.times_host[k:k + 1], self.grads_host[g % len(self.grads_host)]
        return self.cams_dev[k], self.times_dev[k], self.grads_dev[g % len(self.grads_dev)]

    def step(self, s, host_inputs=False):
        """Eager launches (no graph): autograd accumulates the samples' gradients in ``p.grad``."""
        from gaussianprediction_b200.renderer import render
        for p in self.params:
            p.grad = None
        total = None
        for g in self._ids(s):
            cam, t, G = self.sample(g, host_inputs)
            if host_inputs:
                cam, t, G = cam_to(cam, self.dev, True), t.to(self.dev, non_blocking=True), G.to(self.dev, non_blocking=True)
            pkg = render(cam, self.pc, None, self.bg, time=t, it=ITER, capacity=self.capacity)
            loss = (pkg["render"] * G).sum()
            loss.backward()
            total = loss.detach() if total is None else total + loss.detach()
        if self.world > 1:
            self.allreduce()
        return total

    def make_reducer(self):
        from gaussianprediction_b200 import parallel
        d = self.pc.deformation
        live = [p for p in self.params if id(p) in self._live]
        if self.world == 1:
            return parallel.FlatGradAllReduce(live)                 # one bucket -> the step stays one graph
        cls = parallel.FlatGradAllReduce
        first_call = not getattr(self, "_have_acc", False)       # the accumulator is exchanged, the scratch buffer is not
        self._have_acc = True
        if first_call and self.exchange in ("auto", "nvls"):
            if getattr(self, "_nvls_ok", None) is None:
                try:
                    self._nvls_ok = parallel.NvlsFlatGradAllReduce.supported()
                except Exception:
                    self._nvls_ok = False
            if self._nvls_ok:
                cls = parallel.NvlsFlatGradAllReduce
            elif self.exchange == "nvls":
                raise SystemExit("--exchange nvls: no symmetric memory between the ranks here")
        red = cls(live, first=[self.pc._features_dc, self.pc._features_rest],
                  # xyz goes with the keypoints and MLP weights into the LAST bucket: the next step's neighbour search needs
                  # it first, so that bucket is the one joined at the end of the step; the others are joined later
                  second=[d._rotation, d._scaling, d._opacity] + list(d.weights_model.parameters()),
                  aux_int32=self.wl["N"] if first_call else 0)       # the radii ride the same exchange (MAX)
        if first_call:
            if cls is parallel.FlatGradAllReduce:
                self.exchange_kind = "NCCL all_reduce per bucket"
            elif getattr(red, "copy_engine", False):
                self.exchange_kind = ("symmetric memory, 2 ranks: copy engines pull the peer's ranges (cudaMemcpyAsync), "
                                      "gp_add_inplace_kernel in front of each first reader")
            else:
                self.exchange_kind = ("gp_allreduce_sum_kernel over symmetric memory (" +
                                      ("NVSwitch multicast, multimem.ld_reduce" if red.multicast else "peer loads/stores") + ")")
        return red

    def enable_graph(self):
        """Capture the iteration (gaussianprediction_b200.pipeline.BatchStep: one or two captured TrainSteps)."""
        from gaussianprediction_b200.pipeline import BatchStep
        # only parameters that receive a gradient on this path take part (known after the eager pass)
        self._live = {id(p) for p in self.params if p.grad is not None}
        self.exchange_kind = "none"
        self.batch = BatchStep(self.pc, self.cams_dev[0], self.capacity, self.make_reducer, self.local, bg=self.bg,
                               iteration=ITER)
        if os.environ.get("GP_BENCH_NO_EXCHANGE"):       # DIAGNOSTIC ONLY: the compute of a step without its exchange
            self.batch.exchange = False
            self.exchange_kind += " -- DISABLED (GP_BENCH_NO_EXCHANGE): not a valid result"
        # device-resident samples in the packed form the captured step keeps its camera in (view | proj | centre | time)
        self.packed_dev = [torch.cat([c.world_view_transform.reshape(-1), c.full_proj_transform.reshape(-1),
                                      c.camera_center.reshape(-1), t.reshape(-1)]).contiguous()
                           for c, t in zip(self.cams_dev, self.times_dev)]

    def step_graph(self, s, host_inputs=False):
        ids = self._ids(s)
        if host_inputs:
            nxt = self.sample(self._ids(s + 1)[0], True)
            return self.batch.step_host([self.sample(g, True) for g in ids], next_first=nxt)
        return self.batch.step([(self.packed_dev[g % len(self.packed_dev)], self.grads_dev[g % len(self.grads_dev)])
                                for g in ids])

    def allreduce(self):
        if getattr(self, "_grouped", None) is None:
            from gaussianprediction_b200.parallel import GroupedGradAllReduce
            self._grouped = GroupedGradAllReduce(self.params)
        self._grouped.reduce()

    def measured_fps(self, frames, world):
        """Forward-only frames/s of the whole job: every rank renders ``frames`` frames of its shard through the captured
        ``RenderStep`` between two barriers; the reference's span (eval.py:210-214) is one frame between two syncs."""
        import torch.distributed as dist
        from gaussianprediction_b200.pipeline import RenderStep
        rs = RenderStep(self.pc, self.cams_dev[0], self.capacity, bg=self.bg, iteration=ITER)
        for i in range(3):
            rs(self.cams_dev[i % len(self.cams_dev)], self.times_dev[i % len(self.times_dev)])
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        for i in range(frames):
            k = (i * world + self.rank) % len(self.cams_dev)
            rs(self.cams_dev[k], self.times_dev[k])
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        dt = max_over_ranks(dt, world, self.dev)
        del rs
        return frames * world / dt


class RefRunner:
    """The reference's CUDA path (oracle/ref_pipeline.py) on ONE GPU: all ``S`` samples of a step, gradients
    accumulated by autograd as the reference's summed loss does (train.py:113-133)."""

    def __init__(self, wl, dev, ncams, world, samples_per_step=1, exchange=None):
        from oracle.ref_pipeline import RefKeypointModel
        self.wl, self.dev, self.world, self.S = wl, dev, 1, int(samples_per_step)
        gs, kp, kf, logits, cams, times, grads = build_inputs(wl, ncams)
        torch.manual_seed(2024)
        self.pc = RefKeypointModel(model_args(), gs.xyz, gs.rotation, gs.scaling, gs.opacity, gs.motion_feature, kp, kf,
                                   logits, gs.features_dc, gs.features_rest).to(dev)
        self.params = [p for p in self.pc.parameters()]
        self.cams_host = [pin_cam(c) for c in cams]
        self.cams_dev = [cam_to(c, dev) for c in cams]
        self.times_host = times.pin_memory()
        self.times_dev = [times[i:i + 1].to(dev) for i in range(len(times))]
        self.grads_host = [g.pin_memory() for g in grads]
        self.grads_dev = [g.to(dev) for g in grads]
        self.bg = torch.zeros(3, device=dev)
        self.h2d_bytes = self.S * (grads[0].numel() * 4 + (16 + 16 + 3 + 1) * 4)
        self.d2h_bytes = 4

    def step(self, s, host_inputs=False):
        from oracle.ref_pipeline import ref_render
        for p in self.params:
            p.grad = None
        total = None
        for g in range(s * self.S, (s + 1) * self.S):
            k = g % len(self.cams_dev)
            if host_inputs:
                cam = cam_to(self.cams_host[k], self.dev, non_blocking=True)
                t = self.times_host[k:k + 1].to(self.dev, non_blocking=True)
                G = self.grads_host[g % len(self.grads_host)].to(self.dev, non_blocking=True)
            else:
                cam, t, G = self.cams_dev[k], self.times_dev[k], self.grads_dev[g % len(self.grads_dev)]
            pkg = ref_render(cam, self.pc, self.bg, t, ITER)
            loss = (pkg["render"] * G).sum()
            loss.backward()
            total = loss.detach() if total is None else total + loss.detach()
        return total

    step_graph = step

    def measured_fps(self, frames, world):
        from oracle.ref_pipeline import ref_render
        with torch.no_grad():
            for i in range(2):
                ref_render(self.cams_dev[i], self.pc, self.bg, self.times_dev[i], ITER)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for i in range(frames):
                k = i % len(self.cams_dev)
                ref_render(self.cams_dev[k], self.pc, self.bg, self.times_dev[k], ITER)
            torch.cuda.synchronize()
        return frames / (time.perf_counter() - t0)


# ------------------------------------------------------------------------------------------------ timing

def timed_steps(runner, steps, warmup, world, host_inputs, flush, graph=False):
    """W untimed steps, then exactly K steps; every step is bracketed by CUDA events on the current stream, L2 is
    flushed (256 MiB write) between steps outside the events; returns (per-step device ms list, wall ms)."""
    import torch.distributed as dist
    step = runner.step_graph if graph else runner.step
    # End to end: the loss of every step is copied to pinned host memory inside the timed region and read by the host
    # ONE STEP LATER (the reference reads it for its progress bar only, train.py:139-147), so the host keeps enqueueing
    # while the GPU computes; the last step's loss is read before its closing event.
    pinned = [torch.zeros((), dtype=torch.float32).pin_memory() for _ in range(2)] if host_inputs else None
    copied = [torch.cuda.Event(), torch.cuda.Event()] if host_inputs else None
    for i in range(warmup):
        loss = step(i, host_inputs)
        if host_inputs:
            loss.item()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    losses = []
    t0 = time.perf_counter()
    for i in range(steps):
        flush.add_(1.0)                      # evict L2 (126 MB) between timed steps
        evs[i][0].record()
        loss = step(warmup + i, host_inputs)
        if i == steps - 1 and graph and getattr(runner, "batch", None) is not None:
            runner.batch.flush()             # pipelined exchange: the last step's buckets are joined inside its bracket
        if host_inputs:                      # D2H read of every step's result inside the timed region
            pinned[i & 1].copy_(loss.detach(), non_blocking=True)
            copied[i & 1].record()
            if i > 0:
                copied[(i - 1) & 1].synchronize()
                losses.append(float(pinned[(i - 1) & 1]))
            if i == steps - 1:
                copied[i & 1].synchronize()
                losses.append(float(pinned[i & 1]))
        evs[i][1].record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    wall_ms = (time.perf_counter() - t0) * 1e3
    if host_inputs:
        assert len(losses) == steps and all(math.isfinite(x) for x in losses), "a step's loss did not reach the host"
    return [a.elapsed_time(b) for a, b in evs], wall_ms


def max_over_ranks(x, world, dev):
    if world == 1:
        return x
    import torch.distributed as dist
    t = torch.tensor([x], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def profile_pass(runner, steps, warmup):
    """Second pass over the same steps with a CUDA-event pair around every library launch (same stream, eager
    launches); returns {kernel: (launches, ms)}."""
    import ctypes
    from gaussianprediction_b200 import _lib
    L = _lib.lib()
    world, runner.world = runner.world, 1          # no exchange in the profiling pass (per-kernel times only)
    try:
        runner.step(warmup, False)
        torch.cuda.synchronize()
        L.gp_profile_enable(1)
        for i in range(steps):
            runner.step(warmup + i, False)
        torch.cuda.synchronize()
    finally:
        runner.world = world
    buf = ctypes.create_string_buffer(1 << 16)
    L.gp_profile_report(buf, len(buf))
    L.gp_profile_enable(0)
    out = {}
    for line in buf.value.decode().splitlines():
        name, n, ms = line.split()
        out[name] = (int(n), float(ms))
    return out


def workload_stats(runner):
    """Exact-mode pass over every (camera, time) sample of the run: the largest tile-instance count R (sizes the
    sync-free capacity) and, for sample 0, R and the evaluated pixel x Gaussian pairs (algorithmic-work figures)."""
    from gaussianprediction_b200 import rasterizer as R
    from gaussianprediction_b200.renderer import _settings
    out = {}
    max_r = 0
    with torch.no_grad():
        for i, cam in enumerate(runner.cams_dev):
            xyz, q, s, o = runner.pc(runner.times_dev[i], ITER)
            rs = _settings(cam, runner.pc, runner.bg, 1.0)
            color, radii, depth, tidx, st = R.raster_forward(xyz, runner.pc.get_features, None, o, s, q, None, rs)
            max_r = max(max_r, int(st.num_rendered))
            if i == 0:
                ex = R.export_state(st)
                out = dict(R=int(st.num_rendered), visible=int((radii > 0).sum()),
                           pairs=int(ex["n_contrib"].to(torch.int64).sum()))
            del color, radii, depth, tidx, st
    out["max_R"] = max_r
    return out


# dram__bytes_read.sum + dram__bytes_write.sum per launch from the committed ncu --set full captures (profiles/), keyed
# by workload; other workloads report null
# ncu dram__bytes_read.sum + dram__bytes_write.sum per launch of the kernels as committed (profiles/r2_ncu_full.md, C5)
NCU_DRAM_BYTES = {"c5_2m_800x800": {"gp_render_bwd_kernel": 219.5e6, "gp_render_fwd_kernel": 181.4e6,
                                    "gp_knn_kernel": 514.4e6, "gp_preprocess_bwd_kernel": 1074.4e6,
                                    "gp_blend_bwd_kernel": 445.6e6, "gp_preprocess_kernel": 629.6e6,
                                    "gp_duplicate_kernel": 242.4e6, "gp_blend_fwd_kernel": 297.0e6}}
NCU_DRAM_SOURCE = "profiles/r2_ncu_full.md"


def measured_fp32_peak(dev, peaks):
    """FP32 FMA-pipe peak of this GPU, measured (gp_fma_peak: 8 independent FFMA chains per thread, grid = 8 x SMs),
    next to the derived 148 SM x 128 lanes x 2 x clock.  TFLOP/s."""
    import ctypes
    from gaussianprediction_b200 import _lib
    L = _lib.lib()
    derived = 148 * 128 * 2 * peaks.get("sm_max_mhz", 1965.0) * 1e6 / 1e12
    if not hasattr(L, "gp_fma_peak"):
        return derived, derived, "derived 148 SM x 128 lanes x 2 x sm_max_mhz"
    out = torch.zeros(1, device=dev)
    best = 0.0
    iters = 1 << 14
    for _ in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        flops = ctypes.c_double()
        _lib.check(L.gp_fma_peak(iters, out.data_ptr(), ctypes.byref(flops), _lib.stream_ptr()), "gp_fma_peak")
        e1.record()
        torch.cuda.synchronize()
        best = max(best, flops.value / (e0.elapsed_time(e1) * 1e-3) / 1e12)
    return best, derived, "measured in this run (gp_fma_peak, best of 5)"


def rooflines(prof, stats, wl, steps, samples, peaks, fp32_peak, fp32_src):
    """Achieved rate of EVERY kernel of the step against the roofline that bounds it.  Algorithmic work per launch
    (DESIGN.md par. 4): HBM kernels in bytes (SURVEY.md par. 8d per-unit figures), blend kernels in FP32 operations per
    evaluated pixel x Gaussian pair (18 forward, 45 backward, SURVEY.md par. 8d hand counts; pairs = sum of n_contrib),
    the search in sub + FMA per (Gaussian, keypoint, dimension)."""
    if not prof:
        return None, {}, {}
    per = {k: v[1] / max(v[0], 1) for k, v in prof.items()}            # ms per launch
    tot = {k: v[1] / steps for k, v in prof.items()}                    # ms per step
    calls = {k: v[0] / steps for k, v in prof.items()}
    N, R, pairs, Kp = wl["N"], stats["R"], stats["pairs"], wl["Kp"]
    hbm = peaks.get("hbm_gbs", 6650.0)
    mlp_fma = Kp * (104 * 256 + 3 * 256 * 256 + 256 * 7)
    work = {
        "gp_preprocess_kernel": ("hbm", 311.0 * N),
        "gp_preprocess_bwd_kernel": ("hbm", 535.0 * N),
        "gp_rs_pass_kernel": ("hbm", None),
        "gp_render_fwd_kernel": ("fp32", 18.0 * pairs),
        "gp_render_bwd_kernel": ("fp32", 45.0 * pairs),
        # the search prunes (tens of the 512 keypoints survive per warp): its exhaustive-equivalent rate would exceed the
        # pipe's peak, so it gets a time but no roofline fraction
        "gp_knn_kernel": ("fp32", None),
        "gp_blend_fwd_kernel": ("hbm", 128.0 * N),
        "gp_blend_bwd_kernel": ("hbm", 204.0 * N),
        "gp_block_sum_kernel": ("hbm", 8.0 * N),
        "gp_duplicate_kernel": ("hbm", 36.0 * N + 8.0 * R),
        "gp_tile_ranges_kernel": ("hbm", 4.0 * R),
        "gp_kpmlp_fwd_kernel": ("fp32", 2.0 * mlp_fma),
        "gp_kpmlp_bwd_data_kernel": ("fp32", 2.0 * mlp_fma),
        "gp_kpmlp_bwd_weights_kernel": ("fp32", 2.0 * mlp_fma),
    }
    # the sort: 16 B per key per pass (key + value, read + write); 4 passes over P depth keys + the tile-id passes over R
    n_pass = calls.get("gp_rs_pass_kernel", 0) / max(samples, 1)       # sort passes per sample
    if n_pass:
        depth_passes = min(4.0, n_pass)
        work["gp_rs_pass_kernel"] = ("hbm", (depth_passes * 16.0 * N + (n_pass - depth_passes) * 16.0 * R) / n_pass)
    table = {}
    for k in sorted(tot, key=tot.get, reverse=True):
        kind, w = work.get(k, ("hbm", None))
        row = {"ms_per_step": round(tot[k], 4), "ms_per_launch": round(per[k], 4), "launches_per_step": round(calls[k], 2),
               "bound": kind}
        if w is not None:
            sec = per[k] * 1e-3
            if kind == "hbm":
                ach, peak, unit = w / sec / 1e9, hbm, "GB/s"
            else:
                ach, peak, unit = w / sec / 1e12, fp32_peak, "TFLOP/s"
            row.update(achieved=round(ach, 3), peak=round(peak, 3), unit=unit, frac=round(ach / peak, 4))
        if k in NCU_DRAM_BYTES.get(wl.get("name"), {}):
            row["traffic"] = NCU_DRAM_BYTES[wl["name"]][k]      # measured DRAM bytes per launch (ncu), next to the algorithmic ones
        table[k] = row
    # the dominant kernel that HAS a roofline: the neighbour search (pruned, no algorithmic-work figure) is named beside it
    # when it is the longer one
    rated = [k for k in tot if table[k].get("frac") is not None]
    top = max(rated or list(tot), key=tot.get)
    longest = max(tot, key=tot.get)
    r = dict(table[top])
    roof = {"kernel": top, "bound": r["bound"], "achieved": r.get("achieved"), "peak": r.get("peak"), "unit": r.get("unit"),
            "frac": r.get("frac"), "traffic": NCU_DRAM_BYTES.get(wl.get("name"), {}).get(top),
            "traffic_unit": f"bytes/launch (ncu dram read+write, {NCU_DRAM_SOURCE})",
            "ms_per_launch": r["ms_per_launch"], "share_of_kernel_time": round(tot[top] / sum(tot.values()), 4),
            "peak_source": "MEASURED_PEAKS.json hbm_gbs" if r["bound"] == "hbm" else f"FP32 FMA pipe, {fp32_src}"}
    if longest != top:
        roof["note"] = (f"{longest} is longer ({round(per[longest], 4)} ms per launch, "
                        f"{round(tot[longest] / sum(tot.values()), 4)} of the kernel time) but has no roofline fraction: it "
                        "prunes, so its exhaustive-equivalent rate would exceed the pipe's peak (DESIGN.md par. 4)")
    return roof, table, {k: round(v, 4) for k, v in sorted(tot.items(), key=lambda kv: -kv[1])}


# ------------------------------------------------------------------------------------------------ CPU baseline

def cpu_deformation_baseline(dev, budget_s=12.0):
    """BASELINE.md par. 4(B): the deformation stage (kNN + blend, forward, no_grad, float32) on BASELINE.json configs[0]
    (20 k Gaussians x 512 keypoints, nearest_num 6, D = 35) on the host cores with the CPU oracle (oracle/deform_oracle.py,
    pinned against the reference's own source), at 1 thread (the reference pins OMP/MKL to 1, train.py:13-14) and at all
    cores; >= 3 warm-ups, then up to 20 runs within the time budget.  The rasterizer has no CPU implementation.  Also
    timed: the reference's own single-threaded C++ kNN (frnn_bf_cpu) and this library on the same sample."""
    from gaussianprediction_b200 import synthetic as syn
    from gaussianprediction_b200.deform import DeformableField
    from oracle import deform_oracle as do
    N, Kp, K = 20_000, 512, 6
    gs = syn.make_gaussians(N, seed=2024)
    kp, kf = syn.make_keypoints(gs, Kp=Kp, seed=2025)
    logits = syn.make_logits(N, K, seed=2026)
    a = model_args()
    torch.manual_seed(2024)
    net = DeformableField(2 * a.time_freq + 6 * a.xyz_freq + a.feature_dim, 7, a.d, a.w)
    sd = dict(net.state_dict())
    t = torch.tensor([0.37])

    def torch_knn(p1, p2):
        # the reference's accumulation order (sequential over d, SURVEY.md par. 8c), vectorised over the pairs
        acc = torch.zeros(p1.shape[0], p2.shape[0])
        for d in range(p1.shape[1]):
            diff = p2[None, :, d] - p1[:, None, d]
            acc = acc + diff * diff
        return torch.topk(acc, K, dim=1, largest=False, sorted=True)[1]

    def once():
        with torch.no_grad():
            p1, p2 = do.hybrid_points(gs.xyz, gs.motion_feature, kp, kf, a.feature_amplify)
            idx = torch_knn(p1, p2)
            do.keypoint_forward(t, gs.xyz, gs.rotation, gs.scaling, gs.opacity, kp, kf, logits, idx, sd,
                                norm_rotation=a.norm_rotation)

    def timed(threads):
        old = torch.get_num_threads()
        torch.set_num_threads(threads)
        try:
            for _ in range(3):
                once()
            ts, t_end = [], time.perf_counter() + budget_s
            while len(ts) < 20 and (len(ts) < 3 or time.perf_counter() < t_end):
                t0 = time.perf_counter()
                once()
                ts.append((time.perf_counter() - t0) * 1e3)
        finally:
            torch.set_num_threads(old)
        return float(np.median(ts)), len(ts)

    cores = os.cpu_count() or 1
    one_ms, one_runs = timed(1)
    all_ms, all_runs = timed(cores)
    out = {"value": round(one_ms, 2), "unit": "ms per deformation forward (kNN + keypoint MLP + blend)", "cores": 1,
           "kind": "port", "runs": one_runs,
           "sample": "BASELINE.json configs[0]: 20k Gaussians x 512 keypoints, nearest_num 6, D=35, forward only, float32, "
                     "median of the timed runs after 3 warm-ups; torch CPU restatement of the reference ops (oracle/)",
           "all_cores": {"value": round(all_ms, 2), "cores": cores, "runs": all_runs}}
    try:
        from oracle import ref_cuda
        p1, p2 = do.hybrid_points(gs.xyz, gs.motion_feature, kp, kf, a.feature_amplify)
        t0 = time.perf_counter()
        ref_cuda.frnn_bf_cpu(p1[None].contiguous(), p2[None].contiguous(), K, 1e8)
        out["reference_frnn_bf_cpu_ms"] = round((time.perf_counter() - t0) * 1e3, 2)
    except Exception as e:                                              # pragma: no cover
        out["reference_frnn_bf_cpu_ms"] = f"unavailable: {type(e).__name__}"
    try:
        from gaussianprediction_b200.deform import KeypointDeformation
        m = KeypointDeformation(a, gs.xyz, gs.rotation, gs.scaling, gs.opacity, gs.motion_feature, kp, kf, logits=logits).to(dev)
        tt = t.to(dev)
        with torch.no_grad():
            for _ in range(3):
                m(tt, ITER)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(20):
                m(tt, ITER)
            e1.record()
            torch.cuda.synchronize()
        out["gpu_same_sample_ms"] = round(e0.elapsed_time(e1) / 20, 4)
    except Exception as e:                                              # pragma: no cover
        out["gpu_same_sample_ms"] = f"failed: {type(e).__name__}: {e}"
    return out


# ------------------------------------------------------------------------------------------------ checks / modes

def exchange_check(runner, s):
    """Outside the timed region (world > 1): the exchanged gradient buffer of one step must equal the sum over the ranks
    of the per-rank sums, formed independently (all_gather of the un-exchanged buffers, added in rank order)."""
    import torch.distributed as dist
    b = runner.batch
    b.flush()
    b.exchange = False
    b.step([runner.sample(g) for g in runner._ids(s)])
    torch.cuda.synchronize()
    local = b.acc.flat.clone()
    radii_local = b.radii.clone()
    b.exchange = True
    parts = [torch.empty_like(local) for _ in range(runner.world)]
    dist.all_gather(parts, local)
    want = parts[0].double()
    for p in parts[1:]:
        want += p.double()
    rparts = [torch.empty_like(radii_local) for _ in range(runner.world)]
    dist.all_gather(rparts, radii_local)
    rwant = torch.stack(rparts).max(0)[0]
    del parts
    b.step([runner.sample(g) for g in runner._ids(s)])
    b.flush()
    torch.cuda.synchronize()
    got = b.acc.flat.double()
    scale = float(want.abs().max())
    err = float((got - want).abs().max()) / max(scale, 1e-30)
    out = {"max_rel_err": err, "radii_equal": bool(torch.equal(b.radii, rwant)), "bytes": int(local.numel() * 4),
           "kind": getattr(runner, "exchange_kind", "?")}
    # per bucket too (a bucket that is all zeros would hide behind the global scale)
    per = []
    for _, _, e0, e1 in b.acc.bucket_ranges:
        sc = float(want[e0:e1].abs().max())
        per.append(float((got[e0:e1] - want[e0:e1]).abs().max()) / max(sc, 1e-30))
    out["max_rel_err_per_bucket"] = per
    out["ok"] = bool(max(per) <= 1e-4 and out["radii_equal"])      # run-to-run atomic-order noise is ~1e-6
    t = torch.tensor([1.0 if out["ok"] else 0.0], device=local.device)
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    out["ok"] = bool(t.item() > 0)
    return out


def exchange_time(runner, reps=10):
    """The exchange alone (all buckets back to back on idle GPUs), ms: what a step would pay with no overlap."""
    import torch.distributed as dist
    acc = runner.batch.acc
    for _ in range(2):
        for w in [acc.all_reduce_bucket(b) for b in acc.buckets()]:
            if w is not None:
                w.wait()
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        for w in [acc.all_reduce_bucket(b) for b in acc.buckets()]:
            if w is not None:
                w.wait()
    e1.record()
    torch.cuda.synchronize()
    acc.flat.zero_()
    return max_over_ranks(e0.elapsed_time(e1) / reps, runner.world, runner.dev)


def run_single(name, impl, dev, steps, warmup, flush, ncams=8):
    """One secondary workload at one sample per step on one GPU: ms per train iteration, device-resident and end to end."""
    wl = dict(WORKLOADS[name], name=name)
    cls = OurRunner if impl == "ours" else RefRunner
    r = cls(wl, dev, ncams, 1, 1)
    out = {"workload": name, "gaussians": wl["N"], "image": f'{wl["W"]}x{wl["H"]}', "samples_per_step": 1}
    graph = False
    if impl == "ours":
        st = workload_stats(r)
        r.capacity = int(st["max_R"] * 1.25) + 65536
        r.step(0)
        r.enable_graph()
        graph = True
        out.update(tile_instances=st["R"], blend_pairs=st["pairs"])
    per, _ = timed_steps(r, steps, warmup, 1, False, flush, graph=graph)
    per_e, _ = timed_steps(r, steps, warmup, 1, True, flush, graph=graph)
    if impl == "ours" and r.batch.overflowed()[1]:
        raise SystemExit(f"{name}: tile-instance capacity overflowed during the timed region")
    out.update(ms_per_iter=round(sum(per) / steps, 4), ms_per_iter_median=round(float(np.median(per)), 4),
               e2e_ms_per_iter=round(sum(per_e) / steps, 4))
    del r
    torch.cuda.empty_cache()
    return out


def extrapolate_mode(args, dev, rank, world):
    """BASELINE.json configs[3]: 1 M Gaussians, GCN-style keypoint trajectories, ``frames`` = timesteps x views frames
    sharded f = r (mod world), no communication; neighbour indices / blend logits computed once per rank (eval.py:125-126).
    Reports measured aggregate frames/s (barrier, all ranks' frames, barrier)."""
    import torch.distributed as dist
    from gaussianprediction_b200 import synthetic as syn
    name = args.workload if args.workload != HEADLINE else "c4_1m_800x800"
    wl = dict(WORKLOADS[name], name=name)
    n_t, n_v = 200, 50
    frames = min(args.frames, n_t * n_v)
    line = {"metric": "render FPS (motion extrapolation: per-frame blend + rasterize forward)", "unit": "frames/s",
            "n_gpus": world if args.impl == "ours" else args.gpus, "higher_is_better": True, "scaling": "strong",
            "dtype": "f32", "data": "synthetic", "vs_baseline": None,
            "config": {"workload": name + "_extrapolate", "gaussians": wl["N"], "image": f'{wl["W"]}x{wl["H"]}',
                       "keypoints": wl["Kp"], "timesteps": n_t, "views": n_v, "frames": frames}}
    if args.impl == "ours":
        from gaussianprediction_b200.pipeline import MotionStep
        r = OurRunner(wl, dev, n_v, world, world)
        kpts, krot = syn.keypoint_trajectory(r.pc.deformation.super_gaussians.detach().cpu(), frames=n_t)
        kpts, krot = kpts.to(dev), krot.to(dev)
        st = workload_stats(r)
        cap = int(st["max_R"] * 1.6) + 65536            # moving keypoints change the instance count: headroom, checked below
        with torch.no_grad():
            r.pc(r.times_dev[0], ITER)                  # indices + logits once (eval.py:125-126)
        ms = MotionStep(r.pc, r.cams_dev[0], cap, bg=r.bg)
        mine = list(range(rank, frames, world))
        for f in mine[:3]:
            ms(r.cams_dev[f % n_v], kpts[f // n_v % n_t], krot[f // n_v % n_t])
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        over = False
        for i, f in enumerate(mine):
            ms(r.cams_dev[f % n_v], kpts[f // n_v % n_t], krot[f // n_v % n_t])
            if i % 512 == 511:
                over = over or ms.overflowed()[1]
        torch.cuda.synchronize()
        dt = max_over_ranks(time.perf_counter() - t0, world, dev)
        over = over or ms.overflowed()[1]
        if over:
            raise SystemExit("extrapolate: tile-instance capacity overflowed; result invalid")
        line.update(value=round(frames / dt, 2), ms_per_frame_per_gpu=round(dt / len(mine) * 1e3, 4), frames_timed=frames)
    else:
        from oracle import deform_oracle as do
        from oracle.ref_pipeline import ref_render_motion
        r = RefRunner(wl, dev, n_v, 1, 1)
        kpts, krot = syn.keypoint_trajectory(r.pc.super_gaussians.detach().cpu(), frames=n_t)
        kpts, krot = kpts.to(dev), krot.to(dev)
        sample = min(frames, 400)                        # bounded sample of the same frame sequence (stride through it)
        ids = [int(i * frames / sample) for i in range(sample)]
        with torch.no_grad():
            wx, wr = r.pc.dense_weights(r.times_dev[0], ITER)                    # eval.py:125-126, dense [N,Kp] x 2
            for f in ids[:2]:
                ref_render_motion(r.cams_dev[f % n_v], r.pc, r.bg, wx, wr, kpts[f // n_v % n_t], krot[f // n_v % n_t])
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for f in ids:
                ref_render_motion(r.cams_dev[f % n_v], r.pc, r.bg, wx, wr, kpts[f // n_v % n_t], krot[f // n_v % n_t])
            torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        line.update(value=round(sample / dt, 2), ms_per_frame_per_gpu=round(dt / sample * 1e3, 4), frames_timed=sample,
                    impl="reference")
    line["ms_per_step"] = line["ms_per_frame_per_gpu"]
    line["steps"], line["warmup"] = line["frames_timed"], 3
    return line


# ------------------------------------------------------------------------------------------------ main

def main():
    args = parse()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    wl = dict(WORKLOADS[args.workload], name=args.workload)
    steps, warmup = args.steps, max(args.warmup, 3)
    S = args.samples_per_step

    if args.impl == "reference" and rank != 0:
        return 0                                                   # rank 0 alone runs the reference arm
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (the product has no CPU path)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1 and args.impl == "ours":
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    eff_world = world if args.impl == "ours" else 1

    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass

    if args.impl == "ours":
        from gaussianprediction_b200 import _lib
        _lib.lib()                                                 # raises if the CUDA library is missing
    else:
        from oracle import ref_cuda
        for m in ("ref_dgr", "ref_frnn", "ref_prefix_sum"):
            if not ref_cuda.available(m):
                print(json.dumps({"impl": "reference", "unavailable": f"oracle/_ref/{m}.so not built"}))
                return 0

    if args.mode == "extrapolate":
        line = extrapolate_mode(args, dev, rank, eff_world)
        if rank == 0:
            print(json.dumps(line))
        if world > 1 and args.impl == "ours":
            import torch.distributed as dist
            dist.barrier()
            dist.destroy_process_group()
        return 0

    if S % eff_world != 0:
        raise SystemExit(f"--samples-per-step {S} does not divide over {eff_world} ranks")
    cls = OurRunner if args.impl == "ours" else RefRunner
    runner = cls(wl, dev, args.ncams, eff_world, S, exchange=args.exchange)

    stats = None
    if args.impl == "ours":
        stats = workload_stats(runner)
        runner.capacity = int(stats["max_R"] * 1.25) + 65536       # sync-free sizing: >= the exact R of every sample used

    flush = torch.zeros(64 * 1024 * 1024, device=dev)              # 256 MiB > 126 MB L2
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    use_graph, graph_note = False, "eager launches"
    eager_ms = None
    launches = 0
    if args.impl == "ours":
        # eager pass first: it counts the kernels of one step (a graph replay re-issues exactly those launches)
        n_eager = min(steps, 3)
        launches0 = _lib.lib().gp_launch_count()
        eager_per, _ = timed_steps(runner, n_eager, 1, eff_world, False, flush)
        launches = (_lib.lib().gp_launch_count() - launches0) // (n_eager + 1) * steps
        eager_ms = max_over_ranks(sum(eager_per), eff_world, dev) / n_eager / S
        runner.enable_graph()
        use_graph = True
        graph_note = ("every sample replayed as captured CUDA graphs (pipeline.BatchStep); exchange: " + runner.exchange_kind)
    per, wall_ms = timed_steps(runner, steps, warmup, eff_world, False, flush, graph=use_graph)
    per_e, _ = timed_steps(runner, steps, warmup, eff_world, True, flush, graph=use_graph)
    clocks = sampler.stop() if rank == 0 else None
    if use_graph and runner.batch.overflowed()[1]:
        raise SystemExit("tile-instance capacity overflowed during the timed region: result invalid")

    dev_ms = max_over_ranks(sum(per), eff_world, dev)
    e2e_dev_ms = max_over_ranks(sum(per_e), eff_world, dev)
    ms_per_step = dev_ms / steps
    value = ms_per_step / S                                        # whole job: ms per (camera, time) train iteration
    e2e_value = e2e_dev_ms / steps / S

    line = {
        "metric": METRIC, "value": round(value, 4), "unit": "ms/iter", "n_gpus": eff_world if args.impl == "ours" else args.gpus,
        "steps": steps, "warmup": warmup, "ms_per_step": round(ms_per_step, 4), "higher_is_better": False,
        "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": args.workload, "gaussians": wl["N"], "image": f'{wl["W"]}x{wl["H"]}', "keypoints": wl["Kp"],
                   "nearest_num": wl["K"], "knn_type": "hybird", "sh_degree": 3, "samples_per_step": S,
                   "l2": "flushed between timed steps (256 MiB write, outside the events)",
                   "timing": "sum of per-step CUDA-event times, max over ranks; value = step time / samples per step"},
        "e2e": {"value": round(e2e_value, 4), "unit": "ms/iter", "h2d_bytes_per_step": runner.h2d_bytes,
                "d2h_bytes_per_step": runner.d2h_bytes},
        "gpu_launches": int(launches),
        "ms_per_step_median": round(float(np.median(per)), 4),
        "train_iters_per_s": round(1e3 / value, 2),
        "wall_ms_per_step": round(wall_ms / steps, 4),
        "launch": graph_note,
    }
    if eager_ms is not None:
        line["eager_ms_per_iter"] = round(eager_ms, 4)
    if stats:
        line["workload_stats"] = dict(tile_instances=stats["R"], visible=stats["visible"], blend_pairs=stats["pairs"],
                                      max_tile_instances=stats["max_R"])
    if args.impl == "reference":
        line["impl"] = "reference"
        line["reference_gpus"] = 1
        line["cpu_baseline"] = {"value": line["value"], "unit": "ms/iter", "cores": 0, "kind": "reference",
                                "sample": "the reference's CUDA kernels (oracle/_ref, sm_100) + its torch ops on ONE B200 "
                                          "(all samples of the step on that GPU): north_star's bar; the rasterizer has no "
                                          "CPU implementation"}
    if clocks is not None:
        line["clocks"] = clocks

    if args.impl == "ours" and eff_world > 1 and not os.environ.get("GP_BENCH_NO_EXCHANGE"):
        chk = exchange_check(runner, warmup + steps)
        line["exchange_check"] = chk
        line["exchange_alone_ms"] = round(exchange_time(runner), 4)
        # how much of the exchange the step hides: the same steps once more with the exchange switched off (diagnostic
        # pass, never the reported value), max over ranks like the value itself
        runner.batch.flush()
        runner.batch.exchange = False
        k = max(3, min(steps, 10))
        per0, _ = timed_steps(runner, k, 2, eff_world, False, flush, graph=True)
        runner.batch.exchange = True
        compute_only = max_over_ranks(sum(per0), eff_world, dev) / k
        exposed = max(0.0, line["ms_per_step"] - compute_only)
        line["exchange_overlap"] = {"compute_only_ms_per_step": round(compute_only, 4),
                                    "exposed_ms_per_step": round(exposed, 4),
                                    "hidden_frac": round(max(0.0, 1.0 - exposed / max(line["exchange_alone_ms"], 1e-9)), 4),
                                    "note": "exposed = step time - the same steps with the exchange off; hidden = 1 - "
                                            "exposed / exchange_alone (the exchange run alone on idle GPUs)"}
        if not chk["ok"]:
            if rank == 0:
                print(json.dumps(line))
            raise SystemExit("exchange check failed: the reduced gradients are not the sum of the per-rank gradients")

    # forward-only frames/s of the whole job, measured (not extrapolated)
    try:
        line["render_fps"] = round(runner.measured_fps(min(4 * steps, 64) if args.impl == "ours" else min(steps, 8), eff_world), 2)
    except Exception as e:                                          # pragma: no cover
        line["render_fps"] = None
        line["render_fps_error"] = f"{type(e).__name__}: {e}"

    if args.impl == "ours" and not args.no_profile:
        fp32_peak, fp32_derived, fp32_src = measured_fp32_peak(dev, peaks)
        n_prof = min(steps, 4)
        prof = profile_pass(runner, n_prof, warmup)
        roof, table, per_kernel = rooflines(prof, stats, wl, n_prof, runner.local, peaks, fp32_peak, fp32_src)
        line["roofline"] = roof
        line["rooflines"] = table
        line["peaks"] = {"hbm_gbs": peaks.get("hbm_gbs"), "fp32_tflops_measured": round(fp32_peak, 2),
                         "fp32_tflops_derived": round(fp32_derived, 2), "source": fp32_src}

    if rank == 0 and eff_world == 1 and not args.no_secondary and args.workload == HEADLINE:
        sec = []
        del runner
        torch.cuda.empty_cache()
        for name in SECONDARY:
            try:
                sec.append(run_single(name, args.impl, dev, steps, warmup, flush))
            except Exception as e:                                  # pragma: no cover
                sec.append({"workload": name, "error": f"{type(e).__name__}: {e}"})
        line["secondary"] = sec

    if args.impl == "ours" and rank == 0 and eff_world == 1 and not args.no_cpu_baseline:
        try:
            line["cpu_baseline"] = cpu_deformation_baseline(dev)
        except Exception as e:                                      # pragma: no cover
            line["cpu_baseline"] = {"value": None, "unit": "ms", "cores": 1, "kind": "port", "sample": f"failed: {e}"}
    if rank == 0:
        print(json.dumps(line))
    if world > 1 and args.impl == "ours":
        import torch.distributed as dist
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
