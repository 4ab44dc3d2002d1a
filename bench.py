#!/usr/bin/env python
"""Headline benchmark: one train iteration of the per-frame render hot path (keypoint deformation forward ->
rasterize forward -> backward through both), on BASELINE.json configs[1]: D-NeRF-shaped synthetic scene, 800x800,
~100k Gaussians, 512 keypoints, nearest_num 6, SH degree 3.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

One JSON line on stdout (rank 0).  A "step" is one (camera, timestep) sample per rank: weak scaling, samples are
sharded across ranks and the parameter gradients are summed with one NCCL all-reduce per step.

  value        ms per train iteration, whole job (step time / samples processed in it), inputs resident in HBM
  e2e          same, through the public API with the step's camera + target image coming from pinned host memory and
               the loss scalar read back every step
  roofline     the dominant kernel of the step, timed with CUDA events on its stream in a second pass over the
               same steps (gp_profile_enable), against MEASURED_PEAKS.json
  cpu_baseline the CPU oracle (oracle/) on the same workload, bounded sample, rank 0 / N=1 only
  --impl reference   the reference's own CUDA path (oracle/_ref/*.so + its torch ops, oracle/ref_pipeline.py)
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
    "c5_2m_800x800": dict(N=2_000_000, W=800, H=800, Kp=512, K=6, shape="dnerf"),
    "tiny": dict(N=5_000, W=160, H=112, Kp=64, K=6, shape="dnerf"),
}
ITER = 50_000          # > second_stage_iteration + xyz_noise_iteration: keypoint stage, noise factor 0


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c2_dnerf_100k_800x800", choices=list(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-profile", action="store_true")
    ap.add_argument("--ncams", type=int, default=16)
    ap.add_argument("--eager", action="store_true", help="ours: do not capture the iteration into a CUDA graph")
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
    def __init__(self, wl, dev, ncams, world):
        from gaussianprediction_b200.deform import KeypointDeformation
        from gaussianprediction_b200.renderer import SceneModel
        self.wl, self.dev, self.world = wl, dev, world
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
        self.reducer = None
        self.h2d_bytes = grads[0].numel() * 4 + (16 + 16 + 3 + 1) * 4
        self.d2h_bytes = 4

    def step(self, i, host_inputs=False):
        from gaussianprediction_b200.renderer import render
        i = i * self.world + self.rank          # samples r, r + W, r + 2W, ... of the global sequence (parallel.shard_samples)
        k = i % len(self.cams_dev)
        if host_inputs:
            cam = cam_to(self.cams_host[k], self.dev, non_blocking=True)
            t = self.times_host[k:k + 1].to(self.dev, non_blocking=True)
            G = self.grads_host[i % len(self.grads_host)].to(self.dev, non_blocking=True)
        else:
            cam, t, G = self.cams_dev[k], self.times_dev[k], self.grads_dev[i % len(self.grads_dev)]
        for p in self.params:
            p.grad = None
        pkg = render(cam, self.pc, None, self.bg, time=t, it=ITER, capacity=self.capacity)
        loss = (pkg["render"] * G).sum()
        loss.backward()
        if self.world > 1:
            self.allreduce()
        return loss

    def enable_graph(self):
        """Capture the whole iteration into one CUDA graph (gaussianprediction_b200.pipeline.TrainStep)."""
        from gaussianprediction_b200.pipeline import RenderStep, TrainStep
        reducer = None
        if self.world > 1:
            from gaussianprediction_b200.parallel import FlatGradAllReduce
            # only parameters that receive a gradient on this path take part in the exchange (known after the eager pass)
            d = self.pc.deformation
            reducer = FlatGradAllReduce([p for p in self.params if p.grad is not None],
                                        first=[self.pc._features_dc, self.pc._features_rest],
                                        second=[d._xyz, d._rotation, d._scaling, d._opacity] + list(d.weights_model.parameters()))
        self.graph_reducer = reducer
        self.train_step = TrainStep(self.pc, self.cams_dev[0], self.capacity, bg=self.bg, iteration=ITER, reducer=reducer)
        self.render_step = RenderStep(self.pc, self.cams_dev[0], self.capacity, bg=self.bg, iteration=ITER)

    def _host_sample(self, i):
        i = i * self.world + self.rank
        k = i % len(self.cams_dev)
        return self.cams_host[k], self.times_host[k:k + 1], self.grads_host[i % len(self.grads_host)]

    def step_graph(self, i, host_inputs=False):
        if host_inputs:
            # end-to-end: every sample comes from pinned host memory; sample i+1 is copied while step i computes
            if getattr(self, "_prefetched", None) != i:
                self.train_step.prefetch(*self._host_sample(i))
            loss = self.train_step.step_prefetched()
            self.train_step.prefetch(*self._host_sample(i + 1))
            self._prefetched = i + 1
            return loss
        i = i * self.world + self.rank
        k = i % len(self.cams_dev)
        if host_inputs:
            cam, t, G = self.cams_host[k], self.times_host[k:k + 1], self.grads_host[i % len(self.grads_host)]
        else:
            cam, t, G = self.cams_dev[k], self.times_dev[k], self.grads_dev[i % len(self.grads_dev)]
        loss = self.train_step(cam, t, G)       # with world > 1 the step includes the gradient all-reduce
        return loss

    def allreduce(self):
        if self.reducer is None:
            from gaussianprediction_b200.parallel import GroupedGradAllReduce
            self.reducer = GroupedGradAllReduce(self.params)
        self.reducer.reduce()

    def eval_fps(self, steps, graph=False):
        """Forward-only render FPS in the reference's own span (eval.py:210-214: sync, render(), sync)."""
        from gaussianprediction_b200.renderer import render
        ts = []
        with torch.no_grad():
            for i in range(steps + 3):
                k = i % len(self.cams_dev)
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                if graph:
                    self.render_step(self.cams_dev[k], self.times_dev[k])
                else:
                    render(self.cams_dev[k], self.pc, None, self.bg, time=self.times_dev[k], it=ITER, capacity=self.capacity)
                torch.cuda.synchronize()
                if i >= 3:
                    ts.append(time.perf_counter() - t0)
        return 1.0 / float(np.mean(ts))


class RefRunner:
    """The reference's CUDA path (see oracle/ref_pipeline.py)."""

    def __init__(self, wl, dev, ncams, world):
        from oracle.ref_pipeline import RefKeypointModel
        self.wl, self.dev, self.world = wl, dev, 1
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
        self.h2d_bytes = grads[0].numel() * 4 + (16 + 16 + 3 + 1) * 4
        self.d2h_bytes = 4

    def calibrate(self):
        return None

    def step(self, i, host_inputs=False):
        from oracle.ref_pipeline import ref_render
        k = i % len(self.cams_dev)
        if host_inputs:
            cam = cam_to(self.cams_host[k], self.dev, non_blocking=True)
            t = self.times_host[k:k + 1].to(self.dev, non_blocking=True)
            G = self.grads_host[i % len(self.grads_host)].to(self.dev, non_blocking=True)
        else:
            cam, t, G = self.cams_dev[k], self.times_dev[k], self.grads_dev[i % len(self.grads_dev)]
        for p in self.params:
            p.grad = None
        pkg = ref_render(cam, self.pc, self.bg, t, ITER)
        loss = (pkg["render"] * G).sum()
        loss.backward()
        return loss

    def eval_fps(self, steps):
        from oracle.ref_pipeline import ref_render
        ts = []
        with torch.no_grad():
            for i in range(steps + 3):
                k = i % len(self.cams_dev)
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                ref_render(self.cams_dev[k], self.pc, self.bg, self.times_dev[k], ITER)
                torch.cuda.synchronize()
                if i >= 3:
                    ts.append(time.perf_counter() - t0)
        return 1.0 / float(np.mean(ts))


# ------------------------------------------------------------------------------------------------ timing

def timed_steps(runner, steps, warmup, world, host_inputs, flush, graph=False):
    """W untimed steps, then exactly K steps; every step is bracketed by CUDA events on the current stream, L2 is
    flushed (256 MiB write) between steps outside the events; returns (sum of per-step device ms, wall ms)."""
    import torch.distributed as dist
    step = runner.step_graph if graph else runner.step
    for i in range(warmup):
        loss = step(i, host_inputs)
        if host_inputs:
            loss.item()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    t0 = time.perf_counter()
    for i in range(steps):
        flush.add_(1.0)                      # evict L2 (126 MB) between timed iterations
        evs[i][0].record()
        loss = step(warmup + i, host_inputs)
        if host_inputs:
            loss.item()                      # D2H read of the step's result inside the timed region
        evs[i][1].record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    wall_ms = (time.perf_counter() - t0) * 1e3
    dev_ms = sum(a.elapsed_time(b) for a, b in evs)
    return dev_ms, wall_ms


def max_over_ranks(x, world, dev):
    if world == 1:
        return x
    import torch.distributed as dist
    t = torch.tensor([x], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def profile_pass(runner, steps, warmup):
    """Second pass over the same steps with per-kernel CUDA events (same stream); returns {kernel: (launches, ms)}."""
    import ctypes
    from gaussianprediction_b200 import _lib
    L = _lib.lib()
    L.gp_profile_enable(1)
    for i in range(steps):
        runner.step(warmup + i, False)
    torch.cuda.synchronize()
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
    out["max_R"] = max_r
    return out


# dram__bytes_read.sum + dram__bytes_write.sum per launch from the committed ncu --set full captures
# (profiles/r1_ncu_full.md), C2 workload only; other workloads report null
NCU_DRAM_BYTES = {"c2_dnerf_100k_800x800": {"gp_render_bwd_kernel": 24.19e6, "gp_render_fwd_kernel": 6.97e6}}


def roofline_for(prof, stats, wl, steps, peaks):
    """Dominant kernel of the step and its achieved rate.  Algorithmic work per launch:
    HBM kernels in bytes (SURVEY.md §8d per-unit figures), blend kernels in FP32 operations per evaluated
    pixel x Gaussian pair (18 forward, 45 backward, SURVEY.md §8d hand counts; pair count = sum of n_contrib)."""
    if not prof:
        return None, {}
    per = {k: v[1] / max(v[0], 1) for k, v in prof.items()}            # ms per launch
    tot = {k: v[1] / steps for k, v in prof.items()}                    # ms per step
    N, R, pairs = wl["N"], stats["R"], stats["pairs"]
    hbm = peaks.get("hbm_gbs", 6650.0)
    fp32_peak = 148 * 128 * 2 * peaks.get("sm_max_mhz", 1965.0) * 1e6 / 1e12   # TFLOP/s, derived (not measured)
    work = {
        "gp_preprocess_kernel": ("hbm", 311.0 * N),
        "gp_preprocess_bwd_kernel": ("hbm", 535.0 * N),
        "gp_rs_scatter_kernel": ("hbm", None),
        "gp_render_fwd_kernel": ("fp32", 18.0 * pairs),
        "gp_render_bwd_kernel": ("fp32", 45.0 * pairs),
        "gp_knn_kernel": ("fp32", 2.0 * 2 * N * wl["Kp"] * 35),
        "gp_blend_fwd_kernel": ("hbm", 128.0 * N),
        "gp_blend_bwd_kernel": ("hbm", 204.0 * N),
    }
    top = max(tot, key=tot.get)
    table = {k: round(v, 4) for k, v in sorted(tot.items(), key=lambda kv: -kv[1])}
    kind, w = work.get(top, ("hbm", None))
    if w is None:
        return {"kernel": top, "bound": kind, "achieved": None, "peak": None, "unit": None, "frac": None, "traffic": None,
                "ms_per_launch": per[top], "share_of_kernel_time": tot[top] / sum(tot.values())}, table
    sec = per[top] * 1e-3
    if kind == "hbm":
        ach, peak, unit = w / sec / 1e9, hbm, "GB/s"
    else:
        ach, peak, unit = w / sec / 1e12, fp32_peak, "TFLOP/s"
    return {"kernel": top, "bound": kind, "achieved": round(ach, 3), "peak": round(peak, 3), "unit": unit,
            "frac": round(ach / peak, 4), "traffic": NCU_DRAM_BYTES.get(wl.get("name"), {}).get(top),
            "traffic_unit": "bytes/launch (ncu dram read+write, profiles/r1_ncu_full.md)",
            "ms_per_launch": round(per[top], 4),
            "share_of_kernel_time": round(tot[top] / sum(tot.values()), 4),
            "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if kind == "hbm" else
            "derived 148 SM x 128 lanes x 2 x sm_max_mhz (FP32 FMA pipe; not in MEASURED_PEAKS.json)"}, table


# ------------------------------------------------------------------------------------------------ CPU baseline

def cpu_train_iter(wl, sample_note=True):
    """One train iteration of the same workload on the host: deformation in torch CPU (oracle/deform_oracle.py, dense
    formulation, autograd) + rasterizer forward/backward in the single-threaded C oracle (oracle/raster_oracle.c)."""
    from oracle import deform_oracle as do
    from oracle import raster_oracle as ro
    from gaussianprediction_b200.deform import DeformableField
    gs, kp, kf, logits, cams, times, grads = build_inputs(wl, 1)
    a = model_args()
    torch.manual_seed(2024)
    net = DeformableField(2 * a.time_freq + 6 * a.xyz_freq + a.feature_dim, 7, a.d, a.w)
    sd = dict(net.state_dict())
    cam = cams[0]
    leaf = {k: v.clone().requires_grad_(True) for k, v in dict(xyz=gs.xyz, rot=gs.rotation, sc=gs.scaling, op=gs.opacity,
                                                              kp=kp, kf=kf, logits=logits).items()}
    shs = torch.cat([gs.features_dc, gs.features_rest], 1).contiguous()
    t0 = time.perf_counter()
    p1, p2 = do.hybrid_points(gs.xyz, gs.motion_feature, kp, kf, a.feature_amplify)
    # kNN: the reference's own single-threaded CPU brute force on a bounded sample of the queries (it is linear in
    # their number), scaled to all of them; the indices used downstream come from the full-size numpy oracle only
    # when the sample covers everything.
    n_q = p1.shape[0]
    n_s = min(n_q, 20_000)
    scale = n_q / n_s
    t_k0 = time.perf_counter()
    try:
        from oracle import ref_cuda
        _, idx_s = ref_cuda.frnn_bf_cpu(p1[None, :n_s].contiguous(), p2[None], wl["K"], 1e8)   # the reference's own CPU kNN
        idx_s = idx_s[0]
        knn_kind = "reference frnn_bf_cpu"
    except Exception:
        _, idx_s = do.knn(p1[:n_s].numpy(), p2.numpy(), wl["K"], fma=False)
        idx_s = torch.from_numpy(idx_s)
        knn_kind = "oracle knn (numpy)"
    t_knn = (time.perf_counter() - t_k0) * scale
    if n_s < n_q:      # neighbours of the remaining queries: any valid index set keeps the downstream work identical
        idx = torch.cat([idx_s, idx_s.repeat((n_q + n_s - 1) // n_s, 1)[:n_q - n_s]], 0)
    else:
        idx = idx_s
    knn_kind += f" on {n_s} of {n_q} queries, scaled x{scale:.0f}"
    t_rest0 = time.perf_counter()
    xyz_t, q_t, s_t, o_t, _, _ = do.keypoint_forward(times[0:1], leaf["xyz"], leaf["rot"], leaf["sc"], leaf["op"], leaf["kp"],
                                                      leaf["kf"], leaf["logits"], idx, sd, norm_rotation=a.norm_rotation)
    t_def = time.perf_counter() - t_rest0
    o = ro.forward(xyz_t.detach().numpy(), o_t.detach().numpy(), np.zeros(3, np.float32), cam.world_view_transform.numpy(),
                   cam.full_proj_transform.numpy(), cam.camera_center.numpy(), wl["W"], wl["H"], cam.tanfovx, cam.tanfovy,
                   shs=shs.numpy(), scales=s_t.detach().numpy(), rotations=q_t.detach().numpy())
    g = ro.backward(o, grads[0].numpy())
    torch.autograd.backward([xyz_t, q_t, s_t, o_t], [torch.from_numpy(g["means3D"]), torch.from_numpy(g["rotations"]),
                                                     torch.from_numpy(g["scales"]), torch.from_numpy(g["opacity"])])
    total = (time.perf_counter() - t_rest0) + t_knn
    return total * 1e3, dict(knn_ms=t_knn * 1e3, deform_fwd_ms=t_def * 1e3, knn=knn_kind)


# ------------------------------------------------------------------------------------------------ main

def main():
    args = parse()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    wl = dict(WORKLOADS[args.workload], name=args.workload)
    steps, warmup = args.steps, max(args.warmup, 3)

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
        L = _lib.lib()                                             # raises if the CUDA library is missing
        runner = OurRunner(wl, dev, args.ncams, eff_world)
    else:
        from oracle import ref_cuda
        for m in ("ref_dgr", "ref_frnn", "ref_prefix_sum"):
            if not ref_cuda.available(m):
                print(json.dumps({"impl": "reference", "unavailable": f"oracle/_ref/{m}.so not built"}))
                return 0
        runner = RefRunner(wl, dev, args.ncams, 1)

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
    if args.impl == "ours":
        # eager pass first: it counts the kernels of one step (a graph replay re-issues exactly those launches)
        launches0 = _lib.lib().gp_launch_count()
        eager_dev_ms, _ = timed_steps(runner, steps, warmup, eff_world, False, flush)
        launches = _lib.lib().gp_launch_count() - launches0
        launches = launches * steps // (steps + warmup)
        eager_ms = max_over_ranks(eager_dev_ms, eff_world, dev) / steps / eff_world
        if not args.eager:
            try:
                runner.enable_graph()
                use_graph, graph_note = True, "whole iteration replayed as one CUDA graph (pipeline.TrainStep)"
            except Exception as e:                                  # pragma: no cover
                graph_note = f"eager launches (graph capture failed: {type(e).__name__}: {e})"
    else:
        launches = 0
    dev_ms, wall_ms = timed_steps(runner, steps, warmup, eff_world, False, flush, graph=use_graph)
    e2e_dev_ms, e2e_wall_ms = timed_steps(runner, steps, warmup, eff_world, True, flush, graph=use_graph)
    clocks = sampler.stop() if rank == 0 else None
    if use_graph:
        R_last, over = runner.train_step.overflowed()
        if over:
            raise SystemExit("tile-instance capacity overflowed during the timed region: result invalid")

    dev_ms = max_over_ranks(dev_ms, eff_world, dev)
    e2e_dev_ms = max_over_ranks(e2e_dev_ms, eff_world, dev)
    ms_per_step = dev_ms / steps
    e2e_ms_per_step = e2e_dev_ms / steps
    value = ms_per_step / eff_world                                # whole job: ms per (camera, time) sample
    e2e_value = e2e_ms_per_step / eff_world

    fps = runner.eval_fps(min(steps, 20), graph=True) if use_graph else runner.eval_fps(min(steps, 20))

    line = {
        "metric": METRIC, "value": round(value, 4), "unit": "ms/iter", "n_gpus": eff_world if args.impl == "ours" else args.gpus,
        "steps": steps, "warmup": warmup, "ms_per_step": round(ms_per_step, 4), "higher_is_better": False,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": args.workload, "gaussians": wl["N"], "image": f'{wl["W"]}x{wl["H"]}', "keypoints": wl["Kp"],
                   "nearest_num": wl["K"], "knn_type": "hybird", "sh_degree": 3, "samples_per_step_per_gpu": 1,
                   "l2": "flushed between timed steps (256 MiB write, outside the events)",
                   "timing": "sum of per-step CUDA-event times, max over ranks"},
        "e2e": {"value": round(e2e_value, 4), "unit": "ms/iter", "h2d_bytes_per_step": runner.h2d_bytes,
                "d2h_bytes_per_step": runner.d2h_bytes},
        "gpu_launches": int(launches),
        "train_iters_per_s": round(1e3 / value, 2), "render_fps": round(fps * eff_world, 2),
        "wall_ms_per_step": round(wall_ms / steps, 4),
    }
    line["config"]["launch"] = graph_note
    if eager_ms is not None:
        line["eager_ms_per_iter"] = round(eager_ms, 4)
    if args.impl == "reference":
        line["impl"] = "reference"
        line["n_gpus"] = args.gpus
        line["cpu_baseline"] = {"value": line["value"], "unit": "ms/iter", "cores": 0, "kind": "reference",
                                "sample": "the reference's CUDA kernels (oracle/_ref, sm_100) + its torch ops on one B200: "
                                          "north_star's bar; the rasterizer has no CPU implementation"}
        line["e2e"] = {"value": round(e2e_value, 4), "unit": "ms/iter", "h2d_bytes_per_step": runner.h2d_bytes,
                       "d2h_bytes_per_step": runner.d2h_bytes}
    if stats:
        line["config"].update(tile_instances=stats["R"], visible=stats["visible"], blend_pairs=stats["pairs"])
    if clocks is not None:
        line["clocks"] = clocks
    if args.impl == "ours" and not args.no_profile:
        prof = profile_pass(runner, min(steps, 10), warmup)
        roof, table = roofline_for(prof, stats, wl, min(steps, 10), peaks)
        line["roofline"] = roof
        line["kernel_ms_per_step"] = table
    if args.impl == "ours" and rank == 0 and eff_world == 1 and not args.no_cpu_baseline:
        small = dict(wl)
        try:
            ms, parts = cpu_train_iter(small)
            line["cpu_baseline"] = {"value": round(ms, 1), "unit": "ms/iter", "cores": 1, "kind": "port",
                                    "sample": f"1 train iteration of the same workload ({args.workload}) on the host: "
                                              f"kNN {parts['knn']} {parts['knn_ms']:.0f} ms, torch-CPU dense deformation "
                                              f"({torch.get_num_threads()} threads), C oracle rasterizer fwd+bwd (1 thread)"}
        except Exception as e:                                      # pragma: no cover
            line["cpu_baseline"] = {"value": None, "unit": "ms/iter", "cores": 1, "kind": "port", "sample": f"failed: {e}"}
    if rank == 0:
        print(json.dumps(line))
    if world > 1 and args.impl == "ours":
        import torch.distributed as dist
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
