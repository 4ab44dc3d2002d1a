"""Whole-iteration CUDA graphs for the frame loop.

The reference's train iteration (train.py:77-135) issues ~150 small torch/CUDA launches and one blocking D2H per
frame; at 100k Gaussians the GPU work is ~1.5 ms, so launch latency is a large share of the iteration.  Because this
implementation has no host round trip inside a frame (the tile-instance count stays on the device, see
rasterizer.py ``capacity=``), the complete step -- deformation forward, rasterizer forward, loss, rasterizer
backward, deformation backward -- can be captured once into a CUDA graph and replayed with new camera / time /
target tensors.  Kernels and results are the same as the eager path; only the launch mechanism changes.

A graph is tied to (image size, tan-fov, SH degree, parameter shapes, capacity): those are kernel arguments by value.
"""
import os

import torch
from torch.nn.utils.stateless import _reparametrize_module

from . import _lib
from ._sinks import grad_sinks
from .rasterizer import export_state  # noqa: F401  (kept importable for diagnostics)
from .renderer import render


# share of the SH bucket exchanged right after the rasterizer backward in pipelined mode (the rest follows the small
# buckets).  Measured on the 2 M-Gaussian job at 8 ranks: 0 is best (3.65 vs 3.80 ms per step with 0.25) -- under the
# blend backward the head runs at a third of its speed and delays the buckets the next step needs first
EARLY_FRACTION = float(os.environ.get("GP_EXCHANGE_EARLY_FRACTION", "0"))

RESIDENT_EXCHANGE = os.environ.get("GP_EXCHANGE_RESIDENT", "1") != "0"   # A/B knob (see TrainStep._replay)

TRACE = None     # diagnostics (tools/timeline_probe.py): a list that receives (label, timing event, host time) marks


def _mark(label, stream=None):
    if TRACE is not None:
        import time
        e = torch.cuda.Event(enable_timing=True)
        e.record(stream if stream is not None else torch.cuda.current_stream())
        TRACE.append((label, e, time.perf_counter()))


class _StaticCamera:
    def __init__(self, cam, device):
        self.image_width, self.image_height = int(cam.image_width), int(cam.image_height)
        self.FoVx, self.FoVy = float(cam.FoVx), float(cam.FoVy)
        # one packed buffer (view 16 | proj 16 | centre 3 | time 1) so that a prefetched sample lands with ONE small copy
        self.packed = torch.empty(36, device=device)
        self.world_view_transform = self.packed[0:16].view(4, 4)
        self.full_proj_transform = self.packed[16:32].view(4, 4)
        self.camera_center = self.packed[32:35]
        self.time = self.packed[35:36]

    def load(self, cam, non_blocking=True):
        assert int(cam.image_width) == self.image_width and int(cam.image_height) == self.image_height
        assert abs(float(cam.FoVx) - self.FoVx) < 1e-12 and abs(float(cam.FoVy) - self.FoVy) < 1e-12, \
            "a captured step is tied to the camera intrinsics; build another TrainStep for other intrinsics"
        self.world_view_transform.copy_(cam.world_view_transform, non_blocking=non_blocking)
        self.full_proj_transform.copy_(cam.full_proj_transform, non_blocking=non_blocking)
        self.camera_center.copy_(cam.camera_center, non_blocking=non_blocking)


class TrainStep:
    """``loss = step(cam, t, target)``: render(cam, pc, time=t) -> loss_fn(image, target) -> backward, as one graph.

    After the call ``p.grad`` of every parameter of ``pc`` holds the gradients of this sample (static tensors,
    overwritten by the next call).  ``loss_fn`` defaults to ``(image * target).sum()`` (an upstream-gradient probe);
    any capturable torch function of (image, target) works.  ``capacity`` bounds the tile instances of a frame;
    ``overflowed()`` tells whether the last frame exceeded it (that frame's image and gradients are then invalid).
    """

    def __init__(self, pc, cam, capacity, bg=None, iteration=50_000, loss_fn=None, warmup=3, reducer=None,
                 capture_collective=False, accumulate_into=None, split_forward=False, stats=None):
        """``reducer``: a ``parallel.FlatGradAllReduce`` over the same parameters.  The backward kernels then write
        the parameter gradients straight into its flat buffer (``_sinks``: no zero fill, no accumulation pass, no
        packing copies) and the step is captured as one CUDA graph per reducer bucket (up to three, see
        ``_stage_fns``): each graph ends where the gradients of its bucket are final, its all-reduce is issued right
        after the replay and runs over NVLink while the next graph computes.  ``capture_collective=True`` records ONE
        all-reduce into a single graph instead; EXPERIMENTAL -- with torch 2.11 / NCCL 2.28 on the B200 test box
        that capture hung, so it is off by default and not used by bench.py.

        ``split_forward``: the deformation forward becomes two graphs of its own in front of the others, [neighbour
        search + keypoint MLP] and [blend].  The first reads only xyz / keypoints / MLP weights, the second the other
        per-Gaussian parameters and the blend logits, and neither reads an SH coefficient -- so
        ``_replay(wait_before_blend=..., wait_before_raster=...)`` can let the PREVIOUS step's exchange of those
        groups (and the optimizer updates that depend on it) run underneath them.

        ``accumulate_into``: a second ``FlatGradAllReduce`` with the same layout.  Every stage graph then ends with
        ``accumulate_into.bucket[k] += reducer.bucket[k]``: this step object serves samples 2.. of a gradient-
        accumulation step (``BatchStep``), the running sum lives in the other buffer and is what gets exchanged."""
        self._ctor = dict(pc=pc, cam=cam, bg=bg, iteration=iteration, loss_fn=loss_fn, warmup=warmup, reducer=reducer,
                          capture_collective=capture_collective, accumulate_into=accumulate_into,
                          split_forward=split_forward, stats=stats)
        self._stats = stats      # (loss_sum, radii, first): running loss sum / radii max of a BatchStep, updated in-graph
        self.exchange_radii = False
        self.split_forward = bool(split_forward)
        self.deferred_blend = self.deferred_raster = None
        self.pc = pc
        self.reducer = reducer
        self.accumulate_into = accumulate_into
        # A/B knob: 0 = every gradient goes through the scratch buffer and one add per stage
        self.accumulate_in_kernels = os.environ.get("GP_ACCUMULATE_IN_KERNELS", "1") != "0"
        if accumulate_into is not None:
            assert reducer is not None and accumulate_into.flat.shape == reducer.flat.shape and \
                accumulate_into.bucket_ranges == reducer.bucket_ranges, "accumulate_into must mirror the reducer"
        self.capture_collective = bool(capture_collective and reducer is not None)
        self.params = [p for p in pc.parameters() if p.requires_grad]
        if reducer is not None:
            # the reducer may cover a subset (parameters that never receive a gradient on this path, e.g. the
            # per-Gaussian motion_feature in the keypoint stage, need no exchange)
            ids = {id(p) for p in self.params}
            assert all(id(p) in ids for p in reducer.params), "reducer parameters must belong to pc"
        self._reduced = {id(p) for p in reducer.params} if reducer is not None else set()
        dev = self.params[0].device
        self.dev = dev
        self.capacity = int(capacity)
        self.iteration = int(iteration)
        self.bg = torch.zeros(3, device=dev) if bg is None else bg
        self.loss_fn = loss_fn or (lambda image, target: (image * target).sum())
        self.cam = _StaticCamera(cam, dev)
        self.cam.load(cam, non_blocking=False)
        self.t = self.cam.time
        self.t.zero_()
        self.target = torch.zeros(3, self.cam.image_height, self.cam.image_width, device=dev)
        self.loss = None
        self.pkg = None
        _lib.lib()
        self.max_stages = int(os.environ.get("GP_TRAINSTEP_STAGES", "3"))   # A/B knob: 2 merges the last two stages
        self._stages = self._stage_fns() if self.two_stage else None
        self._cut1 = self._cut2 = None
        self.packed = self.in_place = 0
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            for _ in range(max(warmup, 1)):       # also resolves lazy inits (cuBLAS handles, cached tables)
                if accumulate_into is not None:   # the check below compares ONE sample's gradients
                    accumulate_into.flat.zero_()
                self._body()
            if self.two_stage:
                self._verify_staging()
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        if self.reducer is None:
            for p in self.params:
                p.grad = None
        self.graph = torch.cuda.CUDAGraph()
        self.stage_graphs = None
        if self.two_stage:
            self.stage_graphs = []
            for k, fn in enumerate(self._stages):
                g = self.graph if k == 0 else torch.cuda.CUDAGraph()
                if k == 0 and getattr(self, "_n_pre", 0) == 2:
                    g2 = torch.cuda.CUDAGraph()
                    self._capture_split_at_blend(g, g2, fn)
                    self.stage_graphs += [g, g2]
                    continue
                with torch.cuda.graph(g, **({} if k == 0 else dict(pool=self.graph.pool()))):
                    fn()
                self.stage_graphs.append(g)
        else:
            with torch.cuda.graph(self.graph):
                self._body()
        from . import rasterizer as _r
        self._state = _r.LAST_STATE          # workspaces of the captured frame (static addresses)
        self.grads = [p.grad for p in self.params]   # the tensors the graph writes; re-attached after every replay
        torch.cuda.synchronize(dev)

    def _capture_split_at_blend(self, g_a, g_b, fn):
        """Capture ``fn`` (the deformation forward) into TWO graphs cut right before the blend kernel: the neighbour search
        and the keypoint MLP read only xyz / keypoints / MLP weights, the blend reads the other per-Gaussian parameters and
        the blend logits -- so the exchange (and optimizer update) of those may still be running during the first graph."""
        import gc
        model = getattr(self.pc, "deformation", self.pc)
        gc.collect()
        torch.cuda.synchronize(self.dev)
        cs = torch.cuda.Stream(device=self.dev)
        cs.wait_stream(torch.cuda.current_stream(self.dev))
        state = {"cut": False}

        def hook():
            g_a.capture_end()
            g_b.capture_begin(pool=g_a.pool())
            state["cut"] = True

        with torch.cuda.stream(cs):
            g_a.capture_begin()
            model._pre_blend_hook = hook
            try:
                fn()
            finally:
                model._pre_blend_hook = None
            (g_b if state["cut"] else g_a).capture_end()
            if not state["cut"]:                  # a model without the hook: the second graph stays empty
                g_b.capture_begin(pool=g_a.pool())
                g_b.capture_end()
        torch.cuda.current_stream(self.dev).wait_stream(cs)

    def overflowed(self):
        """(R, flag) of the last replay, read from the device header; if flag is set the frame was not rendered."""
        hdr = self._state.geom[:16].view(torch.int32).cpu()
        return int(hdr[_lib.HDR_NUM_RENDERED]), bool(int(hdr[_lib.HDR_OVERFLOW]))

    @property
    def two_stage(self):
        return self.reducer is not None and not self.capture_collective

    # The staged backward.  Stage k ends where the gradients of the reducer's bucket k are final:
    #   0: forward, loss, backward through the rasterizer        -> SH coefficients
    #   1: blend backward                                        -> per-Gaussian deformation parameters, blend logits
    #   2: keypoint MLP backward                                 -> keypoints, MLP weights
    # With fewer buckets (or a model that does not expose its keypoint deltas) the later stages merge.
    def _stage_fns(self):
        r = self.reducer
        nb = len(r.bucket_ranges)
        model = getattr(self.pc, "deformation", None)
        can_cut_mlp = nb >= 3 and hasattr(model, "keep_kp_deltas") and self.max_stages >= 3

        def leaves(k):
            lo, hi, _, _ = r.bucket_ranges[k]
            return r.params[lo:hi]

        def finish(k, last):
            lo, hi, e0, _ = r.bucket_ranges[k]
            hi = len(r.params) if last else hi
            acc = self.accumulate_into
            if acc is None:
                self.packed += r.pack(lo, hi)
                return
            # Running sum of the step's samples (captured with the stage).  Gradients a kernel has already ADDED to the
            # sum (accumulating sinks) are done; the others are packed into the scratch buffer and added run by run.
            runs = []
            for i in range(lo, hi):
                p, av = r.params[i], acc.views[i]
                if p.grad is not None and p.grad.data_ptr() == av.data_ptr():
                    p.grad = av
                    self.in_place += 1
                    continue
                self.packed += r.pack(i, i + 1)
                o0 = r.views[i].storage_offset() - r.flat.storage_offset()
                o1 = o0 + r.views[i].numel()
                if runs and runs[-1][1] >= o0 - 4:      # (alignment padding between two views is zero on both sides)
                    runs[-1][1] = o1
                else:
                    runs.append([o0, o1])
            for o0, o1 in runs:
                acc.flat[o0:o1].add_(r.flat[o0:o1])

        split = self.split_forward and nb > 1

        def forward(phase):
            if phase != "raster":
                for p in self.params:
                    p.grad = None
                self.packed = 0               # diagnostics: gradients that did NOT arrive through a sink
                self.in_place = 0             # ... and gradients a kernel added to the running sum itself
                if can_cut_mlp:
                    model.keep_kp_deltas = True
            acc = self.accumulate_into
            in_place = zip(r.params, acc.views) if (acc is not None and self.accumulate_in_kernels) else ()
            with grad_sinks(zip(r.params, r.views), accumulate=in_place):   # looked up while the autograd graph is built
                self._render_aliased(phase)
            if phase != "raster" and can_cut_mlp:
                model.keep_kp_deltas = False
                self._cut2, model._kp_deltas = model._kp_deltas, None

        def stage_pre():
            forward("deform")

        def stage0():
            forward("raster" if split else "all")
            if self._stats is not None:           # the accumulation step's running sums, captured with the stage
                loss_sum, radii, first = self._stats
                if first:
                    loss_sum.copy_(self.loss.detach())
                    radii.copy_(self.pkg["radii"])
                else:
                    loss_sum.add_(self.loss.detach())
                    torch.maximum(radii, self.pkg["radii"], out=radii)
            deformed = [t for t in self.pkg["deformed"] if t.requires_grad]
            if nb == 1:
                self._grads_into([self.loss], None, self.params, viewspace=True)
                finish(0, True)
                return
            early = leaves(0)
            outs = torch.autograd.grad(self.loss, deformed + self._aliases_of(early) + [self.pkg["viewspace_points"]],
                                       allow_unused=True)
            self._cut1 = [(t, g) for t, g in zip(deformed, outs[:len(deformed)]) if g is not None]
            for p, g in zip(early, outs[len(deformed):len(deformed) + len(early)]):
                p.grad = g
            self.viewspace_grad = outs[-1]
            finish(0, False)

        def stage1():
            ts, gs = [t for t, _ in self._cut1], [g for _, g in self._cut1]
            cut = [t for t in (self._cut2 or ()) if t is not None and t.requires_grad] if can_cut_mlp else []
            if not cut:
                if ts:
                    done = {id(p) for p in leaves(0)}
                    self._grads_into(ts, gs, [p for p in self.params if id(p) not in done])
                finish(1, True)
                return
            # leaves of the LAST bucket whose gradient is already produced here (e.g. xyz, when the caller put it there
            # because the next step needs it first) are taken now; stage 2 asks only for what is still missing
            mine = leaves(1) + leaves(2)
            outs = torch.autograd.grad(ts, cut + self._aliases_of(mine), grad_outputs=gs, allow_unused=True)
            self._cut2 = [(t, g) for t, g in zip(cut, outs[:len(cut)]) if g is not None]
            for p, g in zip(mine, outs[len(cut):]):
                p.grad = g
            finish(1, False)

        def stage2():
            if self._cut2:
                done = {id(p) for p in leaves(0) + leaves(1)} | {id(p) for p in leaves(2) if p.grad is not None}
                self._grads_into([t for t, _ in self._cut2], [g for _, g in self._cut2],
                                 [p for p in self.params if id(p) not in done])
            self._cut2 = None
            finish(2, True)

        self._n_pre = 2 if split else 0          # [search + keypoint MLP | blend] in front of the rasterizer stages
        pre = [stage_pre] if split else []
        if nb == 1:
            return [stage0]
        if not can_cut_mlp:
            return pre + [stage0, stage1]
        return pre + [stage0, stage1, stage2]

    def _render_aliased(self, phase="all"):
        """``render()`` + loss on fresh leaf ALIASES of the parameters (same storage, new autograd identity).  A parameter's
        AccumulateGrad node is bound to the stream it was first used on and stays alive as long as any earlier autograd
        graph does (a loss the caller still holds, say); the engine would then route this backward through that stream
        -- the legacy default stream, typically -- and break the capture.  Aliases made here belong to this stream."""
        if phase != "raster":
            named = list(self.pc.named_parameters())
            self._aliases = {n: p.detach().requires_grad_(p.requires_grad) for n, p in named}
            self._alias = {id(p): self._aliases[n] for n, p in named}
        with _reparametrize_module(self.pc, self._aliases):
            if phase == "deform":             # the deformation forward alone (``split_forward``)
                self._deformed = tuple(self.pc(self.t, self.iteration))
                return
            self.pkg = render(self.cam, self.pc, None, self.bg, time=self.t, it=self.iteration, capacity=self.capacity,
                              deformed=self._deformed if phase == "raster" else None)
            self.loss = self._flag_overflow(self.loss_fn(self.pkg["render"], self.target))
        self._deformed = None

    def _flag_overflow(self, loss):
        """Overflow policy, device side: a frame whose tile instances exceed ``capacity`` is not rendered (the image is
        the empty one, every gradient zero) and its loss becomes NaN -- the caller, who reads the loss anyway
        (train.py:139), sees it without an extra synchronisation and calls ``grow()``."""
        from . import rasterizer as _r
        hdr = _r.LAST_STATE.geom[:16].view(torch.int32)
        if not hasattr(self, "_nan"):
            self._nan = torch.full((), float("nan"), device=self.dev)
        return torch.where(hdr[_lib.HDR_OVERFLOW] != 0, self._nan, loss)

    def grow(self, capacity=None):
        """Re-capture the step with room for more tile instances (default: 1.5 x the instance count of the frame that
        overflowed); returns the new capacity.  The static inputs (camera, time, target) are kept."""
        R, _ = self.overflowed()
        new = int(capacity) if capacity is not None else int(max(R, self.capacity) * 1.5) + 65536
        packed, target = self.cam.packed.clone(), self.target.clone()
        stage = getattr(self, "_stage", None)
        self.__init__(capacity=new, **self._ctor)
        self.cam.packed.copy_(packed)
        self.target.copy_(target)
        if stage is not None:
            del self._stage                     # staging slots are rebuilt against the new static tensors on next use
        return new

    def _aliases_of(self, params):
        return [self._alias[id(p)] for p in params]

    def _grads_into(self, outputs, grad_outputs, params, viewspace=False):
        """``p.grad = dL/dp`` through ``torch.autograd.grad`` rather than ``backward()``: no AccumulateGrad node runs, so a
        stale autograd graph the caller still holds (whose AccumulateGrad nodes are bound to another stream) cannot
        invalidate a capture, and a gradient written through a sink is adopted as is."""
        wanted = self._aliases_of(params) + ([self.pkg["viewspace_points"]] if viewspace else [])
        outs = torch.autograd.grad(outputs, wanted, grad_outputs=grad_outputs, allow_unused=True)
        for p, g in zip(params, outs):
            p.grad = g
        if viewspace:
            self.viewspace_grad = outs[-1]

    def _verify_staging(self):
        """The staged backward only asks autograd for the leaves of the bucket a stage closes.  A parameter that gets
        (part of) its gradient in an earlier stage than its bucket says would silently lose it, so the staged result
        of the warm-up sample is compared once with a plain ``loss.backward()`` before anything is captured."""
        staged = {id(p): (None if p.grad is None else p.grad.clone()) for p in self.params}
        for p in self.params:
            p.grad = None
        pkg = render(self.cam, self.pc, None, self.bg, time=self.t, it=self.iteration, capacity=self.capacity)
        self.loss_fn(pkg["render"], self.target).backward()
        names = {id(p): n for n, p in self.pc.named_parameters()}
        for p in self.reducer.params:
            a, b = staged[id(p)], p.grad
            if b is None:
                continue
            scale = float(b.abs().max())
            if a is None or float((a - b).abs().max()) > 1e-3 * max(scale, 1e-30):
                raise RuntimeError(f"TrainStep: the staged backward loses gradient of '{names.get(id(p), '?')}': put it "
                                   "in the reducer bucket of the stage that produces it (parallel.FlatGradAllReduce "
                                   "first=/second=) or leave it to the last bucket")
        for p in self.params:
            p.grad = None

    def _body(self):
        if self.two_stage:
            for fn in self._stages:
                fn()
            return
        if self.reducer is None:
            for p in self.params:
                p.grad = None
        else:
            self.reducer.flat.zero_()                 # autograd then accumulates in place into the flat views
            for p in self.params:
                if id(p) not in self._reduced:
                    p.grad = None
            for p, v in zip(self.reducer.params, self.reducer.views):
                p.grad = v
        if self.reducer is None:
            self._render_aliased()
            self._grads_into([self.loss], None, self.params, viewspace=True)
        else:
            self.pkg = render(self.cam, self.pc, None, self.bg, time=self.t, it=self.iteration, capacity=self.capacity)
            self.loss = self._flag_overflow(self.loss_fn(self.pkg["render"], self.target))
            self.loss.backward()
        if self.capture_collective:
            import torch.distributed as dist
            dist.all_reduce(self.reducer.flat, op=dist.ReduceOp.SUM, group=self.reducer.group)

    # ---- input prefetch: the next sample's host->device copies overlap the current step's kernels ----
    def prefetch(self, cam, t, target):
        """Start copying the NEXT sample (pinned host tensors) on a side stream into a staging slot; the following
        ``step_prefetched()`` moves it into the graph's static inputs (device-to-device) and replays."""
        if not hasattr(self, "_stage"):
            self._copy_stream = torch.cuda.Stream(device=self.dev)
            self._stage = [dict(host=torch.empty(36).pin_memory(), small=torch.empty(36, device=self.dev),
                                target=torch.empty_like(self.target), ready=torch.cuda.Event(), free=torch.cuda.Event())
                           for _ in range(2)]
            self._slot = 0
            self._pending = []
            for s in self._stage:
                s["free"].record(torch.cuda.current_stream(self.dev))
        s = self._stage[self._slot]
        if s.get("used"):
            s["ready"].synchronize()                         # the slot's previous H2D has left its pinned host buffer
        h = s["host"]                                        # camera + time packed on the host: one H2D, one D2D later
        h[0:16].view(4, 4).copy_(cam.world_view_transform)
        h[16:32].view(4, 4).copy_(cam.full_proj_transform)
        h[32:35].copy_(cam.camera_center)
        h[35:36].copy_(t.reshape(-1)[:1])
        with torch.cuda.stream(self._copy_stream):
            self._copy_stream.wait_event(s["free"])          # the step that consumed this slot has read it
            s["small"].copy_(h, non_blocking=True)
            s["target"].copy_(target, non_blocking=True)
            s["ready"].record(self._copy_stream)
        s["used"] = True
        self._pending.append(s)                               # FIFO: at most two samples are in flight (two slots)
        assert len(self._pending) <= 2, "prefetch() twice without a step_prefetched() in between is the limit"
        self._slot ^= 1

    def step_prefetched(self, exchange=None, **kw):
        s = self._pending.pop(0)
        cur = torch.cuda.current_stream(self.dev)
        cur.wait_event(s["ready"])
        self.cam.packed.copy_(s["small"])
        self.target.copy_(s["target"])
        s["free"].record(cur)
        return self._replay(exchange, **kw)

    def step_packed(self, packed, target, exchange=None, **kw):
        """Device-resident sample in the packed form of ``_StaticCamera`` (view 16 | proj 16 | centre 3 | time 1, float32
        [36]): two copies into the static inputs instead of five."""
        self.cam.packed.copy_(packed, non_blocking=True)
        self.target.copy_(target, non_blocking=True)
        return self._replay(exchange, **kw)

    def __call__(self, cam, t, target, non_blocking=True, exchange=None, **kw):
        self.cam.load(cam, non_blocking)
        self.t.copy_(t, non_blocking=non_blocking)
        self.target.copy_(target, non_blocking=non_blocking)
        return self._replay(exchange, **kw)

    def _replay(self, exchange=None, wait_before_blend=None, wait_before_raster=None, defer=False):
        """``exchange``: None = this step's own reducer (when it has one and a process group of more than one rank
        exists), False = no exchange (a sample in the middle of an accumulation step, or a single-rank run), or the
        ``FlatGradAllReduce`` whose buckets are to be summed over the ranks (the accumulator of a ``BatchStep``).

        Pipelined exchange (``defer=True``, ``split_forward`` steps): only the LAST bucket (what the next step's
        neighbour search and keypoint MLP read: xyz, keypoints, MLP weights) is joined before this call returns; the
        per-Gaussian bucket is handed back in ``deferred_blend`` and the SH bucket in ``deferred_raster`` -- the next
        step joins them through ``wait_before_blend`` / ``wait_before_raster``, right in front of their first reader.
        Exchanges execute in issue order (one NCCL / side stream): last bucket, radii, per-Gaussian bucket, SH bucket
        (optionally a head of the SH bucket first, ``EARLY_FRACTION``)."""
        self.deferred_blend = self.deferred_raster = None
        if self.two_stage:
            import torch.distributed as dist
            if exchange is None:
                exchange = self.reducer if (dist.is_available() and dist.is_initialized() and
                                            dist.get_world_size(self.reducer.group) > 1) else False
            npre = getattr(self, "_n_pre", 0)
            _mark("begin")
            if npre:
                self.stage_graphs[0].replay()     # neighbour search + keypoint MLP
                _mark("search + mlp done")
                if wait_before_blend is not None:
                    wait_before_blend.wait()
                self.stage_graphs[1].replay()     # blend
                _mark("blend done")
            if wait_before_raster is not None:    # the previous step's SH exchange ran underneath all of that
                wait_before_raster.wait()
                _mark("joined previous SH")
            stages = self.stage_graphs[npre:]
            last = len(stages) - 1
            if exchange is False:
                for g in stages:
                    g.replay()
            else:
                buckets, works = exchange.buckets(), []
                pipelined = defer and len(stages) == 3 and len(buckets) == 3
                if not pipelined:
                    for k, g in enumerate(stages):
                        g.replay()                # stage k+1 runs while bucket k is on the wire
                        for b in (buckets[k:] if k == last else buckets[k:k + 1]):
                            works.append(exchange.all_reduce_bucket(b))
                else:
                    sh = buckets[0]
                    stages[0].replay()
                    _mark("stage 0 done")
                    if getattr(exchange, "copy_engine", False):
                        # Two ranks: copy engines.  A copy costs the SMs nothing (the next step's neighbour search runs
                        # at full speed next to it; next to the 148-CTA peer-load kernel it took 1.2 ms instead of
                        # 0.45), so the SH range -- 2/3 of the bytes -- starts crossing the link as soon as it is final,
                        # underneath the rest of this backward, and the other ranges follow on a second copy stream
                        # when the last gradients are final.  Every range is added in place in front of its first
                        # reader (parallel.NvlsFlatGradAllReduce.ce_pull).  What stays exposed is link time: 569 MB at
                        # 730 GB/s do not fit under the 0.3 ms of backward + 0.55 ms of search and blend that precede
                        # the SH range's first reader.  Measured at C5 (step of 4 samples per rank): 10.54 ms; all
                        # ranges after the last stage, one after the other: 10.62; the small first range through the
                        # peer-load kernel instead: 10.60; the resident peer-load kernel for everything: 10.69.
                        from .parallel import _Joins
                        h_sh = exchange.ce_pull(sh, 0)
                        stages[1].replay()
                        _mark("stage 1 done")
                        stages[2].replay()
                        _mark("stage 2 done")
                        works.append(exchange.ce_pull(buckets[2], 1))
                        h_aux = exchange.ce_pull("aux", 1) if self.exchange_radii else None
                        self.deferred_blend = _Joins([h_aux, exchange.ce_pull(buckets[1], 1)])
                        self.deferred_raster = h_sh
                    else:
                        cut = int(sh.numel() * EARLY_FRACTION) // 4 * 4
                        if cut:
                            works.append(exchange.all_reduce_bucket(sh[:cut]))
                        stages[1].replay()
                        _mark("stage 1 done")
                        stages[2].replay()
                        _mark("stage 2 done")
                        if hasattr(exchange, "exchange_segments") and RESIDENT_EXCHANGE:
                            # one resident kernel: it takes its CTAs now, while the GPU is between steps, and is not
                            # locked out by the persistent blocks of the next step's neighbour search
                            segs = [buckets[2], buckets[1]] + (["aux"] if self.exchange_radii else []) + \
                                ([sh[cut:]] if cut < sh.numel() else [])
                            hs = exchange.exchange_segments(segs)
                            works.append(hs[0])
                            self.deferred_blend = hs[2] if self.exchange_radii else hs[1]
                            self.deferred_raster = hs[-1]
                        else:
                            works.append(exchange.all_reduce_bucket(buckets[2]))
                            if self.exchange_radii:   # MAX of the radii: in front of the per-Gaussian bucket, joined with it
                                exchange.all_reduce_max_aux()
                            self.deferred_blend = exchange.all_reduce_bucket(buckets[1])
                            self.deferred_raster = exchange.all_reduce_bucket(sh[cut:]) if cut < sh.numel() else self.deferred_blend
                for w in works:
                    if w is not None:
                        w.wait()
                _mark("joined exchanges")
        else:
            self.graph.replay()
        for p, g in zip(self.params, self.grads):     # p.grad may have been replaced / cleared by the caller
            p.grad = g
        return self.loss

    @property
    def image(self):
        return self.pkg["render"]


class RenderStep:
    """Forward-only ``render()`` as one graph (the eval / extrapolation frame loop, eval.py:192-226)."""

    def __init__(self, pc, cam, capacity, bg=None, iteration=50_000, warmup=3):
        self.pc = pc
        dev = next(pc.parameters()).device
        self.dev, self.capacity, self.iteration = dev, int(capacity), int(iteration)
        self.bg = torch.zeros(3, device=dev) if bg is None else bg
        self.cam = _StaticCamera(cam, dev)
        self.cam.load(cam, non_blocking=False)
        self.t = torch.zeros(1, device=dev)
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side), torch.no_grad():
            for _ in range(max(warmup, 1)):
                self._body()
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        self.graph = torch.cuda.CUDAGraph()
        with torch.no_grad(), torch.cuda.graph(self.graph):
            self._body()
        torch.cuda.synchronize(dev)

    def _body(self):
        self.pkg = render(self.cam, self.pc, None, self.bg, time=self.t, it=self.iteration, capacity=self.capacity)

    def __call__(self, cam, t, non_blocking=True):
        self.cam.load(cam, non_blocking)
        self.t.copy_(t, non_blocking=non_blocking)
        self.graph.replay()
        return self.pkg


class BatchStep:
    """One optimizer step of the reference's ``--batch`` loop (train.py:113-133) on this rank's share of the step's
    (camera, time) samples: every sample is one replay of a captured train iteration, the parameter gradients are
    summed over the samples (the reference sums the losses and calls backward once), ``radii`` are max-reduced and the
    visibility mask or-reduced (train.py:122-123).  With more than one rank the sums are then taken over the ranks too
    (``parallel``: SUM for gradients, MAX for radii), started bucket by bucket while the last sample's backward is still
    running.

    Two captured steps share the work: ``first`` writes its gradients straight into the accumulator (no zero fill, no
    copy), ``rest`` writes into a scratch buffer of the same layout and ends each stage with ``accumulator += scratch``.
    After ``step()`` ``p.grad`` of every exchanged parameter is its view of the (summed) accumulator.
    """

    def __init__(self, pc, cam, capacity, make_reducer, samples_per_rank, bg=None, iteration=50_000, loss_fn=None,
                 group=None):
        import torch.distributed as dist
        self.pc = pc
        self.n = int(samples_per_rank)
        self.world = dist.get_world_size(group) if (dist.is_available() and dist.is_initialized()) else 1
        self.group = group
        self.acc = make_reducer()
        # Pipelined exchange (more than one rank): only the bucket the next step reads FIRST (xyz, keypoints, MLP weights)
        # is joined at the end of a step.  The per-Gaussian bucket is joined in front of the next step's blend kernel and
        # the SH bucket -- three quarters of the bytes -- in front of its rasterizer, their first readers: they (and the
        # optimizer updates of those groups) run underneath the next step's neighbour search, MLP and blend.
        self.pipelined = self.world > 1 and os.environ.get("GP_PIPELINE_EXCHANGE", "1") != "0"
        self._sh_pending = self._blend_pending = None
        dev = next(pc.parameters()).device
        self.dev = dev
        self.loss_sum = torch.zeros((), device=dev)
        P = int(pc.get_xyz.shape[0])
        aux = getattr(self.acc, "aux", None)          # the reducer's own MAX-exchanged buffer, when it was given one
        self._radii_in_acc = aux is not None and aux.numel() == P
        self.radii = aux if self._radii_in_acc else torch.zeros(P, dtype=torch.int32, device=dev)   # max over samples (and ranks)
        # the running loss sum and radii max are updated inside the captured stage 0 of every sample (TrainStep ``stats``)
        self.first = TrainStep(pc, cam, capacity, bg=bg, iteration=iteration, loss_fn=loss_fn, reducer=self.acc,
                               split_forward=self.pipelined, stats=(self.loss_sum, self.radii, True))
        self.rest = None
        if self.n > 1:
            self.scratch = make_reducer()
            self.rest = TrainStep(pc, cam, capacity, bg=bg, iteration=iteration, loss_fn=loss_fn, reducer=self.scratch,
                                  accumulate_into=self.acc, stats=(self.loss_sum, self.radii, False))
        # pipelined + the radii live in the reducer: their MAX exchange is issued inside the last sample's replay, in
        # front of the per-Gaussian bucket, and is joined with it (next step, in front of the blend)
        self._radii_inline = self.pipelined and self._radii_in_acc
        (self.rest if self.rest is not None else self.first).exchange_radii = self._radii_inline

    def _steps(self):
        return [self.first] + [self.rest] * (self.n - 1)

    def _finish(self):
        _mark("finish")
        if self.world > 1 and self.exchange and not self._radii_inline:
            import torch.distributed as dist
            w = self.acc.all_reduce_max_aux() if self._radii_in_acc else \
                dist.all_reduce(self.radii, op=dist.ReduceOp.MAX, group=self.group, async_op=True)
            if self.pipelined:
                self._radii_pending = w       # 4 B per Gaussian: joined by flush() / the next step
            else:
                w.wait()
        for p, v in zip(self.acc.params, self.acc.views):
            p.grad = v
        return self.loss_sum

    def _exchange(self, j):
        return self.acc if (self.world > 1 and self.exchange and j == self.n - 1) else False

    exchange = True     # False: leave the per-rank sums un-exchanged (diagnostics: bench.py's exchange check)

    def _kw(self, j):
        kw = {}
        if j == 0 and self._sh_pending is not None:
            kw["wait_before_raster"], self._sh_pending = self._sh_pending, None
        if j == 0 and self._blend_pending is not None:
            kw["wait_before_blend"], self._blend_pending = self._blend_pending, None
        if self.pipelined and self._exchange(j) is not False:
            kw["defer"] = True
        return kw

    def _keep_deferred(self, ts):
        if ts.deferred_raster is not None:
            self._sh_pending, ts.deferred_raster = ts.deferred_raster, None
        if ts.deferred_blend is not None:
            self._blend_pending, ts.deferred_blend = ts.deferred_blend, None

    def flush(self):
        """Join what the last step left on the wire (pipelined mode) to the current stream: call it before reading
        those gradients (or the radii), e.g. in front of the optimizer update of the per-Gaussian and SH groups.  The
        next ``step()`` joins them by itself, each in front of its first reader."""
        for name in ("_blend_pending", "_sh_pending"):
            w = getattr(self, name)
            if w is not None:
                w.wait()
                setattr(self, name, None)
        self._join_radii()

    def _join_radii(self):
        w, self._radii_pending = getattr(self, "_radii_pending", None), None
        if w is not None:
            w.wait()

    def _before_first(self, ts):
        # stage 0 of sample 0 overwrites the radii: a MAX exchange of the previous step that is still pending must be
        # done (inline mode: it sits in front of the per-Gaussian bucket on the side stream, which is joined before the
        # blend, i.e. before stage 0)
        if not ts.split_forward:
            self.flush()
        else:
            self._join_radii()

    def step(self, samples):
        """``samples``: this rank's ``samples_per_rank`` samples, device tensors: tuples (camera, time [1], target
        [3,H,W]) or, cheaper, (packed camera + time [36], target) -- see ``TrainStep.step_packed``."""
        assert len(samples) == self.n
        for j, (ts, sample) in enumerate(zip(self._steps(), samples)):
            if j == 0:
                self._before_first(ts)
            if len(sample) == 2:
                ts.step_packed(sample[0], sample[1], exchange=self._exchange(j), **self._kw(j))
            else:
                ts(sample[0], sample[1], sample[2], exchange=self._exchange(j), **self._kw(j))
            self._keep_deferred(ts)
        return self._finish()

    def step_host(self, samples, next_first=None):
        """Same, the samples being PINNED HOST tensors: the copy of sample j + 1 is issued on a side stream before
        sample j is replayed (``TrainStep.prefetch``); ``next_first`` is the first sample of the NEXT step, whose copy
        then overlaps this step's last sample."""
        assert len(samples) == self.n
        steps = self._steps()
        if not getattr(self, "_primed", False):
            steps[0].prefetch(*samples[0])
        for j, ts in enumerate(steps):
            if j + 1 < self.n:
                steps[j + 1].prefetch(*samples[j + 1])
            elif next_first is not None:
                steps[0].prefetch(*next_first)
            if j == 0:
                self._before_first(ts)
            ts.step_prefetched(exchange=self._exchange(j), **self._kw(j))
            self._keep_deferred(ts)
        self._primed = next_first is not None
        return self._finish()

    def overflowed(self):
        a = self.first.overflowed()
        b = self.rest.overflowed() if self.rest is not None else (0, False)
        return max(a[0], b[0]), (a[1] or b[1])

    def grow(self, capacity=None):
        """Both captured steps re-captured with the same, larger capacity (see ``TrainStep.grow``)."""
        self.flush()
        R, _ = self.overflowed()
        new = int(capacity) if capacity is not None else int(max(R, self.first.capacity) * 1.5) + 65536
        self.first.grow(new)
        if self.rest is not None:
            self.rest.grow(new)
        self._primed = False
        return new

    def step_checked(self, samples):
        """``step()`` with the overflow policy applied: the loss of an overflowing frame is NaN (device side); if the
        summed loss comes back NaN and a frame did overflow, the steps are re-captured with more room and the step is
        run again.  One host read of the loss per step -- the read the training loop does anyway (train.py:139)."""
        loss = self.step(samples)
        if torch.isnan(loss).item() and self.overflowed()[1]:
            self.grow()
            loss = self.step(samples)
        return loss


class MotionStep:
    """One frame of the motion-extrapolation loop (eval.py:120-157, ``render_kpts``) as one graph: the per-frame blend
    ``xyz + W_x (kpts_i - kpts)``, ``normalize(quat_mul(normalize(W_r rot_i), rotation))`` (eval.py:140-141, sparse: the
    neighbour indices and blend logits of one earlier ``pc(time, iteration)`` call are reused, as the reference reuses
    its dense weights) followed by ``render_motion`` (gaussian_renderer/__init__.py:117-191)."""

    def __init__(self, pc, cam, capacity, bg=None, warmup=3, opacity=None):
        from .renderer import render_motion
        self.pc, self._render_motion = pc, render_motion
        self.model = getattr(pc, "deformation", pc)
        assert self.model._idx32 is not None, "call pc(time, iteration) once first (eval.py:125-126)"
        dev = next(pc.parameters()).device
        self.dev, self.capacity = dev, int(capacity)
        self.bg = torch.zeros(3, device=dev) if bg is None else bg
        self.opacity = opacity                      # ``lifecycle_opacity`` of that call, or None (eval.py:127)
        self.cam = _StaticCamera(cam, dev)
        self.cam.load(cam, non_blocking=False)
        Kp = int(self.model.super_gaussians.shape[0])
        self.kpts = torch.zeros(Kp, 3, device=dev)
        self.kpts_rotation = torch.zeros(Kp, 4, device=dev)
        self.kpts_rotation[:, 0] = 1.0
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side), torch.no_grad():
            for _ in range(max(warmup, 1)):
                self._body()
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        self.graph = torch.cuda.CUDAGraph()
        with torch.no_grad(), torch.cuda.graph(self.graph):
            self._body()
        from . import rasterizer as _r
        self._state = _r.LAST_STATE
        torch.cuda.synchronize(dev)

    def _body(self):
        xyz_t, q_t = self.model.extrapolate(self.kpts, self.kpts_rotation)
        self.pkg = self._render_motion(self.cam, self.pc, None, self.bg, xyz_t=xyz_t, r_t=q_t, opacity=self.opacity,
                                       capacity=self.capacity)

    def __call__(self, cam, kpts_i, kpts_rotation_i, non_blocking=True):
        self.cam.load(cam, non_blocking)
        self.kpts.copy_(kpts_i, non_blocking=non_blocking)
        self.kpts_rotation.copy_(kpts_rotation_i, non_blocking=non_blocking)
        self.graph.replay()
        return self.pkg

    def overflowed(self):
        hdr = self._state.geom[:16].view(torch.int32).cpu()
        return int(hdr[_lib.HDR_NUM_RENDERED]), bool(int(hdr[_lib.HDR_OVERFLOW]))
