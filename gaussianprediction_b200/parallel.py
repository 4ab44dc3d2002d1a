"""Data-parallel plumbing for the frame loop (one process per GPU, torch.distributed over NCCL / NVLink).

The path shards by (camera, timestep) sample: parameters are replicated, every rank renders its own
samples, and one exchange per optimizer step sums the parameter gradients.  The reference has no distributed code;
its single-GPU analogue is the ``--batch`` accumulation of train.py:113-133, which SUMS the losses of consecutive
samples -- hence ``all_reduce(SUM)`` with no averaging -- and keeps ``max`` of the radii and ``or`` of the visibility
masks (train.py:122-123).  Evaluation / extrapolation frames shard with no communication at all.
"""
import torch
import torch.distributed as dist


def shard_samples(n_samples, rank, world_size):
    """Indices of the samples rank ``rank`` renders: r, r + W, r + 2W, ... (frames f = r mod W)."""
    return list(range(rank, n_samples, world_size))


class FlatGradAllReduce:
    """One flat buffer for all parameter gradients -> a single all-reduce per step.

    ``reduce()`` packs ``p.grad`` of every parameter into the buffer (zeros where a parameter got no gradient on
    this rank), sums it over the process group and points ``p.grad`` at views of the reduced buffer, so every rank
    ends the step with identical gradients (and, with identical optimizer state, identical parameters)."""

    def __init__(self, params, group=None, first=(), second=(), aux_int32=0):
        """``first`` / ``second``: parameters whose gradients are complete early in the backward, in that order -- the
        SH coefficients (written by the rasterizer backward), then the per-Gaussian deformation parameters (written
        by the blend backward).  Each group becomes its own bucket at the front of the buffer so that
        ``pipeline.TrainStep`` can start its all-reduce while the rest of the backward is still running; everything
        else forms the last bucket.  ``reduce()`` itself always exchanges the whole buffer."""
        params = [p for p in params if p.requires_grad]
        self._aux_n = int(aux_int32)      # optional int32 side buffer exchanged with MAX (radii, train.py:122)
        groups, seen = [], set()
        for grp in (first, second):
            ids = {id(p) for p in grp}
            groups.append([p for p in params if id(p) in ids and id(p) not in seen])
            seen |= ids
        groups.append([p for p in params if id(p) not in seen])
        self.params = [p for grp in groups for p in grp]
        self.group = group
        # Every view starts on a 16-byte boundary (ALIGN elements): the backward kernels write gradients straight into
        # these views with 128-bit stores (pipeline.TrainStep, _sinks), and after densification the Gaussian count is
        # rarely a multiple of 4.  The padding elements stay zero and travel with the exchange (< 16 B per parameter).
        ALIGN = 4
        pad = lambda k: (k + ALIGN - 1) // ALIGN * ALIGN
        offs, off = [], 0
        for p in self.params:
            offs.append(off)
            off = pad(off + p.numel())
        p0 = self.params[0]
        self.flat = self._alloc_flat(max(off, ALIGN), p0.dtype, p0.device)
        self.views = [self.flat[o:o + p.numel()].view_as(p) for o, p in zip(offs, self.params)]
        # per bucket: (first parameter index, one past the last, first element, one past the last); empty ones dropped
        self.bucket_ranges = []
        i = 0
        for grp in groups:
            if grp:
                e0 = offs[i]
                e1 = offs[i + len(grp)] if i + len(grp) < len(offs) else self.flat.numel()
                self.bucket_ranges.append((i, i + len(grp), e0, e1))
            i += len(grp)

    def _alloc_flat(self, n, dtype, device):
        self.aux = torch.zeros(self._aux_n, dtype=torch.int32, device=device) if self._aux_n else None
        return torch.zeros(n, dtype=dtype, device=device)

    def all_reduce_max_aux(self):
        """Element-wise MAX of ``aux`` over the ranks; returns a handle with ``wait()``."""
        return dist.all_reduce(self.aux, op=dist.ReduceOp.MAX, group=self.group, async_op=True)

    def buckets(self):
        """The flat buffer bucket by bucket, in the order their gradients complete."""
        return [self.flat[e0:e1] for _, _, e0, e1 in self.bucket_ranges]

    def pack(self, lo=0, hi=None):
        """Make ``p.grad`` of parameters [lo, hi) the parameter's view of the flat buffer (zero / copy where it is not
        already: gradients the kernels wrote through a grad sink are in place and cost nothing here)."""
        moved = 0
        for p, v in list(zip(self.params, self.views))[lo:hi]:
            if p.grad is None:
                v.zero_()
                moved += 1
            elif p.grad.data_ptr() != v.data_ptr():
                v.copy_(p.grad)
                moved += 1
            p.grad = v
        return moved                         # how many gradients had to be filled in / copied

    @property
    def nbytes(self):
        return self.flat.numel() * self.flat.element_size()

    def all_reduce_bucket(self, bucket):
        """Sum one bucket (a slice of ``flat``) over the ranks; returns a work handle or None.  Subclasses replace the
        transport (``NvlsFlatGradAllReduce``: an in-stream kernel over NVSwitch multicast memory)."""
        return dist.all_reduce(bucket, op=dist.ReduceOp.SUM, group=self.group, async_op=True)

    def reduce(self, async_op=False):
        self.pack()
        return dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=self.group, async_op=async_op)


class GroupedGradAllReduce:
    """Sum the parameter gradients over the ranks WITHOUT packing them: all tensors go into one NCCL group call
    (ncclGroupStart/End through torch's coalescing manager), i.e. one fused collective over NVLink with no staging
    copy of the ~236 B per Gaussian.  Falls back to ``FlatGradAllReduce`` on backends without coalescing (gloo)."""

    def __init__(self, params, group=None):
        self.params = [p for p in params if p.requires_grad]
        self.group = group
        self._flat = None

    @property
    def nbytes(self):
        return sum(p.numel() * p.element_size() for p in self.params)

    def reduce(self):
        backend = dist.get_backend(self.group)
        if backend != "nccl":
            if self._flat is None:
                self._flat = FlatGradAllReduce(self.params, self.group)
            return self._flat.reduce()
        # Parameters without a gradient are skipped: which leaves receive one is a property of the model graph, so it
        # is the same on every rank and the ranks still issue identical collectives.
        live = [p for p in self.params if p.grad is not None]
        grads = [p.grad if p.grad.is_contiguous() else p.grad.contiguous() for p in live]
        if not grads:
            return None
        from torch.distributed.distributed_c10d import _coalescing_manager
        with _coalescing_manager(group=self.group, device=grads[0].device, async_ops=False):
            for g in grads:
                dist.all_reduce(g, op=dist.ReduceOp.SUM, group=self.group)
        for p, g in zip(live, grads):
            if p.grad is not g:
                p.grad = g
        return None


def reduce_densification_stats(radii, visibility_filter, group=None):
    """``radii = max over samples``, ``visibility = or over samples`` (train.py:122-123) across ranks, in place."""
    dist.all_reduce(radii, op=dist.ReduceOp.MAX, group=group)
    vis = visibility_filter.to(torch.int32)
    dist.all_reduce(vis, op=dist.ReduceOp.MAX, group=group)
    return radii, vis.bool()


class NvlsFlatGradAllReduce(FlatGradAllReduce):
    """``FlatGradAllReduce`` whose exchange is this library's own kernel (csrc/exchange.cu) instead of NCCL: the flat
    buffer lives in a symmetric allocation (``torch.distributed._symmetric_memory`` provides the allocation and the
    address exchange -- plumbing), every bucket is summed in place by ONE kernel per rank that reduces through the
    NVSwitch multicast mapping (``multimem.ld_reduce`` / ``multimem.st``: in-switch reduction, NVLS) or, where no
    multicast mapping exists, through plain peer loads and stores.  The kernel runs on a side stream next to the
    backward kernels of the following stage; ``all_reduce_bucket`` returns a handle whose ``wait()`` joins it to the
    current stream.  Results are bit-identical on every rank (each word is summed once by its owner and broadcast)."""

    BLOCKS = int(__import__("os").environ.get("GP_EXCHANGE_BLOCKS", "24"))   # CTAs of the exchange kernel (tools/exchange_probe.py)

    @staticmethod
    def supported(group=None):
        """Collective: True on every rank iff a symmetric allocation can be set up on all of them."""
        if not (dist.is_available() and dist.is_initialized()) or dist.get_backend(group) != "nccl":
            return False
        ok = 1
        try:
            import torch.distributed._symmetric_memory as symm
            t = symm.empty(1024, dtype=torch.float32, device=torch.device("cuda", torch.cuda.current_device()))
            symm.rendezvous(t, group if group is not None else dist.group.WORLD)
        except Exception:
            ok = 0
        flag = torch.tensor([ok], device="cuda")
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)
        return bool(flag.item())

    def _alloc_flat(self, n, dtype, device):
        import ctypes
        import os
        import torch.distributed._symmetric_memory as symm
        from . import _lib
        assert dtype == torch.float32
        L = _lib.lib()
        self._flag_bytes = int(L.gp_allreduce_flag_bytes(max(self.BLOCKS, 148) + 4))   # + the copy-engine path's own slots
        data_bytes = (n * 4 + 255) // 256 * 256
        self._aux_off = data_bytes
        data_bytes += (self._aux_n * 4 + 255) // 256 * 256
        self._flags_off = data_bytes
        total = data_bytes + self._flag_bytes
        self._symm = symm.empty(total // 4, dtype=torch.float32, device=device)
        self._symm.zero_()
        self._hdl = symm.rendezvous(self._symm, self.group if self.group is not None else dist.group.WORLD)
        torch.cuda.synchronize(device)
        dist.barrier(group=self.group)              # every rank's flags are zero before anybody signals
        self._rank, self._world = int(self._hdl.rank), int(self._hdl.world_size)
        self._peers = (ctypes.c_void_p * self._world)(*[int(p) for p in self._hdl.buffer_ptrs])
        mc = int(getattr(self._hdl, "multicast_ptr", 0) or 0)
        # measured (tools/exchange_probe.py, B200 NVSwitch box, 512 MiB): 8 ranks multimem 1.13 ms on 24 CTAs vs NCCL 1.34,
        # peer ld/st 1.51; 2 ranks peer ld/st 0.81 ms on 148 CTAs vs NCCL 1.00, multimem 1.33 -- the in-switch reduction
        # pays from 3 ranks up
        self.multicast = mc != 0 and os.environ.get("GP_EXCHANGE_NO_MULTICAST") is None and self._world > 2
        self._mc = mc if self.multicast else None
        if "GP_EXCHANGE_BLOCKS" not in os.environ:
            self.BLOCKS = 24 if self.multicast else 148
        self._side = torch.cuda.Stream(device=device)
        # Two ranks: the copy engines move the bytes (``ce_pull``): the neighbour search of the next step runs at full
        # speed next to a cudaMemcpyAsync, while it crawls next to the 148-CTA peer-load kernel (tools/ce_copy_probe.py)
        self.copy_engine = self._world == 2 and os.environ.get("GP_EXCHANGE_CE", "1") != "0"
        self._ce_streams, self._stage = None, None
        self._ce_slot0 = max(self.BLOCKS, 148)       # flag slots of the copy-engine path (the kernels use the ones below)
        self.aux = self._symm[self._aux_off // 4:self._aux_off // 4 + self._aux_n].view(torch.int32) if self._aux_n else None
        return self._symm[:n]

    def ce_pull(self, segment, lane):
        """Two ranks.  Start pulling the peer's copy of ``segment`` (a slice of ``flat``, or "aux") into local staging
        on copy stream ``lane`` (0 or 1): a flag barrier with the peer (its gradients are final), then one
        cudaMemcpyAsync over NVLink -- no SM takes part.  Returns a handle whose ``wait()``, on the current stream,
        joins the copy, meets the peer once more (it has finished reading THIS rank's range, which is about to change)
        and adds the staged range in place (``gp_add_inplace_f32`` / ``gp_max_inplace_s32``).  a + b is commutative, so
        both ranks end with the same bits.  Calls and waits must come in the same order on both ranks."""
        from . import _lib
        L = _lib.lib()
        dev = self.flat.device
        if self._ce_streams is None:
            self._ce_streams = [torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)]
            self._stage = torch.empty(self._flags_off // 4, dtype=torch.float32, device=dev)
        if isinstance(segment, str):
            off, nbytes, op = self._aux_off, self._aux_n * 4, 1
        else:
            off, nbytes, op = segment.data_ptr() - self._symm.data_ptr(), segment.numel() * 4, 0
            assert off % 16 == 0 and nbytes % 16 == 0 and 0 <= off and off + nbytes <= self._aux_off
        st = self._ce_streams[lane]
        ready = torch.cuda.Event()
        ready.record(torch.cuda.current_stream(dev))
        st.wait_event(ready)
        _lib.check(L.gp_peer_barrier_launch(self._rank, self._world, self._peers, self._flags_off, self._ce_slot0 + lane,
                                            st.cuda_stream), "gp_peer_barrier_launch")
        peer_base = int(self._peers[1 - self._rank])
        _lib.check(L.gp_copy_async(self._stage.data_ptr() + off, peer_base + off, nbytes, st.cuda_stream), "gp_copy_async")
        done = torch.cuda.Event()
        done.record(st)
        return _CeJoin(self, done, off, nbytes, op)

    def _launch(self, fn, what, byte_off, n):
        from . import _lib
        from . import pipeline as _p
        cur = torch.cuda.current_stream(self.flat.device)
        ready, done = torch.cuda.Event(), torch.cuda.Event()
        ready.record(cur)
        self._side.wait_event(ready)
        _p._mark(f"{what} {n * 4 >> 20} MiB start", self._side)
        _lib.check(fn(self._rank, self._world, self._mc, self._peers, byte_off, n, self._flags_off, self.BLOCKS,
                      self._side.cuda_stream), what)
        _p._mark(f"{what} {n * 4 >> 20} MiB end", self._side)
        done.record(self._side)
        return _StreamJoin(done, self.flat.device)

    def all_reduce_bucket(self, bucket):
        """Sum any 16-byte aligned slice of ``flat`` over the ranks on the side stream.  The calls of one reducer run
        one after the other (same stream), so they share one flag region: epochs only grow, and a rank that is already
        in the next call cannot release a peer that still waits in this one before that peer's own data is complete."""
        from . import _lib
        off = bucket.data_ptr() - self._symm.data_ptr()
        n = bucket.numel()
        assert off % 16 == 0 and n % 4 == 0 and 0 <= off and off + 4 * n <= self._aux_off
        return self._launch(_lib.lib().gp_allreduce_sum_f32, "gp_allreduce_sum_f32", off, n)

    def all_reduce_max_aux(self):
        from . import _lib
        return self._launch(_lib.lib().gp_allreduce_max_s32, "gp_allreduce_max_s32", self._aux_off, self._aux_n)

    def exchange_segments(self, segments):
        """``segments``: slices of ``flat`` (summed) and/or the string "aux" (the int32 side buffer, MAX), in the order
        their readers come.  ONE resident kernel on the side stream works through them (csrc/exchange.cu,
        gp_exchange_segments_kernel); returns one handle per segment whose ``wait()`` makes the current stream wait (a
        one-thread kernel polling a device word) until that segment is complete on every rank."""
        import ctypes
        from . import _lib
        from . import pipeline as _p
        L = _lib.lib()
        if not hasattr(self, "_sync"):
            self._sync = torch.zeros(int(L.gp_exchange_sync_words()), dtype=torch.int32, device=self.flat.device)
            self._call_id = 0
        offs, counts, ops = [], [], []
        for sgm in segments:
            if isinstance(sgm, str):
                offs.append(self._aux_off); counts.append(self._aux_n); ops.append(1)
            else:
                off = sgm.data_ptr() - self._symm.data_ptr()
                assert off % 16 == 0 and sgm.numel() % 4 == 0 and 0 <= off and off + 4 * sgm.numel() <= self._aux_off
                offs.append(off); counts.append(sgm.numel()); ops.append(0)
        n = len(offs)
        self._call_id += 1
        cur = torch.cuda.current_stream(self.flat.device)
        ready = torch.cuda.Event()
        ready.record(cur)
        self._side.wait_event(ready)
        _p._mark(f"resident exchange ({n} ranges) start", self._side)
        _lib.check(L.gp_exchange_segments(self._rank, self._world, self._mc, self._peers, self._flags_off, self.BLOCKS, n,
                                          (ctypes.c_size_t * n)(*offs), (ctypes.c_size_t * n)(*counts), (ctypes.c_int * n)(*ops),
                                          self._sync.data_ptr(), self._call_id, self._side.cuda_stream), "gp_exchange_segments")
        _p._mark("resident exchange end", self._side)
        half = self._sync.numel() // 2
        return [_WordJoin(self._sync.data_ptr() + 4 * (half + k), self._call_id, self.flat.device) for k in range(n)]

    def reduce(self, async_op=False):
        self.pack()
        works = [self.all_reduce_bucket(b) for b in self.buckets()]
        for w in works:
            w.wait()
        return None


class _CeJoin:
    """Second half of ``NvlsFlatGradAllReduce.ce_pull`` (runs on the stream that calls ``wait()``)."""

    def __init__(self, red, done, off, nbytes, op):
        self.red, self.done, self.off, self.nbytes, self.op = red, done, off, nbytes, op

    def wait(self):
        from . import _lib
        L, r = _lib.lib(), self.red
        cur = torch.cuda.current_stream(r.flat.device)
        cur.wait_event(self.done)
        _lib.check(L.gp_peer_barrier_launch(r._rank, r._world, r._peers, r._flags_off, r._ce_slot0 + 2, cur.cuda_stream),
                   "gp_peer_barrier_launch")
        dst, src = r._symm.data_ptr() + self.off, r._stage.data_ptr() + self.off
        if self.op == 0:
            _lib.check(L.gp_add_inplace_f32(dst, src, self.nbytes // 4, cur.cuda_stream), "gp_add_inplace_f32")
        else:
            _lib.check(L.gp_max_inplace_s32(dst, src, self.nbytes // 4, cur.cuda_stream), "gp_max_inplace_s32")


class _Joins:
    """Several handles joined in order."""

    def __init__(self, handles):
        self.handles = [h for h in handles if h is not None]

    def wait(self):
        for h in self.handles:
            h.wait()


class _StreamJoin:
    def __init__(self, event, device):
        self.event, self.device = event, device

    def wait(self):
        torch.cuda.current_stream(self.device).wait_event(self.event)


class _WordJoin:
    """``wait()``: the current stream waits until a device word reaches ``value`` (gp_stream_wait_word)."""

    def __init__(self, word_ptr, value, device):
        self.word_ptr, self.value, self.device = word_ptr, value, device

    def wait(self):
        from . import _lib
        _lib.check(_lib.lib().gp_stream_wait_word(self.word_ptr, self.value,
                                                  torch.cuda.current_stream(self.device).cuda_stream), "gp_stream_wait_word")
