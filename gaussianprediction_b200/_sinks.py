"""Gradient sinks: let the backward kernels write parameter gradients straight into caller-owned memory.

Data-parallel training keeps one flat buffer per all-reduce bucket (parallel.FlatGradAllReduce).  Without help, every
gradient is first produced in a fresh tensor and then added into its slice of that buffer (a zero fill, a read-modify-
write and a launch per parameter and step).  While a sink map is installed, the autograd Functions of this package
look up the tensor a gradient belongs to (by storage address) and, if it has a sink, hand the sink to the kernel as
its output array.  Autograd then finds ``p.grad`` empty, keeps the returned tensor as ``p.grad`` -- which *is* the
slice of the flat buffer -- and nothing is copied.  Parameters that get their gradient through torch ops instead simply
miss the sink and are packed the usual way by the reducer.  A sink holds what ONE producer wrote: a parameter that feeds
two of this package's Functions at once would have its sink overwritten rather than summed, which is why the only
user, ``pipeline.TrainStep``, compares the result with a plain ``loss.backward()`` once before capturing anything.

Accumulating sinks (``accumulate=`` of ``grad_sinks``): destinations the kernels ADD to instead of overwriting -- the
running sum over the samples of a gradient-accumulation step (pipeline.BatchStep).  Only Functions whose kernels can add
in place look them up (``acc_sink_for``: the rasterizer's SH gradients, the keypoint blend's per-Gaussian gradients);
everything else keeps using the plain sink and is summed by the caller.
"""
import contextlib

_SINKS = {}
_ACC_SINKS = {}


def _key(t):
    return (t.data_ptr(), t.numel())


@contextlib.contextmanager
def grad_sinks(pairs, accumulate=()):
    """``pairs`` / ``accumulate``: iterables of (parameter, destination tensor of the same shape, contiguous float32)."""
    global _SINKS, _ACC_SINKS
    old, old_acc = _SINKS, _ACC_SINKS
    new, new_acc = dict(old), dict(old_acc)
    for table, items in ((new, pairs), (new_acc, accumulate)):
        for p, dst in items:
            assert dst.is_contiguous() and dst.shape == p.shape and dst.dtype == p.dtype
            table[_key(p)] = dst
    _SINKS, _ACC_SINKS = new, new_acc          # installed only once every pair has been checked
    try:
        yield
    finally:
        _SINKS, _ACC_SINKS = old, old_acc


def sink_for(t):
    """The destination registered for the tensor whose storage ``t`` is (None if there is none / no map installed)."""
    if not _SINKS or t is None or t.numel() == 0:
        return None
    return _SINKS.get(_key(t))


def acc_sink_for(t):
    """The accumulating destination registered for ``t`` (None if there is none)."""
    if not _ACC_SINKS or t is None or t.numel() == 0:
        return None
    return _ACC_SINKS.get(_key(t))


def out_like(sink, shape, **kw):
    """The sink if there is one (a fresh view object of it, so that autograd may adopt it as .grad), else a new tensor."""
    import torch
    if sink is not None:
        return sink.view(shape)
    return torch.empty(shape, **kw)
