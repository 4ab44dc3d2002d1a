"""The blend-weight network of the keypoint stage: multiresolution hash-grid encoding + small ReLU MLP.

Mirrors ``self.weights_model = tcnn.NetworkWithInputEncoding(...)`` (scene/gaussian_model.py:370-392) and its call
``weights = self.weights_model(self.get_xyz.detach())`` (:257): input [N,3] -> logits [N, 2 * nearest_num].  The encoding
runs in csrc/hashgrid.cu, the MLP layer by layer on the tensor cores (csrc/linear_tc.cu).  tinycudann itself is not part of
the reference tree, so this implements the published algorithm with the reference's hyper-parameters; parity is
unpinned (oracle/hashgrid_oracle.py restates it on the CPU).

The input is time-invariant (canonical positions, detached): at evaluation time call ``freeze()`` once and every frame
reuses the logits (the reference recomputes them per frame; its own extrapolation loop hoists them, eval.py:125-126).
"""
import math

import torch
import torch.nn as nn

from . import _lib
from ._lib import check, ptr
from ._sinks import out_like, sink_for


class _HashEncode(torch.autograd.Function):
    @staticmethod
    def forward(ctx, xyz, table, cfg, order):
        L = _lib.lib()
        n_levels, base, scale, log2_t = cfg
        N = int(xyz.shape[0])
        xyz_c = xyz.detach().float().contiguous()
        enc = torch.empty((N, 4 * n_levels), dtype=torch.float32, device=xyz.device)
        check(L.gp_hashgrid_forward(N, n_levels, base, scale, log2_t, ptr(xyz_c), ptr(order), ptr(table.detach()), ptr(enc),
                                    _lib.stream_ptr()), "gp_hashgrid_forward")
        ctx.save_for_backward(xyz_c, order if order is not None else torch.empty(0, dtype=torch.int32, device=xyz.device))
        ctx.cfg, ctx.table_shape = cfg, tuple(table.shape)
        ctx.sink = sink_for(table)
        return enc

    @staticmethod
    def backward(ctx, g_enc):
        L = _lib.lib()
        xyz_c, order = ctx.saved_tensors
        order = order if order.numel() else None
        n_levels, base, scale, log2_t = ctx.cfg
        g_table = out_like(ctx.sink, ctx.table_shape, dtype=torch.float32, device=xyz_c.device)
        g_table.zero_()
        g = g_enc.contiguous()
        check(L.gp_hashgrid_backward(int(xyz_c.shape[0]), n_levels, base, scale, log2_t, ptr(xyz_c), ptr(order), ptr(g),
                                     ptr(g_table), _lib.stream_ptr()), "gp_hashgrid_backward")
        return None, g_table, None, None


class HashGridWeights(nn.Module):
    """``weights_model``: xyz [N,3] -> blend logits [N, n_out]."""

    def __init__(self, n_out=12, n_levels=16, base_resolution=16, finest_resolution=2048, log2_hashmap_size=19, width=64,
                 hidden_layers=2):
        super().__init__()
        per_level_scale = math.exp(math.log(finest_resolution / base_resolution) / (n_levels - 1))
        self.cfg = (int(n_levels), int(base_resolution), float(per_level_scale), int(log2_hashmap_size))
        entries = int(_lib.lib().gp_hashgrid_entries(*self.cfg))
        self.table = nn.Parameter(torch.empty(entries, 4).uniform_(-1e-4, 1e-4))      # tiny-cuda-nn's initialisation
        dims = [4 * n_levels] + [width] * hidden_layers + [n_out]
        self.weights = nn.ParameterList([nn.Parameter(nn.init.xavier_uniform_(torch.empty(dims[i + 1], dims[i])))
                                         for i in range(len(dims) - 1)])
        self.order = None          # optional Morton permutation of the points (deform.spatial_order): speed only
        self._frozen = None

    def freeze(self, xyz):
        """Evaluate once and serve the result from then on (evaluation: the input never changes)."""
        with torch.no_grad():
            self._frozen = None
            self._frozen = self.forward(xyz)
        return self._frozen

    def unfreeze(self):
        self._frozen = None

    def forward(self, xyz):
        if self._frozen is not None and self._frozen.shape[0] == xyz.shape[0]:
            return self._frozen
        if not xyz.is_cuda:
            raise TypeError("HashGridWeights needs CUDA tensors (there is no CPU path)")
        from .deform import _GaussianMLP
        order = self.order if (self.order is not None and self.order.shape[0] == xyz.shape[0]) else None
        enc = _HashEncode.apply(xyz, self.table, self.cfg, order)
        n = len(self.weights)
        return _GaussianMLP.apply(enc, n, *self.weights, *([None] * n))
