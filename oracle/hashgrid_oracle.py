"""CPU restatement (torch / numpy) of the blend-weight network: multiresolution hash-grid encoding + 64-wide ReLU MLP.

TEST INFRASTRUCTURE ONLY (tests/): never imported by gaussianprediction_b200/.

PARITY UNPINNED.  The reference gets this network from tinycudann (``tcnn.NetworkWithInputEncoding``,
scene/gaussian_model.py:370-392, called at :257), installed from an unpinned fork
(``BoMingZhao/tiny-cuda-nn-float32``, scripts/utils/env.sh:7) that is neither in /root/reference nor in this image, and
the reference holds no vector of its output.  What is restated here is the PUBLISHED algorithm (Mueller et al.,
"Instant Neural Graphics Primitives with a Multiresolution Hash Encoding", 2022; tiny-cuda-nn's grid encoding) with
the reference's hyper-parameters: L = 16, F = 4, T = 2^19, N_min = 16, b = exp(ln(2048/16)/15), linear interpolation,
then FullyFusedMLP 64 x 2 hidden, ReLU, no biases, linear output of 2 * nearest_num.
"""
import numpy as np
import torch

PRIMES = (1, 2654435761, 805459861)


def levels(n_levels=16, base_resolution=16, per_level_scale=float(np.exp(np.log(2048 / 16) / 15)), log2_hashmap_size=19):
    """Per level: (scale, resolution, offset, size); float32 arithmetic like the CUDA host code."""
    out, off = [], 0
    cap = 1 << log2_hashmap_size
    for l in range(n_levels):
        b = float(np.float32(per_level_scale))       # the scale crosses the C ABI as a float
        scale = np.float32(np.exp2(float(l) * np.log2(b)) * float(base_resolution) - 1.0)   # double, rounded once
        res = int(np.ceil(scale)) + 1
        dense = (res ** 3 + 7) // 8 * 8
        size = min(dense, cap)
        out.append((float(scale), res, off, size))
        off += size
    return out, off


def grid_index(c, res, size):
    """c: int64 [...,3] holding uint32 cell coordinates (wrapped).  Dense while the stride fits, else the spatial hash."""
    M = (1 << 32) - 1
    x, y, z = c[..., 0] & M, c[..., 1] & M, c[..., 2] & M
    stride, dense_dims = 1, 0
    idx = torch.zeros_like(x)
    for d, v in enumerate((x, y, z)):
        if stride > size:
            break
        idx = (idx + v * stride) & M
        stride *= res
        dense_dims = d + 1
    if size < stride:
        idx = (x * PRIMES[0]) & M
        idx = idx ^ ((y * PRIMES[1]) & M)
        idx = idx ^ ((z * PRIMES[2]) & M)
    return idx % size


def encode(xyz, table, **kw):
    """[N,3] float32 -> [N, 4 L]; ``table`` [entries, 4] (requires_grad for the table gradient)."""
    lv, total = levels(**kw)
    assert table.shape[0] == total
    outs = []
    for scale, res, off, size in lv:
        # fmaf(scale, x, 0.5) as tiny-cuda-nn's pos_fract does it: one rounding (emulated in float64).  At the finest
        # levels pos ~ 2000, so the fused and the two-step forms differ by ~1e-4 in the interpolation weight.
        pos = (xyz.double() * float(np.float32(scale)) + 0.5).float()
        cell = torch.floor(pos)
        w = pos - cell
        c = cell.to(torch.int64)                    # (uint32)(int)floor: negative cells wrap modulo 2^32 in grid_index
        acc = torch.zeros(xyz.shape[0], 4, dtype=table.dtype)
        for k in range(8):
            d = torch.tensor([k & 1, (k >> 1) & 1, k >> 2])
            wt = torch.where(d.bool(), w, 1 - w).prod(dim=1, keepdim=True)
            idx = grid_index(c + d, res, size)
            acc = acc + wt * table[off + idx]
        outs.append(acc)
    return torch.cat(outs, dim=1)


def network(xyz, table, w0, w1, w2, **kw):
    h = encode(xyz, table, **kw)
    h = torch.relu(h @ w0.t())
    h = torch.relu(h @ w1.t())
    return h @ w2.t()
