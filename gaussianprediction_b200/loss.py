"""Image loss of the train iteration on fused B200 kernels.

Mirrors utils/loss_utils.py: ``l1_loss`` (:54-55), ``ssim`` (:70-100, window 11, sigma 1.5, zero padding,
``size_average=True``) and the combination of train.py:108,
``loss = (1 - lambda_dssim) * l1_loss(image, gt) + lambda_dssim * (1 - ssim(image, gt))``.
One forward launch computes both means and the derivative maps, one backward launch the image gradient; the
reference runs five grouped 11x11 convolutions and ~15 elementwise kernels forward plus their autograd backward.
"""
import torch

from . import _lib
from ._lib import check, ptr


def _planes(img):
    if not img.is_cuda:
        raise RuntimeError("gaussianprediction_b200.loss needs CUDA tensors (there is no CPU path)")
    if img.dim() < 2:
        raise ValueError("image must have at least 2 dimensions")
    H, W = int(img.shape[-2]), int(img.shape[-1])
    C = int(img.numel() // (H * W)) if H * W else 0
    return C, H, W


class _ImageLoss(torch.autograd.Function):
    """w_l1 * mean|img - gt| + w_ssim * mean SSIM(img, gt) + const; differentiable w.r.t. ``img`` only (like the
    reference's use: gt is data)."""

    @staticmethod
    def forward(ctx, img, gt, w_l1, w_ssim, const):
        if img.shape != gt.shape:
            raise ValueError("image and target must have the same shape")
        L = _lib.lib()
        C, H, W = _planes(img)
        x = img.detach().float().contiguous()
        y = gt.detach().float().contiguous()
        dev = x.device
        need_bwd = ctx.needs_input_grad[0]
        dmaps = torch.empty((3 * C * H * W,), dtype=torch.float32, device=dev) if need_bwd else None
        ws = torch.empty((L.gp_image_loss_ws_bytes(C, H, W),), dtype=torch.uint8, device=dev)
        sums = torch.empty((2,), dtype=torch.float32, device=dev)
        check(L.gp_image_loss_forward(C, H, W, ptr(x), ptr(y), ptr(dmaps), ptr(ws), ptr(sums), _lib.stream_ptr()),
              "gp_image_loss_forward")
        ctx.save_for_backward(x, y, dmaps)
        ctx.cfg = (C, H, W, float(w_l1), float(w_ssim), img.shape)
        return w_l1 * sums[0] + w_ssim * sums[1] + const

    @staticmethod
    def backward(ctx, g):
        x, y, dmaps = ctx.saved_tensors
        C, H, W, w_l1, w_ssim, shape = ctx.cfg
        grad = torch.empty_like(x)
        gs = g.detach().float().contiguous().reshape(1)
        check(_lib.lib().gp_image_loss_backward(C, H, W, ptr(x), ptr(y), ptr(dmaps), w_l1, w_ssim, ptr(gs), ptr(grad),
                                                _lib.stream_ptr()), "gp_image_loss_backward")
        return grad.view(shape), None, None, None, None


def l1_loss(network_output, gt):
    return _ImageLoss.apply(network_output, gt, 1.0, 0.0, 0.0)


def ssim(img1, img2, window_size=11, size_average=True):
    if window_size != 11 or not size_average:
        raise NotImplementedError("only the reference's use is implemented: window_size=11, size_average=True")
    return _ImageLoss.apply(img1, img2, 0.0, 1.0, 0.0)


def image_loss(image, gt, lambda_dssim=0.2):
    """train.py:108: (1 - lambda) * L1 + lambda * (1 - SSIM), as one fused forward/backward pair."""
    return _ImageLoss.apply(image, gt, 1.0 - lambda_dssim, -lambda_dssim, lambda_dssim)
