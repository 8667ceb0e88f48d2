"""Fused photometric loss of the training iteration.

Mirrors the reference's call sites: `l1_loss(image, gt)` / `ssim(image, gt)` (utils/loss_utils.py:16-17, 32-62)
combined at train.py:113-114 as `(1 - lambda_dssim) * Ll1 + lambda_dssim * (1 - ssim(image, gt_image))`.
One C-ABI call (gsm_photometric_loss: two hand-written kernels, csrc/loss.cu) produces the loss AND
dloss/dimage; the autograd wrapper only hands the stored gradient back.  No CPU or torch fallback:
a missing library or a non-CUDA tensor raises.
"""
import ctypes

import torch

from . import _lib


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr())


class LossWorkspace:
    """Caller-owned scratch of the fused loss (sums + three partial-derivative maps), reusable across iterations."""

    def __init__(self):
        self.buf = None

    def get(self, nbytes, device):
        if self.buf is None or self.buf.numel() < nbytes or self.buf.device != device:
            self.buf = torch.empty(nbytes, dtype=torch.uint8, device=device)
        return self.buf


_default_ws = LossWorkspace()


def photometric_loss_and_grad(image, gt, lambda_dssim=0.2, grad_scale=1.0, out=None, grad=None, workspace=None):
    """(out[3] = loss, l1, ssim; grad[C,H,W] = grad_scale * dloss/dimage), both on the device, no sync."""
    lib = _lib.load()
    for name, t in (("image", image), ("gt", gt)):
        if not t.is_cuda:
            raise RuntimeError(f"gaussmart_b200: {name} must be a CUDA tensor (there is no CPU path)")
        if t.dtype != torch.float32 or t.dim() != 3 or not t.is_contiguous():
            raise RuntimeError(f"{name} must be a contiguous float32 [C,H,W] tensor")
    if image.shape != gt.shape:
        raise RuntimeError("image and gt must have the same shape")
    C, H, W = image.shape
    dev = image.device
    if out is None:
        out = torch.empty(3, dtype=torch.float32, device=dev)
    if grad is None:
        grad = torch.empty_like(image)
    ws = (workspace or _default_ws).get(lib.gsm_photometric_loss_workspace_bytes(C, H, W), dev)
    with torch.cuda.device(dev):
        stream = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        _lib.check(lib.gsm_photometric_loss(_ptr(image), _ptr(gt), C, H, W, float(lambda_dssim), float(grad_scale),
                                            _ptr(out), _ptr(grad), _ptr(ws), ws.numel(), stream), "photometric_loss")
    return out, grad


class _PhotometricLoss(torch.autograd.Function):
    @staticmethod
    def forward(ctx, image, gt, lambda_dssim):
        out, grad = photometric_loss_and_grad(image.contiguous(), gt.contiguous(), lambda_dssim)
        ctx.save_for_backward(grad)
        return out[0]

    @staticmethod
    def backward(ctx, g):
        (grad,) = ctx.saved_tensors
        return grad * g, None, None


def photometric_loss(image, gt, lambda_dssim=0.2):
    """Scalar `(1 - lambda) * l1_loss(image, gt) + lambda * (1 - ssim(image, gt))`, differentiable w.r.t. image."""
    return _PhotometricLoss.apply(image, gt, lambda_dssim)


def l1_loss(image, gt):
    """utils/loss_utils.py:16-17 through the fused kernel (lambda = 0)."""
    return _PhotometricLoss.apply(image, gt, 0.0)


def ssim(image, gt):
    """utils/loss_utils.py:32-62 (window 11, size_average=True) through the fused kernel: loss at lambda = 1 is 1 - ssim."""
    return 1.0 - _PhotometricLoss.apply(image, gt, 1.0)
