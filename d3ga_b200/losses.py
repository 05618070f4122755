"""Fused image losses of the D3GA training step, same names and meaning as
utils/loss_utils.py: ``l1_loss`` (:29-30), ``ssim`` (:59-90, window 11, sigma 1.5), plus the combination train.py:190-193
builds from them, ``l1_ssim_loss(pred, gt, lambda_dssim) = (1 - l) * l1 + l * (1 - ssim)``, as ONE forward and ONE
backward kernel (the reference spends five grouped convolutions and ~15 elementwise kernels each way)."""
from __future__ import annotations

import torch

from d3ga_b200 import _lib


class _SsimL1(torch.autograd.Function):
    @staticmethod
    def forward(ctx, pred, gt, w_ssim, w_l1):
        """Returns w_ssim * mean(SSIM map) + w_l1 * mean(|pred - gt|)."""
        lib = _lib.load()
        if not pred.is_cuda or not gt.is_cuda:
            raise RuntimeError("pred and gt must be CUDA tensors")
        if pred.dtype != torch.float32 or gt.dtype != torch.float32 or pred.shape != gt.shape:
            raise RuntimeError("pred and gt must be float32 tensors of the same shape")
        p, g = pred.contiguous(), gt.contiguous()
        H, W = int(p.shape[-2]), int(p.shape[-1])
        C = p.numel() // (H * W)
        dev = p.device
        need = ctx.needs_input_grad[0]
        with _lib.device_guard(dev):
            sums = torch.empty(2, dtype=torch.float32, device=dev)
            maps = torch.empty((3,) + tuple(p.shape), dtype=torch.float32, device=dev) if need else None
            _lib.check(lib.d3ga_ssim_l1_forward(
                C, H, W, p.data_ptr(), g.data_ptr(), sums.data_ptr(), _lib.ptr(maps[0]) if need else None,
                _lib.ptr(maps[1]) if need else None, _lib.ptr(maps[2]) if need else None, _lib.raw_stream(dev)),
                "ssim_l1_forward")
        ctx.save_for_backward(p, g, maps)
        ctx.w = (float(w_ssim), float(w_l1), p.numel())
        return (sums[0] * (w_ssim / p.numel()) + sums[1] * (w_l1 / p.numel()))

    @staticmethod
    def backward(ctx, grad_out):
        lib = _lib.load()
        p, g, maps = ctx.saved_tensors
        w_ssim, w_l1, n = ctx.w
        H, W = int(p.shape[-2]), int(p.shape[-1])
        C = p.numel() // (H * W)
        dev = p.device
        go = grad_out.to(torch.float32).reshape(1).contiguous()   # dL/dloss stays on the device: no host sync in backward
        with _lib.device_guard(dev):
            out = torch.empty_like(p)
            _lib.check(lib.d3ga_ssim_l1_backward(
                C, H, W, p.data_ptr(), g.data_ptr(), maps[0].data_ptr(), maps[1].data_ptr(), maps[2].data_ptr(),
                w_ssim / n, w_l1 / n, go.data_ptr(), out.data_ptr(), _lib.raw_stream(dev)), "ssim_l1_backward")
        return out, None, None, None


def ssim(img1, img2, window_size=11, size_average=True):
    """utils/loss_utils.py:59-68.  Only the configuration D3GA uses (window 11, mean over the map) is fused."""
    if window_size != 11 or not size_average:
        raise NotImplementedError("fused ssim supports window_size=11, size_average=True (what train.py:192 calls)")
    return _SsimL1.apply(img1, img2, 1.0, 0.0)


def l1_loss(network_output, gt):
    """utils/loss_utils.py:29-30."""
    return _SsimL1.apply(network_output, gt, 0.0, 1.0)


def l1_ssim_loss(pred, gt, lambda_dssim):
    """train.py:190-193: (1 - lambda) * l1_loss + lambda * (1 - ssim), one fused forward and backward."""
    return lambda_dssim + _SsimL1.apply(pred, gt, -float(lambda_dssim), 1.0 - float(lambda_dssim))
