"""oracle/loss_oracle.py -- TEST INFRASTRUCTURE ONLY (torch CPU; autograd gives the backward).

Restatement of the image losses of utils/loss_utils.py: l1_loss (:29-30), gaussian / create_window (:46-55) and
ssim / _ssim (:59-90).  Pinned against the reference's own functions: tests/golden/make_loss_golden.py executes
utils/loss_utils.py from /root/reference (stub modules for its kornia / torchvision / renderer imports) and stores its
outputs and autograd gradients in tests/golden/loss_golden.npz; tests/test_oracle_loss.py checks this file against them.
"""
from math import exp

import torch as th
import torch.nn.functional as F


def l1_loss(network_output, gt):
    return th.abs(network_output - gt).mean()                                   # loss_utils.py:29-30


def gaussian(window_size, sigma):
    gauss = th.Tensor([exp(-((x - window_size // 2) ** 2) / float(2 * sigma ** 2)) for x in range(window_size)])
    return gauss / gauss.sum()                                                   # loss_utils.py:46-48


def create_window(window_size, channel):
    _1d = gaussian(window_size, 1.5).unsqueeze(1)
    _2d = _1d.mm(_1d.t()).float().unsqueeze(0).unsqueeze(0)
    return _2d.expand(channel, 1, window_size, window_size).contiguous()        # loss_utils.py:51-55


def ssim(img1, img2, window_size=11, size_average=True):
    channel = img1.size(-3)
    window = create_window(window_size, channel).type_as(img1)
    pad = window_size // 2
    mu1 = F.conv2d(img1, window, padding=pad, groups=channel)                    # loss_utils.py:71-72
    mu2 = F.conv2d(img2, window, padding=pad, groups=channel)
    mu1_sq, mu2_sq, mu1_mu2 = mu1.pow(2), mu2.pow(2), mu1 * mu2
    sigma1_sq = F.conv2d(img1 * img1, window, padding=pad, groups=channel) - mu1_sq
    sigma2_sq = F.conv2d(img2 * img2, window, padding=pad, groups=channel) - mu2_sq
    sigma12 = F.conv2d(img1 * img2, window, padding=pad, groups=channel) - mu1_mu2
    C1, C2 = 0.01 ** 2, 0.03 ** 2
    ssim_map = ((2 * mu1_mu2 + C1) * (2 * sigma12 + C2)) / ((mu1_sq + mu2_sq + C1) * (sigma1_sq + sigma2_sq + C2))
    return ssim_map.mean() if size_average else ssim_map.mean(1).mean(1).mean(1)  # loss_utils.py:85-90


def l1_ssim_loss(pred, gt, lambda_dssim):
    """train.py:190-193."""
    return (1.0 - lambda_dssim) * l1_loss(pred, gt) + lambda_dssim * (1.0 - ssim(pred, gt))
