"""-m gpu parity tests of the fused L1 + SSIM loss through the C-ABI, against the vectors the
reference's own utils/loss_utils.py produced and against the torch restatement at the benchmark image size."""
import os

import numpy as np
import pytest
import torch

import parity_util as util
from d3ga_b200 import losses
from oracle import loss_oracle as lo

pytestmark = pytest.mark.gpu
G = np.load(os.path.join(util.GOLDEN, "loss_golden.npz"))


@pytest.mark.parametrize("name", list("abc"))
def test_fused_loss_matches_reference_golden(cuda_lib, name):
    pred = torch.from_numpy(G[f"{name}_pred"]).cuda().requires_grad_(True)
    gt = torch.from_numpy(G[f"{name}_gt"]).cuda()
    assert np.isclose(float(losses.ssim(pred, gt).detach()), float(G[f"{name}_ssim"]), atol=2e-6)
    assert np.isclose(float(losses.l1_loss(pred, gt).detach()), float(G[f"{name}_l1"]), atol=1e-6)
    loss = losses.l1_ssim_loss(pred, gt, 0.2)
    assert np.isclose(float(loss.detach()), float(G[f"{name}_loss"]), atol=2e-6)
    loss.backward()
    assert util.rel_l2(pred.grad.cpu().numpy(), G[f"{name}_grad"]) < 1e-4


def test_fused_loss_full_size_vs_oracle(cuda_lib):
    g = torch.Generator().manual_seed(3)
    gt = torch.rand(3, 1024, 1024, generator=g)
    pred = (gt + 0.1 * torch.randn(3, 1024, 1024, generator=g)).clamp(0, 1)
    p_cpu = pred.clone().requires_grad_(True)
    ref = lo.l1_ssim_loss(p_cpu, gt, 0.2)
    ref.backward()
    p_gpu = pred.cuda().requires_grad_(True)
    got = losses.l1_ssim_loss(p_gpu, gt.cuda(), 0.2)
    got.backward()
    assert abs(float(got.detach()) - float(ref.detach())) < 1e-5
    assert util.rel_l2(p_gpu.grad.cpu().numpy(), p_cpu.grad.numpy()) < 1e-4
    # the ssim drop-in refuses configurations it does not fuse
    with pytest.raises(NotImplementedError):
        losses.ssim(p_gpu, gt.cuda(), window_size=7)
