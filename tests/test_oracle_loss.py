"""CPU tests pinning oracle/loss_oracle.py against vectors produced by the REFERENCE's own utils/loss_utils.py
(tests/golden/loss_golden.npz, see tests/golden/make_loss_golden.py)."""
import os

import numpy as np
import torch as th

import parity_util as util
from oracle import loss_oracle as lo

G = np.load(os.path.join(util.GOLDEN, "loss_golden.npz"))


def test_window_matches_reference():
    assert np.array_equal(lo.gaussian(11, 1.5).numpy(), G["window"])


def test_losses_and_gradients_match_reference():
    for name in "abc":
        pred = th.from_numpy(G[f"{name}_pred"]).requires_grad_(True)
        gt = th.from_numpy(G[f"{name}_gt"])
        assert np.isclose(float(lo.ssim(pred, gt).detach()), float(G[f"{name}_ssim"]), atol=1e-6)
        assert np.isclose(float(lo.l1_loss(pred, gt).detach()), float(G[f"{name}_l1"]), atol=1e-7)
        loss = lo.l1_ssim_loss(pred, gt, 0.2)
        assert np.isclose(float(loss.detach()), float(G[f"{name}_loss"]), atol=1e-6)
        (grad,) = th.autograd.grad(loss, pred)
        assert util.rel_l2(grad.numpy(), G[f"{name}_grad"]) < 1e-5
