"""Generates tests/golden/loss_golden.npz by executing the REFERENCE's own utils/loss_utils.py (l1_loss, ssim) on CPU.
Its module-level imports that are absent here (kornia, torchvision, renderer, lib.cameras) are replaced by empty stub
modules; the functions executed are the reference's, unmodified.

    python tests/golden/make_loss_golden.py      (needs /root/reference)
"""
import importlib.util
import os
import sys
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
REF = os.environ.get("D3GA_REFERENCE", "/root/reference")
for name in ("kornia", "torchvision", "torchvision.transforms", "torchvision.utils", "torchvision.models", "lib",
             "lib.cameras", "renderer"):
    sys.modules.setdefault(name, types.ModuleType(name))
sys.modules["lib.cameras"].batch_to_camera = None
sys.modules["renderer"].render = None
tv = sys.modules["torchvision"]
tv.transforms, tv.utils, tv.models = (sys.modules["torchvision." + n] for n in ("transforms", "utils", "models"))
spec = importlib.util.spec_from_file_location("ref_loss_utils", os.path.join(REF, "utils", "loss_utils.py"))
ref = importlib.util.module_from_spec(spec)
spec.loader.exec_module(ref)

g = torch.Generator().manual_seed(11)
out = {}
for name, (H, W) in {"a": (37, 50), "b": (16, 16), "c": (64, 33)}.items():
    gt = torch.rand(3, H, W, generator=g)
    pred = (gt + 0.2 * torch.randn(3, H, W, generator=g)).clamp(0, 1).requires_grad_(True)
    s = ref.ssim(pred, gt)
    l1 = ref.l1_loss(pred, gt)
    lam = 0.2
    loss = (1.0 - lam) * l1 + lam * (1.0 - s)      # train.py:193
    (grad,) = torch.autograd.grad(loss, pred)
    out.update({f"{name}_pred": pred.detach().numpy(), f"{name}_gt": gt.numpy(), f"{name}_ssim": s.detach().numpy(),
                f"{name}_l1": l1.detach().numpy(), f"{name}_loss": loss.detach().numpy(), f"{name}_grad": grad.numpy()})
out["window"] = ref.gaussian(11, 1.5).numpy()
np.savez(os.path.join(ROOT, "tests", "golden", "loss_golden.npz"), **out)
print("wrote loss_golden.npz", {k: float(v) for k, v in out.items() if k.endswith("_ssim")})
