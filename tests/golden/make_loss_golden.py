"""Generates tests/golden/loss_ref_*.npz by running the REFERENCE's own utils/loss_utils.py (imported from
/root/reference, CPU torch) on seeded inputs: l1, ssim, the train.py:113-114 combination and its autograd
gradient.  Run in the build container only (the GPU box has no /root/reference):
    python tests/golden/make_loss_golden.py
"""
import importlib.util
import os

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
spec = importlib.util.spec_from_file_location("ref_loss_utils", "/root/reference/utils/loss_utils.py")
ref = importlib.util.module_from_spec(spec)
spec.loader.exec_module(ref)


def make(tag, C, H, W, seed, lam, mode):
    g = torch.Generator().manual_seed(seed)
    gt = torch.rand(C, H, W, generator=g)
    if mode == "noise":
        img = torch.rand(C, H, W, generator=g)
    else:  # a rendering close to its target, with flat black regions (zero-variance windows)
        img = (gt + 0.05 * torch.randn(C, H, W, generator=g)).clamp(0, 1)
        img[:, : H // 3, : W // 4] = 0.0
        gt[:, : H // 3, : W // 4] = 0.0
    img.requires_grad_(True)
    l1 = ref.l1_loss(img, gt)
    ss = ref.ssim(img, gt)
    loss = (1.0 - lam) * l1 + lam * (1.0 - ss)
    loss.backward()
    np.savez_compressed(os.path.join(HERE, f"loss_ref_{tag}.npz"), image=img.detach().numpy(), gt=gt.numpy(),
                        lam=np.float32(lam), l1=l1.detach().numpy(), ssim=ss.detach().numpy(),
                        loss=loss.detach().numpy(), grad=img.grad.numpy())
    print(tag, float(loss), float(l1), float(ss))


if __name__ == "__main__":
    make("noise_3x37x53", 3, 37, 53, 1, 0.2, "noise")
    make("render_3x96x130", 3, 96, 130, 2, 0.2, "render")
    make("small_1x7x9", 1, 7, 9, 3, 0.5, "noise")
