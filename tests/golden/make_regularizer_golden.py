"""Generates tests/golden/regularizer_ref_*.npz with the REFERENCE's own utils/point_utils.py depth_to_normal (imported
from /root/reference) on seeded allmaps.  The reference module hard-codes `.cuda()` / device='cuda' and imports cv2 and
matplotlib at module level; this container has no GPU, so those are shimmed (Tensor.cuda -> identity, device kw dropped,
empty stand-in modules) -- the arithmetic that runs is the reference's.  The render() glue around it
(gaussian_renderer/__init__.py:117-147) and the loss terms (train.py:131-140) cannot be called in isolation; they come
from oracle/regularizer_oracle.py with the reference's depth_to_normal plugged in.
    python tests/golden/make_regularizer_golden.py
"""
import importlib.util
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
for name in ("cv2", "matplotlib", "matplotlib.pyplot"):
    sys.modules.setdefault(name, types.ModuleType(name))
torch.Tensor.cuda = lambda self, *a, **k: self
_arange = torch.arange
torch.arange = lambda *a, **k: _arange(*a, **{kk: vv for kk, vv in k.items() if kk != "device"})
spec = importlib.util.spec_from_file_location("ref_point_utils", "/root/reference/utils/point_utils.py")
ref = importlib.util.module_from_spec(spec)
spec.loader.exec_module(ref)

import regularizer_oracle as ro  # noqa: E402
from gaussmart_b200 import synthetic  # noqa: E402


def make(tag, W, H, theta, seed, depth_ratio, lam_n, lam_d):
    g = torch.Generator().manual_seed(seed)
    cam = synthetic.make_camera(W, H, theta)
    yy, xx = torch.meshgrid(torch.arange(H).float(), torch.arange(W).float(), indexing="ij")
    alpha = (0.2 + 0.8 * torch.rand(H, W, generator=g))
    alpha[: H // 5, : W // 4] = 0.0                       # background: nothing blended
    depth = 3.0 + 0.4 * torch.sin(xx / 7.0) * torch.cos(yy / 5.0) + 0.02 * torch.randn(H, W, generator=g)
    nrm = torch.nn.functional.normalize(torch.randn(3, H, W, generator=g), dim=0)
    allmap = torch.stack([depth * alpha, alpha, nrm[0] * alpha, nrm[1] * alpha, nrm[2] * alpha,
                          depth + 0.01 * torch.randn(H, W, generator=g), 0.01 * torch.rand(H, W, generator=g)]).float()
    allmap[5][alpha == 0] = 0.0
    allmap.requires_grad_(True)
    n_loss, d_loss, maps = ro.regularizer_terms(allmap, cam, depth_ratio, lam_n, lam_d, normal_fn=ref.depth_to_normal)
    (n_loss + d_loss).backward()
    grad = torch.nan_to_num(allmap.grad, 0.0, 0.0, 0.0)   # torch: 0/0 where alpha == 0
    np.savez_compressed(os.path.join(HERE, f"regularizer_ref_{tag}.npz"), allmap=allmap.detach().numpy(), W=W, H=H, theta=theta,
                        depth_ratio=np.float32(depth_ratio), lambda_normal=np.float32(lam_n), lambda_dist=np.float32(lam_d),
                        normal_loss=n_loss.detach().numpy(), dist_loss=d_loss.detach().numpy(),
                        surf_normal=maps["surf_normal"].detach().numpy(), rend_normal=maps["rend_normal"].detach().numpy(),
                        surf_depth=maps["surf_depth"].detach().numpy(), grad=grad.numpy())
    print(tag, float(n_loss.detach()), float(d_loss.detach()))


if __name__ == "__main__":
    make("expected_61x47", 61, 47, 0.3, 1, 0.0, 0.05, 100.0)
    make("median_96x64", 96, 64, 2.1, 2, 1.0, 0.05, 1000.0)
    make("mixed_33x40", 33, 40, 4.0, 3, 0.4, 0.1, 10.0)
