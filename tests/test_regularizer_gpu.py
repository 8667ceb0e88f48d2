"""gsm_surface_regularizer (csrc/regularizer.cu: world-space normal, surf_depth, depth_to_normal, normal + distortion
loss, value and gradient w.r.t. allmap) against goldens made with the reference's depth_to_normal and against the torch
restatement (oracle/regularizer_oracle.py) run on the same device.

Tolerances (floating point, written here).  The normal comes from differences of back-projected points, so fp32
round-off is amplified by (depth / pixel footprint); the torch fp32 path itself is only that accurate.  Small goldens:
maps within 2e-5 abs, losses within 1e-6 abs + 1e-6 rel, gradient within 1e-4 of its largest entry.  Full size: our
error against an fp64 evaluation must not exceed 2x the torch fp32 path's own error against it (+1e-6)."""
import ctypes
import glob
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import regularizer_oracle as ro  # noqa: E402

from gaussmart_b200 import _lib, regularizer as gr, synthetic  # noqa: E402

pytestmark = pytest.mark.gpu
GOLDEN = sorted(glob.glob(os.path.join(ROOT, "tests", "golden", "regularizer_ref_*.npz")))


def cam_to(cam, device, dtype=torch.float32):
    return cam._replace(world_view_transform=cam.world_view_transform.to(device, dtype),
                        full_proj_transform=cam.full_proj_transform.to(device, dtype))


def synthetic_allmap(W, H, seed):
    g = torch.Generator().manual_seed(seed)
    yy, xx = torch.meshgrid(torch.arange(H).float(), torch.arange(W).float(), indexing="ij")
    alpha = 0.2 + 0.8 * torch.rand(H, W, generator=g)
    alpha[: H // 5, : W // 4] = 0.0
    depth = 3.0 + 0.4 * torch.sin(xx / 37.0) * torch.cos(yy / 25.0)
    nrm = torch.nn.functional.normalize(torch.randn(3, H, W, generator=g), dim=0)
    allmap = torch.stack([depth * alpha, alpha, nrm[0] * alpha, nrm[1] * alpha, nrm[2] * alpha, depth * 1.01,
                          0.01 * torch.rand(H, W, generator=g)]).float()
    allmap[5][alpha == 0] = 0.0
    return allmap


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p) for p in GOLDEN])
def test_fused_regularizer_matches_reference_goldens(path):
    z = np.load(path)
    cam = synthetic.make_camera(int(z["W"]), int(z["H"]), float(z["theta"]))
    allmap = torch.from_numpy(z["allmap"]).cuda()
    out, grad, maps = gr.surface_regularizer_and_grad(allmap, cam, float(z["depth_ratio"]), float(z["lambda_normal"]),
                                                      float(z["lambda_dist"]), want_maps=True)
    out = out.cpu()
    assert abs(float(out[1]) - float(z["normal_loss"])) <= 1e-6 + 1e-6 * abs(float(z["normal_loss"]))
    assert abs(float(out[2]) - float(z["dist_loss"])) <= 1e-6 + 1e-6 * abs(float(z["dist_loss"]))
    assert abs(float(out[0]) - float(z["normal_loss"]) - float(z["dist_loss"])) <= 2e-6 + 2e-6 * abs(float(out[0]))
    for k in ("surf_normal", "rend_normal", "surf_depth"):
        assert (maps[k].cpu() - torch.from_numpy(z[k])).abs().max() <= 2e-5, k
    g = torch.from_numpy(z["grad"])
    assert torch.isfinite(grad).all()
    assert (grad.cpu() - g).abs().max() <= 1e-4 * g.abs().max()


@pytest.mark.parametrize("W,H,ratio", [(3, 3, 0.0), (2, 5, 1.0), (64, 48, 0.0), (257, 131, 1.0), (1237, 822, 0.0)])
def test_fused_regularizer_no_worse_than_torch_fp32(W, H, ratio):
    cam = synthetic.make_camera(W, H, 0.7)
    allmap = synthetic_allmap(W, H, W + H).cuda()
    lam_n, lam_d = 0.05, 100.0
    a32 = allmap.clone().requires_grad_(True)
    n32, d32, m32 = ro.regularizer_terms(a32, cam_to(cam, "cuda"), ratio, lam_n, lam_d)
    (n32 + d32).backward()
    a64 = allmap.double().requires_grad_(True)
    n64, d64, m64 = ro.regularizer_terms(a64, cam_to(cam, "cuda", torch.float64), ratio, lam_n, lam_d)
    (n64 + d64).backward()
    x = allmap.clone().requires_grad_(True)
    loss = gr.surface_regularizer(x, cam, ratio, lam_n, lam_d)
    (loss * 2.0).backward()
    truth = torch.nan_to_num(a64.grad, 0.0, 0.0, 0.0)
    err_torch = (torch.nan_to_num(a32.grad, 0.0, 0.0, 0.0).double() - truth).abs().max()
    err_ours = (x.grad.double() / 2.0 - truth).abs().max()
    assert torch.isfinite(x.grad).all()
    assert float(err_ours) <= 2.0 * float(err_torch) + 1e-6 * float(truth.abs().max()) + 1e-12, (float(err_ours), float(err_torch))
    ref_loss = float((n64 + d64).detach())
    assert abs(float(loss.detach()) - ref_loss) <= 2.0 * abs(float((n32 + d32).detach()) - ref_loss) + 2e-6 * abs(ref_loss)
    _, _, maps = gr.surface_regularizer_and_grad(allmap, cam, ratio, lam_n, lam_d, want_maps=True)
    e_t = (m32["surf_normal"].detach().double() - m64["surf_normal"].detach()).abs().max()
    e_o = (maps["surf_normal"].double() - m64["surf_normal"].detach()).abs().max()
    assert float(e_o) <= 2.0 * float(e_t) + 1e-6


def test_regularizer_argument_errors():
    lib = _lib.load()
    assert lib.gsm_surface_regularizer(None, None) == _lib.GSM_ERR_INVALID_ARGUMENT
    a = _lib.RegularizerArgs(H=0, W=4)
    assert lib.gsm_surface_regularizer(ctypes.byref(a), None) == _lib.GSM_ERR_INVALID_ARGUMENT
    t = torch.zeros(7, 4, 4).cuda()
    a = _lib.RegularizerArgs(allmap=t.data_ptr(), H=4, W=4, dL_dallmap=t.data_ptr())
    assert lib.gsm_surface_regularizer(ctypes.byref(a), None) == _lib.GSM_ERR_WORKSPACE_TOO_SMALL
    with pytest.raises(RuntimeError):
        gr.surface_regularizer(torch.zeros(7, 4, 4), synthetic.make_camera(4, 4, 0.0))  # no CPU path
