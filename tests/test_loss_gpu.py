"""gsm_photometric_loss (csrc/loss.cu, fused L1 + SSIM value and gradient) against the torch fp32 restatement of
utils/loss_utils.py (oracle/loss_oracle.py: same op on the same device, autograd for the gradient) and against
the reference's own outputs (tests/golden/loss_ref_*.npz).

Tolerances (floating point, written here): loss / l1 / ssim within 2e-6 absolute of the golden values (fp32
reductions in a different order); gradient within 1e-4 relative to its largest entry + 1e-9."""
import ctypes
import glob
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import loss_oracle  # noqa: E402

from gaussmart_b200 import _lib, loss as gl  # noqa: E402

pytestmark = pytest.mark.gpu
GOLDEN = sorted(glob.glob(os.path.join(ROOT, "tests", "golden", "loss_ref_*.npz")))
VALUE_ATOL, GRAD_RTOL = 2e-6, 1e-4


def grad_close(got, want):
    return float((got - want).abs().max()) <= GRAD_RTOL * float(want.abs().max()) + 1e-9


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p) for p in GOLDEN])
def test_fused_loss_matches_reference_outputs(path):
    z = np.load(path)
    img, gt = torch.from_numpy(z["image"]).cuda(), torch.from_numpy(z["gt"]).cuda()
    out, grad = gl.photometric_loss_and_grad(img, gt, float(z["lam"]))
    out = out.cpu()
    assert abs(float(out[0]) - float(z["loss"])) <= VALUE_ATOL
    assert abs(float(out[1]) - float(z["l1"])) <= VALUE_ATOL
    assert abs(float(out[2]) - float(z["ssim"])) <= VALUE_ATOL
    assert grad_close(grad.cpu(), torch.from_numpy(z["grad"]))


@pytest.mark.parametrize("C,H,W", [(3, 1, 1), (3, 5, 300), (1, 32, 32), (3, 33, 31), (3, 64, 64), (3, 545, 980),
                                   (3, 1088, 1920)])
@pytest.mark.parametrize("lam", [0.2, 0.0, 1.0])
def test_fused_loss_matches_torch_restatement(C, H, W, lam):
    g = torch.Generator(device="cpu").manual_seed(C * 1000 + H + W)
    gt = torch.rand(C, H, W, generator=g).cuda()
    img = (gt + 0.1 * torch.randn(C, H, W, generator=g).cuda()).clamp(0, 1)
    if H > 8 and W > 8:  # exact zeros of |x - y| and zero-variance windows
        img[:, : H // 4, : W // 4] = gt[:, : H // 4, : W // 4] = 0.25
    ref_img = img.clone().requires_grad_(True)
    ref = loss_oracle.photometric_loss(ref_img, gt, lam)
    ref.backward()
    x = img.clone().requires_grad_(True)
    loss = gl.photometric_loss(x, gt, lam)
    (loss * 3.0).backward()  # upstream gradient is applied
    assert abs(float(loss.detach()) - float(ref.detach())) <= VALUE_ATOL + 1e-5 * abs(float(ref.detach()))
    assert grad_close(x.grad / 3.0, ref_img.grad)


def test_l1_and_ssim_wrappers():
    g = torch.Generator(device="cpu").manual_seed(5)
    a, b = torch.rand(3, 70, 90, generator=g).cuda(), torch.rand(3, 70, 90, generator=g).cuda()
    assert abs(float(gl.l1_loss(a, b)) - float(loss_oracle.l1_loss(a, b))) <= VALUE_ATOL
    assert abs(float(gl.ssim(a, b)) - float(loss_oracle.ssim(a, b))) <= VALUE_ATOL
    assert abs(float(gl.ssim(a, a)) - 1.0) <= VALUE_ATOL  # identical images


def test_loss_linearity_in_grad_scale_and_workspace_reuse():
    g = torch.Generator(device="cpu").manual_seed(6)
    a, b = torch.rand(3, 200, 333, generator=g).cuda(), torch.rand(3, 200, 333, generator=g).cuda()
    ws = gl.LossWorkspace()
    o1, g1 = gl.photometric_loss_and_grad(a, b, 0.2, 1.0, workspace=ws)
    o2, g2 = gl.photometric_loss_and_grad(a, b, 0.2, -2.5, workspace=ws)
    assert torch.equal(o1, o2)  # double-precision sums: the value does not depend on the atomics' order
    assert torch.allclose(g2, -2.5 * g1, rtol=1e-6, atol=1e-12)


def test_loss_argument_errors():
    lib = _lib.load()
    a = torch.rand(3, 8, 8).cuda()
    p = lambda t: ctypes.c_void_p(t.data_ptr())
    assert lib.gsm_photometric_loss(p(a), p(a), 3, 0, 8, 0.2, 1.0, None, p(a), None, 0, None) == _lib.GSM_ERR_INVALID_ARGUMENT
    assert lib.gsm_photometric_loss(p(a), p(a), 3, 8, 8, 0.2, 1.0, None, None, None, 0, None) == _lib.GSM_ERR_INVALID_ARGUMENT
    assert lib.gsm_photometric_loss(p(a), p(a), 3, 8, 8, 0.2, 1.0, None, p(a), None, 0, None) == _lib.GSM_ERR_WORKSPACE_TOO_SMALL
    with pytest.raises(RuntimeError):
        gl.photometric_loss(torch.rand(3, 8, 8), torch.rand(3, 8, 8))  # no CPU path
