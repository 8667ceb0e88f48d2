"""GPU: gsm_photometric_loss_static -- the allocation-free variant of gsm_photometric_loss (csrc/loss.cu; scratch in a
file-scope `static __device__` array, limit GSM_PHOTOMETRIC_LOSS_STATIC_MAX_ELEMENTS pixel-channels) -- against the
existing entry point with a caller-owned workspace, through the C ABI, on two calls in a row that share the array.

No counterpart in the reference.  This file is named to be collected
LAST: a failure here must not hide any other test.

Tolerances (the ones tests/test_loss_gpu.py uses): loss / l1 / ssim within 2e-6 absolute; gradient within 1e-4 relative
to its largest entry + 1e-9.  (Both entry points launch the same two kernels, so the gradients are in fact expected to
agree bit for bit; the assertion is the stated tolerance.)"""
import ctypes
import re
import os

import pytest
import torch

from gaussmart_b200 import _lib

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
VALUE_ATOL, GRAD_RTOL = 2e-6, 1e-4


def _p(t):
    return ctypes.c_void_p(t.data_ptr())


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _images(C, H, W, seed):
    g = torch.Generator(device="cpu").manual_seed(seed)
    gt = torch.rand(C, H, W, generator=g).cuda()
    img = (gt + 0.1 * torch.randn(C, H, W, generator=g).cuda()).clamp(0, 1)
    if H > 8 and W > 8:  # exact zeros of |x - y| and zero-variance windows, as in test_loss_gpu.py
        img[:, : H // 4, : W // 4] = gt[:, : H // 4, : W // 4] = 0.25
    return img.contiguous(), gt.contiguous()


def _with_workspace(lib, img, gt, lam, scale):
    C, H, W = img.shape
    ws = torch.empty(lib.gsm_photometric_loss_workspace_bytes(C, H, W), dtype=torch.uint8, device=img.device)
    out, grad = torch.full((3,), -7.0, device=img.device), torch.full_like(img, -7.0)
    _lib.check(lib.gsm_photometric_loss(_p(img), _p(gt), C, H, W, lam, scale, _p(out), _p(grad), _p(ws), ws.numel(),
                                        _stream()), "photometric_loss")
    return out, grad


def _agree(got, want):
    (o1, g1), (o0, g0) = got, want
    assert torch.isfinite(o1).all() and torch.isfinite(g1).all()
    assert float((o1 - o0).abs().max()) <= VALUE_ATOL, (o1.tolist(), o0.tolist())
    assert float((g1 - g0).abs().max()) <= GRAD_RTOL * float(g0.abs().max()) + 1e-9


def test_static_scratch_loss_matches_the_workspace_entry_point_on_two_calls_in_a_row():
    lib = _lib.load()
    limit = lib.gsm_photometric_loss_static_max_elements()
    src = open(os.path.join(ROOT, "include", "gaussmart_b200.h")).read()
    assert limit == int(re.search(r"#define GSM_PHOTOMETRIC_LOSS_STATIC_MAX_ELEMENTS (\d+)", src).group(1))
    # call 1: the workload's image (BASELINE.json configs[3]: 1237x822); call 2: a smaller, ragged image with other
    # weights -- it finds call 1's sums and maps in the shared array and must not be affected by them; call 3: the first
    # image again after the small one
    cases = [(3, 822, 1237, 0.2, 1.0, 11), (3, 67, 101, 0.7, -2.5, 12), (3, 822, 1237, 0.2, 1.0, 11)]
    assert 3 * 822 * 1237 <= limit
    inputs = [_images(C, H, W, seed) for C, H, W, _, _, seed in cases]
    want = [_with_workspace(lib, img, gt, lam, scale) for (img, gt), (_, _, _, lam, scale, _) in zip(inputs, cases)]
    outs = [(torch.full((3,), -7.0, device="cuda"), torch.full_like(img, -7.0)) for img, _ in inputs]
    torch.cuda.synchronize()
    got, free = [], []
    for (img, gt), (C, H, W, lam, scale, _), (out, grad) in zip(inputs, cases, outs):
        free.append(torch.cuda.mem_get_info()[0])
        rc = lib.gsm_photometric_loss_static(_p(img), _p(gt), C, H, W, lam, scale, _p(out), _p(grad), _stream())
        _lib.check(rc, "photometric_loss_static")
        got.append((out, grad))
    torch.cuda.synchronize()
    free.append(torch.cuda.mem_get_info()[0])
    for g, w in zip(got, want):
        _agree(g, w)
    assert torch.equal(got[0][1], got[2][1]) and torch.equal(got[0][0], got[2][0])   # same inputs -> same bits
    # nothing was allocated by calls 2 and 3 (call 1 may load the module and run the one-time set-up of this device)
    assert free[2] >= free[1] and free[3] >= free[1], free


def test_static_scratch_loss_refuses_sizes_beyond_its_limit_and_bad_arguments():
    lib = _lib.load()
    limit = lib.gsm_photometric_loss_static_max_elements()
    img, gt = _images(3, 16, 16, 1)
    out, grad = torch.zeros(3, device="cuda"), torch.full_like(img, 5.0)
    H = limit // (3 * 2048) + 1                                         # 3 x H x 2048 > limit; rejected before any launch
    rc = lib.gsm_photometric_loss_static(_p(img), _p(gt), 3, H, 2048, 0.2, 1.0, _p(out), _p(grad), _stream())
    assert rc == _lib.GSM_ERR_WORKSPACE_TOO_SMALL
    msg = lib.gsm_last_error().decode()
    assert str(limit) in msg and "exceeds" in msg, msg
    assert lib.gsm_photometric_loss_static(_p(img), _p(gt), 3, 0, 16, 0.2, 1.0, _p(out), _p(grad), _stream()) == _lib.GSM_ERR_INVALID_ARGUMENT
    assert lib.gsm_photometric_loss_static(None, _p(gt), 3, 16, 16, 0.2, 1.0, _p(out), _p(grad), _stream()) == _lib.GSM_ERR_INVALID_ARGUMENT
    torch.cuda.synchronize()
    assert float(grad.min()) == 5.0 and float(grad.max()) == 5.0      # refused calls write nothing
    # the largest accepted size runs (a 1 x 2048 x 2048 image = exactly the limit) and agrees with the workspace form
    if limit == 2048 * 2048:
        big = _images(1, 2048, 2048, 3)
        o, g = torch.zeros(3, device="cuda"), torch.empty_like(big[0])
        _lib.check(lib.gsm_photometric_loss_static(_p(big[0]), _p(big[1]), 1, 2048, 2048, 0.2, 1.0, _p(o), _p(g), _stream()), "static")
        _agree((o, g), _with_workspace(lib, big[0], big[1], 0.2, 1.0))
