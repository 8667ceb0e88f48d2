"""oracle/loss_oracle.py (torch restatement of utils/loss_utils.py + train.py:113-114) against outputs of the
reference module itself (tests/golden/loss_ref_*.npz, made by tests/golden/make_loss_golden.py)."""
import glob
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import loss_oracle  # noqa: E402

GOLDEN = sorted(glob.glob(os.path.join(ROOT, "tests", "golden", "loss_ref_*.npz")))


def test_golden_fixtures_present():
    assert len(GOLDEN) >= 3


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p) for p in GOLDEN])
def test_loss_oracle_matches_reference_outputs(path):
    z = np.load(path)
    img = torch.from_numpy(z["image"]).requires_grad_(True)
    gt = torch.from_numpy(z["gt"])
    lam = float(z["lam"])
    loss = loss_oracle.photometric_loss(img, gt, lam)
    loss.backward()
    # same torch ops in the same order on the same CPU: equal to the last bits
    assert abs(float(loss_oracle.l1_loss(img.detach(), gt)) - float(z["l1"])) <= 1e-7
    assert abs(float(loss_oracle.ssim(img.detach(), gt)) - float(z["ssim"])) <= 1e-6
    assert abs(float(loss.detach()) - float(z["loss"])) <= 1e-6
    g = torch.from_numpy(z["grad"])
    assert (img.grad - g).abs().max() <= 1e-6 * g.abs().max() + 1e-12


def test_window_is_normalised_and_separable():
    g = loss_oracle.gaussian_1d()
    assert abs(float(g.sum()) - 1.0) < 1e-6
    w = loss_oracle.window_2d(3)
    assert w.shape == (3, 1, 11, 11)
    assert torch.allclose(w[1, 0], torch.outer(g, g), atol=1e-9)
