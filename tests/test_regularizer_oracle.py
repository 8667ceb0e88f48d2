"""oracle/regularizer_oracle.py (torch restatement of render()'s allmap post-processing, depth_to_normal and the normal /
distortion loss terms) against goldens made with the reference's own depth_to_normal
(tests/golden/regularizer_ref_*.npz, tests/golden/make_regularizer_golden.py)."""
import glob
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import regularizer_oracle as ro  # noqa: E402

from gaussmart_b200 import synthetic  # noqa: E402

GOLDEN = sorted(glob.glob(os.path.join(ROOT, "tests", "golden", "regularizer_ref_*.npz")))


def test_golden_fixtures_present():
    assert len(GOLDEN) >= 3


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p) for p in GOLDEN])
def test_regularizer_oracle_matches_reference_depth_to_normal(path):
    z = np.load(path)
    cam = synthetic.make_camera(int(z["W"]), int(z["H"]), float(z["theta"]))
    allmap = torch.from_numpy(z["allmap"]).requires_grad_(True)
    n, d, m = ro.regularizer_terms(allmap, cam, float(z["depth_ratio"]), float(z["lambda_normal"]), float(z["lambda_dist"]))
    (n + d).backward()
    # same torch ops in the same order on the same CPU
    assert (m["surf_normal"].detach() - torch.from_numpy(z["surf_normal"])).abs().max() <= 1e-6
    assert (m["rend_normal"].detach() - torch.from_numpy(z["rend_normal"])).abs().max() <= 1e-6
    assert abs(float(n.detach()) - float(z["normal_loss"])) <= 1e-7
    assert abs(float(d.detach()) - float(z["dist_loss"])) <= 1e-6 * abs(float(z["dist_loss"]))
    g = torch.from_numpy(z["grad"])
    assert (torch.nan_to_num(allmap.grad, 0.0, 0.0, 0.0) - g).abs().max() <= 1e-6 * g.abs().max()
    # torch's backward of D/alpha is NaN exactly where alpha == 0 and nowhere else in channels 0 and 1
    assert torch.equal(torch.isnan(allmap.grad[0]), allmap.detach()[1] == 0)
