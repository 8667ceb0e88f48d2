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


def test_camera_constants_reproduce_the_reference_rays():
    """gaussmart_b200.regularizer.camera_constants (host side of gsm_surface_regularizer: the per-camera ray matrix,
    origin and view rotation passed by value) against depths_to_points' per-pixel rays (utils/point_utils.py:9-22)."""
    from gaussmart_b200.regularizer import camera_constants
    W, H = 57, 41
    cam = synthetic.make_camera(W, H, 1.3)
    ray, origin, rot = camera_constants(cam)
    dirs, o = ro.pixel_rays(cam, "cpu")
    M = torch.tensor(ray).view(3, 3)
    gx, gy = torch.meshgrid(torch.arange(W).float(), torch.arange(H).float(), indexing="xy")
    pix = torch.stack([gx, gy, torch.ones_like(gx)], dim=-1).reshape(-1, 3)
    assert (pix @ M - dirs).abs().max() <= 1e-6 * dirs.abs().max()
    assert torch.allclose(torch.tensor(origin), o, atol=1e-7)
    assert torch.equal(torch.tensor(rot).view(3, 3), cam.world_view_transform[:3, :3])
    assert camera_constants(cam) is camera_constants(cam)   # cached per camera


def test_camera_constants_cache_cannot_serve_another_camera():
    """the cache used to be keyed on data_ptr() / _version, which the caching allocator hands to the NEXT
    camera (MiniCam per GUI frame) -- it is keyed on the identity of the matrix tensors now, and keeps them alive."""
    from gaussmart_b200 import regularizer as gr
    W, H = 33, 21
    seen = []
    for th in (0.1, 0.9, 1.7, 0.1):
        cam = synthetic.make_camera(W, H, th)         # the previous camera's tensors are garbage by now
        ray = gr.camera_constants(cam)[0]
        dirs, _ = ro.pixel_rays(cam, "cpu")
        M = torch.tensor(ray).view(3, 3)
        pix = torch.tensor([[0.0, 0.0, 1.0], [W - 1.0, H - 1.0, 1.0]])
        want = torch.stack([dirs[0], dirs[-1]])
        assert (pix @ M - want).abs().max() <= 1e-6 * want.abs().max()
        seen.append(ray)
    assert seen[0] != seen[1] and seen[0] == seen[3]
    # an in-place pose update of the SAME tensor (version bump) must miss as well
    cam = synthetic.make_camera(W, H, 0.4)
    a = gr.camera_constants(cam)
    other = synthetic.make_camera(W, H, 1.1)
    cam.world_view_transform.copy_(other.world_view_transform)
    cam.full_proj_transform.copy_(other.full_proj_transform)
    b = gr.camera_constants(cam)
    assert a is not b and b[2] == gr.camera_constants(other)[2]


def test_host_mirrors_refuse_cpu_tensors_and_unsupported_settings():
    """No CPU fallback anywhere on the widened path: the host mirrors raise before touching the library."""
    from gaussmart_b200 import loss as gl, regularizer as gr
    from gaussmart_b200.optim import FusedAdam
    with pytest.raises(RuntimeError):
        gl.photometric_loss(torch.rand(3, 8, 8), torch.rand(3, 8, 8))
    with pytest.raises(RuntimeError):
        gr.surface_regularizer(torch.zeros(7, 4, 4), synthetic.make_camera(4, 4, 0.0))
    p = torch.nn.Parameter(torch.ones(4))
    with pytest.raises(ValueError):
        FusedAdam([p], weight_decay=0.1)
    with pytest.raises(ValueError):
        FusedAdam([p], amsgrad=True)
    p.grad = torch.ones(4)
    with pytest.raises(RuntimeError):
        FusedAdam([p], lr=0.1).step()
