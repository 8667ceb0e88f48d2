"""Raw-parameter entry (gsm_forward_args.raw_params / shs_rest, rasterize_gaussians_raw): the getters of
scene/gaussian_model.py:104-123 (exp, F.normalize, sigmoid, torch.cat of features_dc / features_rest) fused into the
preprocess kernels, against the same getters as torch ops in front of the standard entry point (+ autograd behind it).

Bars: radii identical, forward outputs within 1e-5 abs (the activations are evaluated as torch's kernels evaluate them,
so in practice every bit agrees -- also asserted as a fraction), gradients w.r.t. the RAW parameters within 1e-4 rel
(helpers.grad_violations; fp32 atomics order differs between two backward runs)."""
import pytest
import torch

import gaussmart_b200 as gb
from gaussmart_b200 import synthetic
from gaussmart_b200.rasterizer import rasterize_gaussians_raw
from helpers import fwd_close, grad_violations

pytestmark = pytest.mark.gpu


def raw_scene(P, W, deg, seed):
    scene = synthetic.make_scene(P, W, deg, seed)
    scene = scene._replace(scales=scene.scales * 2.0)
    g = torch.Generator().manual_seed(seed + 1)
    rot_raw = scene.rotations * (0.2 + 3.0 * torch.rand(P, 1, generator=g))     # unnormalised, as _rotation is stored
    o = scene.opacities.clamp(1e-4, 1 - 1e-4)
    return dict(xyz=scene.means3D.cuda(), f_dc=scene.shs[:, :1].contiguous().cuda(), f_rest=scene.shs[:, 1:].contiguous().cuda(),
                opacity=torch.log(o / (1 - o)).cuda(), scaling=torch.log(scene.scales).cuda(), rotation=rot_raw.cuda())


def getters(raw):
    # scene/gaussian_model.py:104-123
    return dict(means3D=raw["xyz"], shs=torch.cat((raw["f_dc"], raw["f_rest"]), dim=1), opacities=torch.sigmoid(raw["opacity"]),
                scales=torch.exp(raw["scaling"]), rotations=torch.nn.functional.normalize(raw["rotation"]))


@pytest.mark.parametrize("P,W,H,deg", [(20000, 200, 150, 3), (4999, 96, 64, 1), (3000, 64, 64, 0)])
def test_raw_entry_matches_torch_getters(P, W, H, deg):
    raw0 = raw_scene(P, W, deg, 5 + deg)
    cam = synthetic.make_camera(W, H, 0.4)
    cam_d = synthetic.Camera(*[x.cuda() if isinstance(x, torch.Tensor) else x for x in cam])
    S = gb.GaussianRasterizationSettings(**synthetic.settings_kwargs(cam_d, deg, torch.zeros(3).cuda()))
    cot = synthetic.make_cotangents(W, H, 3)
    cot = (cot[0].cuda(), cot[1].cuda())

    ra = {k: v.clone().requires_grad_(True) for k, v in raw0.items()}
    m2a = torch.zeros_like(ra["xyz"], requires_grad=True)
    ca, rada, alla = gb.GaussianRasterizer(S)(means2D=m2a, **getters(ra))
    torch.autograd.backward([ca, alla], list(cot))

    rb = {k: v.clone().requires_grad_(True) for k, v in raw0.items()}
    m2b = torch.zeros_like(rb["xyz"], requires_grad=True)
    cb, radb, allb = rasterize_gaussians_raw(rb["xyz"], m2b, rb["f_dc"], rb["f_rest"], rb["opacity"], rb["scaling"],
                                             rb["rotation"], S)
    torch.autograd.backward([cb, allb], list(cot))

    assert torch.equal(rada, radb)
    assert bool(fwd_close(cb, ca).all()) and bool(fwd_close(allb, alla).all())
    same = float((cb == ca).float().mean())
    assert same >= 0.999, same
    for k in ("xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation"):
        if deg == 0 and k == "f_rest":
            continue
        bad, scale = grad_violations(rb[k].grad, ra[k].grad)
        assert bad <= 1e-4 * rb[k].numel(), (k, bad, scale)   # (two backward runs differ by their atomics' order)
    assert grad_violations(m2b.grad, m2a.grad)[0] <= 1e-4 * m2b.numel()
    if deg == 0:
        assert rb["f_rest"].grad is None or rb["f_rest"].numel() == 0 or float(rb["f_rest"].grad.abs().max()) == 0.0


def test_raw_entry_inference_equals_training_forward():
    raw0 = raw_scene(8000, 128, 3, 21)
    cam = synthetic.make_camera(128, 96, 1.1)
    cam_d = synthetic.Camera(*[x.cuda() if isinstance(x, torch.Tensor) else x for x in cam])
    S = gb.GaussianRasterizationSettings(**synthetic.settings_kwargs(cam_d, 3, torch.zeros(3).cuda()))
    args = (raw0["xyz"], torch.zeros_like(raw0["xyz"]), raw0["f_dc"], raw0["f_rest"], raw0["opacity"], raw0["scaling"],
            raw0["rotation"])
    with torch.no_grad():
        c0, r0, a0 = rasterize_gaussians_raw(*args, S)
    leaves = [t.clone().requires_grad_(True) for t in args]
    c1, r1, a1 = rasterize_gaussians_raw(*leaves, S)
    assert torch.equal(c0, c1.detach()) and torch.equal(r0, r1) and torch.equal(a0, a1.detach())
