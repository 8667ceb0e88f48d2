"""GPU: densify_and_prune as kernels (gaussmart_b200.densify, csrc/densify.cu; SURVEY.md 8f row 3, last part) against
the reference's own GaussianModel.densify_and_prune (scene/gaussian_model.py:535-549, install copy loaded by
tests/ref_callers.py) on the same model state, optimizer state and CUDA generator seed: same survivors in the same
order, parameters / Adam moments / segments bit-identical -- except the positions of split children, whose 3-term
rotation product goes through cuBLAS in the reference (summation order unspecified): 1e-6 relative."""
from argparse import ArgumentParser

import pytest
import torch

import ref_callers
from gaussmart_b200.densify import densify_and_prune

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not ref_callers.available(), reason="oracle/_ref (reference build + caller install copy) not present")]

NAMES = ("_xyz", "_features_dc", "_features_rest", "_opacity", "_scaling", "_rotation")


def _make_model(C, P, seed, extent, steps=True):
    """A GaussianModel mid-training: random parameters, an Adam state with history, densification statistics."""
    g = torch.Generator(device="cuda").manual_seed(seed)
    r = lambda *s: torch.randn(*s, device="cuda", generator=g)
    gm = C.GaussianModel(3)
    gm._xyz = torch.nn.Parameter(r(P, 3) * 2.0)
    gm._features_dc = torch.nn.Parameter(r(P, 1, 3) * 0.5)
    gm._features_rest = torch.nn.Parameter(r(P, 15, 3) * 0.05)
    gm._opacity = torch.nn.Parameter(r(P, 1) * 2.5 - 1.0)                      # sigmoid spans both sides of min_opacity
    gm._scaling = torch.nn.Parameter(r(P, 2) * 0.9 + torch.log(torch.tensor(0.01 * extent)).item())   # both sides of percent_dense * extent
    gm._rotation = torch.nn.Parameter(r(P, 4))
    gm._segments = torch.randint(0, 50, (P,), device="cuda", generator=g)
    gm.max_radii2D = torch.rand(P, device="cuda", generator=g) * 40.0
    opt = C.arguments.OptimizationParams(ArgumentParser())
    gm.spatial_lr_scale = 1.0
    gm.training_setup(opt)
    if steps:
        for _ in range(2):  # give Adam a history (exp_avg / exp_avg_sq / step)
            for n in NAMES:
                p = getattr(gm, n)
                p.grad = r(*p.shape) * 1e-3
            gm.optimizer.step()
    seen = torch.rand(P, device="cuda", generator=g) > 0.2
    gm.denom = (seen.float() * torch.randint(1, 40, (P,), device="cuda", generator=g).float())[:, None].contiguous()
    gm.xyz_gradient_accum = (gm.denom * torch.rand(P, 1, device="cuda", generator=g) * 4e-4).contiguous()   # ratio around max_grad = 2e-4; 0/0 where never seen
    return gm


def _snapshot(gm):
    st = {n: getattr(gm, n).detach().clone() for n in NAMES}
    st["segments"] = gm._segments.clone()
    st["max_radii2D"], st["accum"], st["denom"] = gm.max_radii2D.clone(), gm.xyz_gradient_accum.clone(), gm.denom.clone()
    for group in gm.optimizer.param_groups:
        s = gm.optimizer.state.get(group["params"][0])
        if s:
            st["m." + group["name"]], st["v." + group["name"]] = s["exp_avg"].clone(), s["exp_avg_sq"].clone()
            st["step." + group["name"]] = s["step"].clone() if torch.is_tensor(s["step"]) else torch.tensor(s["step"])
    return st


@pytest.mark.parametrize("P,size_threshold,steps", [(50_000, 20, True), (50_000, None, True), (4097, 20, False), (1_000_000, 20, True)])
def test_densify_and_prune_matches_reference(P, size_threshold, steps):
    C = ref_callers.load("reference")
    extent, max_grad, min_opacity = 4.4, 0.0002, 0.05
    a, b = _make_model(C, P, 77 + P, extent, steps), _make_model(C, P, 77 + P, extent, steps)
    before = _snapshot(a)
    torch.manual_seed(4321)
    a.densify_and_prune(max_grad, min_opacity, extent, size_threshold)          # the reference's statements
    torch.manual_seed(4321)
    P0, P1 = densify_and_prune(b, max_grad, min_opacity, extent, size_threshold)   # the kernels
    sa, sb = _snapshot(a), _snapshot(b)
    assert P0 == P and P1 == sa["_xyz"].shape[0]
    assert P1 != P and P1 > P // 4                               # clone, split and prune all happened
    for k in sa:
        assert sa[k].shape == sb[k].shape, (k, sa[k].shape, sb[k].shape)
    diff = (sa["_xyz"] != sb["_xyz"]).any(dim=1)
    # every tensor except the children's positions is bit-identical
    for k in sa:
        if k == "_xyz":
            continue
        assert torch.equal(sa[k], sb[k]), k
    rel = (sa["_xyz"] - sb["_xyz"]).abs().max().item() / sa["_xyz"].abs().max().item()
    assert rel <= 1e-6, rel
    # ... and positions differ (if at all) only where the Adam moments are zero and the scaling changed: new children
    zero_m = (sb["m.xyz"] == 0).all(dim=1) if "m.xyz" in sb else torch.ones(P1, dtype=torch.bool, device="cuda")
    assert (~diff | zero_m).all()
    # the optimizer really holds the new parameters (gaussian_model.py:413-466 surgery)
    for group in b.optimizer.param_groups:
        p = group["params"][0]
        assert p is getattr(b, dict(xyz="_xyz", f_dc="_features_dc", f_rest="_features_rest", opacity="_opacity",
                                    scaling="_scaling", rotation="_rotation")[group["name"]])
        assert p.requires_grad and p.shape[0] == P1
        if steps:
            assert b.optimizer.state[p]["exp_avg"].shape == p.shape
    assert sb["max_radii2D"].abs().max().item() == 0 and sb["accum"].abs().max().item() == 0 and sb["denom"].abs().max().item() == 0
    # training continues on the result: one more optimizer step on both, still identical apart from the children
    for gm in (a, b):
        for n in NAMES:
            p = getattr(gm, n)
            p.grad = torch.full_like(p, 1e-3)
        gm.optimizer.step()
    assert torch.equal(a._opacity, b._opacity) and torch.equal(a._features_rest, b._features_rest)
    del before


def test_densify_without_any_selection_is_a_pure_prune():
    C = ref_callers.load("reference")
    a, b = _make_model(C, 20_000, 5, 4.4), _make_model(C, 20_000, 5, 4.4)
    torch.manual_seed(1)
    a.densify_and_prune(1e9, 0.05, 4.4, None)      # max_grad out of reach: nothing cloned or split
    torch.manual_seed(1)
    densify_and_prune(b, 1e9, 0.05, 4.4, None)
    sa, sb = _snapshot(a), _snapshot(b)
    assert sa["_xyz"].shape[0] < 20_000
    for k in sa:
        assert torch.equal(sa[k], sb[k]), k
