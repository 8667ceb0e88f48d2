"""gsm_adam_step / gaussmart_b200.optim.FusedAdam against torch.optim.Adam in the reference's configuration
(scene/gaussian_model.py:283-292: six groups, lr=0.0 default, eps=1e-15; arguments/__init__.py:77-84 learning rates;
xyz learning rate rescheduled every iteration, gaussian_model.py:298-304).

Tolerance (floating point, written here): after every step, parameters and both moments within 2 ulp-ish of torch's
foreach path: |a - b| <= 3e-7 * max|b| element-wise ... in practice the result is bit-identical, which is also
asserted as a fraction (>= 99.9 % of the elements equal to the last bit)."""
import ctypes

import pytest
import torch
from torch import nn

from gaussmart_b200 import _lib
from gaussmart_b200.optim import FusedAdam

pytestmark = pytest.mark.gpu

LRS = {"xyz": 0.00016 * 5.0, "f_dc": 0.0025, "f_rest": 0.0025 / 20.0, "opacity": 0.05, "scaling": 0.005, "rotation": 0.001}
SHAPES = {"xyz": (3,), "f_dc": (1, 3), "f_rest": (15, 3), "opacity": (1,), "scaling": (2,), "rotation": (4,)}


def make_groups(P, seed):
    g = torch.Generator(device="cpu").manual_seed(seed)
    return [{"params": [nn.Parameter(torch.randn(P, *SHAPES[n], generator=g).cuda())], "lr": LRS[n], "name": n} for n in SHAPES]


def clone_groups(groups):
    return [{"params": [nn.Parameter(gr["params"][0].detach().clone())], "lr": gr["lr"], "name": gr["name"]} for gr in groups]


def set_grads(ga, gb, seed, sparsity=0.3):
    g = torch.Generator(device="cpu").manual_seed(seed)
    for a, b in zip(ga, gb):
        pa, pb = a["params"][0], b["params"][0]
        # magnitudes down to where g*g and the second moment are denormal (barely visible surfels)
        grad = torch.randn(pa.shape, generator=g).cuda() * (10.0 ** float(torch.randint(-22, 1, (1,), generator=g)))
        invisible = (torch.rand(pa.shape[0], generator=g) < sparsity).cuda()
        grad[invisible] = 0.0  # surfels outside the view: zero gradient, still updated by their momentum
        grad[: pa.shape[0] // 10] = 0.0  # surfels no view has touched yet: grad = exp_avg = exp_avg_sq = 0 throughout
        pa.grad, pb.grad = grad.clone(), grad.clone()


def compare(opt_a, opt_b, ga, gb):
    worst, same, total = 0.0, 0, 0
    for a, b in zip(ga, gb):
        pa, pb = a["params"][0], b["params"][0]
        sa, sb = opt_a.state[pa], opt_b.state[pb]
        assert float(sa["step"]) == float(sb["step"])
        for x, y in ((pa.detach(), pb.detach()), (sa["exp_avg"], sb["exp_avg"]), (sa["exp_avg_sq"], sb["exp_avg_sq"])):
            worst = max(worst, float((x - y).abs().max() / y.abs().max().clamp_min(1e-30)))
            same += int((x.view(torch.int32) == y.view(torch.int32)).sum())
            total += x.numel()
    return worst, same / total


@pytest.mark.parametrize("P", [1, 7, 1365, 1366, 100_003])
def test_fused_adam_tracks_torch_adam(P):
    ga = make_groups(P, 11)
    gb = clone_groups(ga)
    ours = FusedAdam(ga, lr=0.0, eps=1e-15)
    ref = torch.optim.Adam(gb, lr=0.0, eps=1e-15)
    for it in range(1, 13):
        set_grads(ga, gb, 100 + it)
        for opt in (ours, ref):  # update_learning_rate(iteration): only the xyz group changes
            for group in opt.param_groups:
                if group["name"] == "xyz":
                    group["lr"] = LRS["xyz"] * (0.97 ** it)
        ours.step()
        ref.step()
        worst, same = compare(ours, ref, ga, gb)
        assert worst <= 3e-7, (it, worst)
        assert same >= 0.999, (it, same)


def test_fused_adam_survives_the_reference_densification_surgery():
    """_prune_optimizer / cat_tensors_to_optimizer (gaussian_model.py:372-420) edit optimizer.state in place."""
    P = 5000
    ga = make_groups(P, 3)
    gb = clone_groups(ga)
    ours, ref = FusedAdam(ga, lr=0.0, eps=1e-15), torch.optim.Adam(gb, lr=0.0, eps=1e-15)
    for it in range(3):
        set_grads(ga, gb, it)
        ours.step(); ref.step()
    mask = (torch.rand(P, generator=torch.Generator().manual_seed(9)) < 0.7).cuda()
    for opt in (ours, ref):
        for group in opt.param_groups:
            old = group["params"][0]
            st = opt.state.get(old, None)
            ext = torch.full((17, *old.shape[1:]), 0.5, device="cuda")
            st["exp_avg"] = torch.cat((st["exp_avg"][mask], torch.zeros_like(ext)), dim=0)
            st["exp_avg_sq"] = torch.cat((st["exp_avg_sq"][mask], torch.zeros_like(ext)), dim=0)
            del opt.state[old]
            group["params"][0] = nn.Parameter(torch.cat((old.detach()[mask], ext), dim=0).requires_grad_(True))
            opt.state[group["params"][0]] = st
    ga, gb = ours.param_groups, ref.param_groups
    for it in range(3):
        set_grads(ga, gb, 50 + it)
        ours.step(); ref.step()
    worst, same = compare(ours, ref, ga, gb)
    assert worst <= 3e-7 and same >= 0.999


def test_params_without_grad_are_skipped_and_errors_are_loud():
    a = nn.Parameter(torch.ones(10).cuda())
    b = nn.Parameter(torch.ones(10).cuda())
    opt = FusedAdam([a, b], lr=0.1)
    a.grad = torch.ones(10).cuda()
    opt.step()
    assert torch.all(b == 1.0) and torch.all(a < 1.0) and b not in opt.state
    lib = _lib.load()
    t = _lib.AdamTensor(param=a.data_ptr(), grad=a.grad.data_ptr(), exp_avg=0, exp_avg_sq=0, n=10, lr=0.1, beta1=0.9,
                        beta2=0.999, eps=1e-8, step=1)
    assert lib.gsm_adam_step((_lib.AdamTensor * 1)(t), 1, None) == _lib.GSM_ERR_INVALID_ARGUMENT
    t.exp_avg = t.exp_avg_sq = a.data_ptr() + 4   # misaligned
    assert lib.gsm_adam_step((_lib.AdamTensor * 1)(t), 1, None) == _lib.GSM_ERR_INVALID_ARGUMENT
    assert lib.gsm_adam_step(None, 0, None) == _lib.GSM_OK
    with pytest.raises(RuntimeError):
        c = nn.Parameter(torch.ones(4))
        c.grad = torch.ones(4)
        FusedAdam([c], lr=0.1).step()   # no CPU path
    with pytest.raises(ValueError):
        FusedAdam([a], lr=0.1, amsgrad=True)
