"""Fused multi-tensor Adam: a drop-in for the optimizer GauSSmart builds at
scene/gaussian_model.py:292 -- `torch.optim.Adam(l, lr=0.0, eps=1e-15)` over the six surfel parameter groups -- whose
`step()` (train.py:212) is ONE hand-written kernel launch over all tensors (gsm_adam_step, csrc/adam.cu) instead of
torch's seven foreach passes.

The optimizer state has torch.optim.Adam's own layout (`state[p] = {"step", "exp_avg", "exp_avg_sq"}`, `param_groups`
with "lr", "betas", "eps", "name"), so the reference's densification code that edits it in place
(gaussian_model.py:355-420: replace_tensor_to_optimizer, _prune_optimizer, cat_tensors_to_optimizer) and
update_learning_rate (gaussian_model.py:298-304) run unchanged.  No CPU or torch fallback: CUDA fp32 parameters only.
"""
import ctypes

import torch

from . import _lib


class FusedAdam(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0, amsgrad=False):
        if weight_decay != 0 or amsgrad:
            raise ValueError("gaussmart_b200.FusedAdam covers the reference's configuration only: weight_decay=0, amsgrad=False")
        if lr < 0.0 or eps < 0.0 or not 0.0 <= betas[0] < 1.0 or not 0.0 <= betas[1] < 1.0:
            raise ValueError("invalid Adam hyper-parameters")
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps, weight_decay=0, amsgrad=False))

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        per_device = {}
        keep = []
        for group in self.param_groups:
            beta1, beta2 = group["betas"]
            for p in group["params"]:
                if p.grad is None:
                    continue
                g = p.grad
                if not p.is_cuda:
                    raise RuntimeError("gaussmart_b200.FusedAdam: parameters must be CUDA tensors (there is no CPU path)")
                if p.dtype != torch.float32 or g.dtype != torch.float32 or g.is_sparse:
                    raise RuntimeError("gaussmart_b200.FusedAdam: dense float32 parameters and gradients only")
                if not p.is_contiguous():
                    raise RuntimeError("gaussmart_b200.FusedAdam: parameters must be contiguous")
                if not g.is_contiguous():
                    g = g.contiguous()
                    keep.append(g)
                state = self.state[p]
                if len(state) == 0:  # torch/optim/adam.py _init_group
                    state["step"] = torch.tensor(0.0, dtype=torch.float32)
                    state["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                    state["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                state["step"] += 1
                t = _lib.AdamTensor(param=p.data_ptr(), grad=g.data_ptr(), exp_avg=state["exp_avg"].data_ptr(),
                                    exp_avg_sq=state["exp_avg_sq"].data_ptr(), n=p.numel(), lr=float(group["lr"]),
                                    beta1=float(beta1), beta2=float(beta2), eps=float(group["eps"]),
                                    step=int(state["step"].item()))
                per_device.setdefault(p.device, []).append(t)
        lib = _lib.load()
        for dev, tensors in per_device.items():
            arr = (_lib.AdamTensor * len(tensors))(*tensors)
            with torch.cuda.device(dev):
                stream = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
                _lib.check(lib.gsm_adam_step(arr, len(tensors), stream), "adam_step")
        return loss


@torch.no_grad()
def add_densification_stats(max_radii2D, xyz_gradient_accum, denom, radii, viewspace_grad, update_filter=None):
    """train.py:199-203 + GaussianModel.add_densification_stats (scene/gaussian_model.py:551-553) in one launch
    (gsm_densification_stats, csrc/adam.cu), in place on the model's three statistics tensors:

        max_radii2D[f] = max(max_radii2D[f], radii[f]);  xyz_gradient_accum[f] += ||viewspace_grad[f]||;  denom[f] += 1

    `update_filter` None means radii > 0 (render()'s visibility_filter).  A maintainer's call site:
        add_densification_stats(g.max_radii2D, g.xyz_gradient_accum, g.denom, radii, viewspace_point_tensor.grad)
    No CPU or torch fallback."""
    P = radii.shape[0]
    for name, t, dt, n in (("max_radii2D", max_radii2D, torch.float32, P), ("xyz_gradient_accum", xyz_gradient_accum, torch.float32, P),
                           ("denom", denom, torch.float32, P), ("radii", radii, torch.int32, P),
                           ("viewspace_grad", viewspace_grad, torch.float32, 3 * P)):
        if not t.is_cuda:
            raise RuntimeError(f"gaussmart_b200.add_densification_stats: {name} must be a CUDA tensor (there is no CPU path)")
        if t.dtype != dt or t.numel() != n or not t.is_contiguous():
            raise ValueError(f"gaussmart_b200.add_densification_stats: {name} must be contiguous {dt} with {n} elements")
    f = None
    if update_filter is not None:
        if not update_filter.is_cuda:
            raise RuntimeError("gaussmart_b200.add_densification_stats: update_filter must be a CUDA tensor")
        if update_filter.dtype not in (torch.bool, torch.uint8) or update_filter.numel() != P or not update_filter.is_contiguous():
            raise ValueError(f"gaussmart_b200.add_densification_stats: update_filter must be contiguous bool with {P} elements")
        f = update_filter.data_ptr()
    lib = _lib.load()
    with torch.cuda.device(radii.device):
        stream = ctypes.c_void_p(torch.cuda.current_stream(radii.device).cuda_stream)
        _lib.check(lib.gsm_densification_stats(radii.data_ptr(), viewspace_grad.data_ptr(), f, max_radii2D.data_ptr(),
                                               xyz_gradient_accum.data_ptr(), denom.data_ptr(), P, stream), "densification_stats")
