"""densify_and_prune as stream-compaction kernels (last part): a drop-in for
GaussianModel.densify_and_prune (scene/gaussian_model.py:535-549), the call train.py:205-208 makes every
`densification_interval` iterations:

    from gaussmart_b200.densify import densify_and_prune
    densify_and_prune(gaussians, opt.densify_grad_threshold, opt.opacity_cull, scene.cameras_extent, size_threshold)

Same arguments, same effect on the model object: `_xyz, _features_dc, _features_rest, _opacity, _scaling, _rotation`
become new nn.Parameters holding [kept originals | clones | first split children | second split children] -- the order
the reference's clone / split / prune sequence leaves --, the optimizer's param_groups and state (`exp_avg`,
`exp_avg_sq`, `step`) are re-pointed exactly as cat_tensors_to_optimizer / _prune_optimizer do (gaussian_model.py:413-466),
`xyz_gradient_accum`, `denom`, `max_radii2D` are reset to zeros and `_segments` follows the surfels.  The reference runs
~60 torch ops with a `nonzero` host sync per boolean-mask gather and copies every parameter and both Adam moments three
times; here: two launches, ONE host read (the new surfel count sizes the new tensors), one gather launch over all 19
tensors (csrc/densify.cu).  The split children's random offsets come from the SAME draw the reference makes
(`torch.normal(mean, std)` = N(0,1) of that shape from the current CUDA generator, scaled in-kernel), so a seeded run
produces the same surfels.  No CPU or torch fallback.
"""
import ctypes

import numpy as np
import torch
from torch import nn

from . import _lib

# name of the optimizer group -> attribute of GaussianModel (gaussian_model.py:283-290), in group order
_GROUPS = (("xyz", "_xyz"), ("f_dc", "_features_dc"), ("f_rest", "_features_rest"), ("opacity", "_opacity"),
           ("scaling", "_scaling"), ("rotation", "_rotation"))
_ROLE = {"xyz": _lib.GSM_DENSIFY_ROLE_XYZ, "scaling": _lib.GSM_DENSIFY_ROLE_SCALING}
_ws = {}
_totals = {}


def _f32(x):
    return float(np.float32(x))


@torch.no_grad()
def densify_and_prune(gaussians, max_grad, min_opacity, extent, max_screen_size, N=2):
    """GaussianModel.densify_and_prune(max_grad, min_opacity, extent, max_screen_size) on `gaussians` (any object with the
    reference model's attributes and a torch.optim.Adam-layout optimizer).  Returns (P_before, P_after)."""
    lib = _lib.load()
    xyz = gaussians._xyz
    if not xyz.is_cuda:
        raise RuntimeError("gaussmart_b200.densify_and_prune: the model must live on a CUDA device (there is no CPU path)")
    dev = xyz.device
    P = xyz.shape[0]
    if P == 0:
        return 0, 0
    params = {name: getattr(gaussians, attr) for name, attr in _GROUPS}
    for name, t in params.items():
        if t.dtype != torch.float32 or not t.is_contiguous() or t.shape[0] != P or t.device != dev:
            raise RuntimeError(f"gaussmart_b200.densify_and_prune: parameter {name} must be contiguous float32 [P, ...] on {dev}")
    accum = gaussians.xyz_gradient_accum.reshape(-1)
    denom = gaussians.denom.reshape(-1)
    if accum.numel() != P or denom.numel() != P or not accum.is_contiguous() or not denom.is_contiguous():
        raise RuntimeError("gaussmart_b200.densify_and_prune: xyz_gradient_accum / denom must be contiguous [P,1]")
    percent_dense = float(gaussians.percent_dense)

    with torch.cuda.device(dev):
        stream = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        nbytes = lib.gsm_densify_workspace_bytes(P)
        ws = _ws.get(dev)
        if ws is None or ws.numel() < nbytes:
            ws = _ws[dev] = torch.empty(int(nbytes * 1.5), dtype=torch.uint8, device=dev)
        totals = _totals.get(dev)
        if totals is None:
            totals = _totals[dev] = torch.zeros(4, dtype=torch.int64).pin_memory()
        a = _lib.DensifyArgs()
        a.P = P
        a.xyz, a.opacity = params["xyz"].data_ptr(), params["opacity"].data_ptr()
        a.scaling, a.rotation = params["scaling"].data_ptr(), params["rotation"].data_ptr()
        a.xyz_gradient_accum, a.denom = accum.data_ptr(), denom.data_ptr()
        # Python scalars meet fp32 tensors in the reference's comparisons: they are rounded to fp32 there as here
        a.grad_threshold = _f32(max_grad)
        a.dense_limit = _f32(percent_dense * extent)
        a.min_opacity = _f32(min_opacity)
        a.world_limit = _f32(0.1 * extent)
        a.split_inv = float(np.float32(1.0) / np.float32(0.8 * N))   # `x / (0.8 * N)`: torch multiplies by the fp32 reciprocal
        a.has_size_limit = 1 if max_screen_size else 0
        a.split_count = int(N)
        a.totals_host = totals.data_ptr()
        _lib.check(lib.gsm_densify_plan(ctypes.byref(a), ws.data_ptr(), ws.numel(), stream), "densify_plan")
        torch.cuda.current_stream(dev).synchronize()   # the one host read: the new tensors are sized from it
        nA, nB, nS, n_split = (int(v) for v in totals.tolist())
        P_new = nA + nB + N * nS

        # the reference's draw (gaussian_model.py:504-507): samples = torch.normal(mean=zeros, std=stds), stds of shape
        # [N * n_split, 3] -> at::normal fills N(0,1) of that shape from the current generator, then scales by std
        normal_draw = torch.empty((N * n_split, 3), dtype=torch.float32, device=dev).normal_()

        f32 = dict(dtype=torch.float32, device=dev)
        new_params = {name: torch.empty((P_new,) + tuple(t.shape[1:]), **f32) for name, t in params.items()}
        opt = gaussians.optimizer
        groups = {g["name"]: g for g in opt.param_groups} if opt is not None else {}
        tensors, new_state, keep = [], {}, []
        for name, _ in _GROUPS:
            t = params[name]
            words = t[0].numel() if P else 1
            tensors.append(_lib.DensifyTensor(t.data_ptr(), new_params[name].data_ptr(), words, _ROLE.get(name, _lib.GSM_DENSIFY_ROLE_COPY)))
            g = groups.get(name)
            st = opt.state.get(g["params"][0], None) if g is not None else None
            if st is not None and "exp_avg" in st:
                m_new, v_new = torch.empty_like(new_params[name]), torch.empty_like(new_params[name])
                m_old, v_old = st["exp_avg"].contiguous(), st["exp_avg_sq"].contiguous()
                keep += [m_old, v_old]
                tensors.append(_lib.DensifyTensor(m_old.data_ptr(), m_new.data_ptr(), words, _lib.GSM_DENSIFY_ROLE_MOMENT))
                tensors.append(_lib.DensifyTensor(v_old.data_ptr(), v_new.data_ptr(), words, _lib.GSM_DENSIFY_ROLE_MOMENT))
                new_state[name] = (m_new, v_new)
        seg = getattr(gaussians, "_segments", None)
        seg_new = None
        if isinstance(seg, torch.Tensor) and seg.numel() == P and seg.is_cuda:
            seg = seg.contiguous()
            seg_new = torch.empty((P_new,), dtype=seg.dtype, device=dev)
            tensors.append(_lib.DensifyTensor(seg.data_ptr(), seg_new.data_ptr(), seg.element_size() // 4, _lib.GSM_DENSIFY_ROLE_COPY))
        src_map = torch.empty((max(P_new, 1),), dtype=torch.int32, device=dev)
        zrow = torch.empty((max(P_new, 1),), dtype=torch.int32, device=dev)
        arr = (_lib.DensifyTensor * len(tensors))(*tensors)
        _lib.check(lib.gsm_densify_apply(ctypes.byref(a), ws.data_ptr(), arr, len(tensors), P_new, src_map.data_ptr(),
                                         zrow.data_ptr(), normal_draw.data_ptr(), stream), "densify_apply")

    # ---- the optimizer surgery of cat_tensors_to_optimizer / _prune_optimizer, once (gaussian_model.py:413-466)
    for name, attr in _GROUPS:
        p_new = nn.Parameter(new_params[name].requires_grad_(True))
        g = groups.get(name)
        if g is not None:
            old = g["params"][0]
            st = opt.state.get(old, None)
            if st is not None:
                if name in new_state:
                    st["exp_avg"], st["exp_avg_sq"] = new_state[name]
                del opt.state[old]
                g["params"][0] = p_new
                opt.state[p_new] = st
            else:
                g["params"][0] = p_new
        setattr(gaussians, attr, p_new)
    if seg_new is not None:
        gaussians._segments = seg_new
    # densification_postfix (gaussian_model.py:487-489): fresh statistics for every surfel
    gaussians.xyz_gradient_accum = torch.zeros((P_new, 1), device=dev)
    gaussians.denom = torch.zeros((P_new, 1), device=dev)
    gaussians.max_radii2D = torch.zeros((P_new,), device=dev)
    return P, P_new
