"""Shared helpers of the parity tests: tolerances (BASELINE.json north_star) and scene plumbing."""
import math

import numpy as np
import torch

FWD_ABS_TOL = 1e-5      # forward outputs: <= 1e-5 abs in fp32 (north_star)
GRAD_REL_TOL = 1e-4     # gradients: <= 1e-4 rel (north_star)
GRAD_ABS_FLOOR = 1e-7   # x tensor scale, so exact zeros do not divide


def grad_violations(a, b, rel=GRAD_REL_TOL, floor=GRAD_ABS_FLOOR):
    """# of elements with |a-b| > rel*max(|a|,|b|) + floor*scale, scale = max |.| over both tensors."""
    a = torch.as_tensor(a).double().reshape(-1)
    b = torch.as_tensor(b).double().reshape(-1)
    scale = max(a.abs().max().item() if a.numel() else 0.0, b.abs().max().item() if b.numel() else 0.0, 1e-30)
    tol = rel * torch.maximum(a.abs(), b.abs()) + floor * scale
    return int(((a - b).abs() > tol).sum().item()), scale


def rel_fro(a, b):
    a = torch.as_tensor(a).double()
    b = torch.as_tensor(b).double()
    return ((a - b).norm() / max(b.norm().item(), 1e-30)).item()


def fwd_close(a, b, tol=FWD_ABS_TOL):
    """Forward tolerance: 1e-5 abs, scaled by max(1,|ref|) so that channels holding view depths of
    tens of units (allmap 0 and 5) are judged at fp32 resolution."""
    a = torch.as_tensor(a).double()
    b = torch.as_tensor(b).double()
    return (a - b).abs() <= tol * torch.clamp(b.abs(), min=1.0)


def precomp_T(scene, cam, mod=1.0):
    """[P,9] rows Tu,Tv,Tw in float64 on the CPU (what gaussian_renderer/__init__.py:64-75 would pass)."""
    q = scene.rotations.double()
    w, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    R = torch.stack([
        torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)], -1),
        torch.stack([2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)], -1),
        torch.stack([2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)], -1)], 1)
    s = scene.scales.double() * mod
    P = scene.means3D.shape[0]
    Mt = torch.zeros(P, 3, 4, dtype=torch.float64)
    Mt[:, 0, :3] = R[:, :, 0] * s[:, 0:1]
    Mt[:, 1, :3] = R[:, :, 1] * s[:, 1:2]
    Mt[:, 2, :3] = scene.means3D.double()
    Mt[:, 2, 3] = 1
    W, H = cam.image_width, cam.image_height
    N = torch.tensor([[W / 2, 0, 0], [0, H / 2, 0], [0, 0, 0], [(W - 1) / 2, (H - 1) / 2, 1]], dtype=torch.float64)
    T = Mt @ cam.full_proj_transform.double() @ N
    return torch.cat([T[:, :, 0], T[:, :, 1], T[:, :, 2]], dim=1).float().contiguous()


def variant_inputs(scene, cam, variant, scale_modifier=1.0):
    """kwargs (CPU tensors) of GaussianRasterizer.forward for one of the four input variants."""
    kw = {}
    if variant in ("sh", "sh_cov"):
        kw["shs"] = scene.shs
    else:
        kw["colors_precomp"] = torch.sigmoid(scene.shs[:, 0, :]).contiguous()
    if variant in ("sh", "colors"):
        kw["scales"] = scene.scales
        kw["rotations"] = scene.rotations
    else:
        kw["cov3D_precomp"] = precomp_T(scene, cam, scale_modifier)
    return kw


def run_cuda(mod, scene, cam, bg_value, cot, variant="sh", scale_modifier=1.0, backward=True, device="cuda",
             debug=False):
    """Runs `mod` (gaussmart_b200 or the reference build: same API) on the GPU; returns detached outputs + grads."""
    from gaussmart_b200 import synthetic
    cam_d = synthetic.Camera(*[x.to(device) if isinstance(x, torch.Tensor) else x for x in cam])
    bg = torch.full((3,), float(bg_value), device=device)
    S = mod.GaussianRasterizationSettings(**synthetic.settings_kwargs(cam_d, scene.sh_degree, bg, scale_modifier, debug))
    rast = mod.GaussianRasterizer(S)
    means3D = scene.means3D.to(device).requires_grad_(backward)
    means2D = torch.zeros_like(means3D, requires_grad=backward)
    opac = scene.opacities.to(device).requires_grad_(backward)
    leaves = dict(means3D=means3D, means2D=means2D, opacities=opac)
    kw = {}
    for k, v in variant_inputs(scene, cam, variant, scale_modifier).items():
        kw[k] = v.to(device).requires_grad_(backward)
        leaves[k] = kw[k]
    color, radii, allmap = rast(means3D=means3D, means2D=means2D, opacities=opac, **kw)
    out = dict(color=color.detach(), radii=radii.detach(), allmap=allmap.detach(), settings=S, leaves=leaves)
    if backward:
        torch.autograd.backward([color, allmap], [cot[0].to(device), cot[1].to(device)])
        for k, v in leaves.items():
            out["grad_" + k] = None if v.grad is None else v.grad.detach()
    return out


def run_oracle(oracle, scene, cam, bg_value, cot=None, variant="sh", scale_modifier=1.0):
    kw = variant_inputs(scene, cam, variant, scale_modifier)
    f = oracle.rasterize_forward(
        scene.means3D, scene.opacities, cam.world_view_transform, cam.full_proj_transform, cam.camera_center,
        cam.image_width, cam.image_height, cam.tanfovx, cam.tanfovy, torch.full((3,), float(bg_value)),
        sh_degree=scene.sh_degree, shs=kw.get("shs"), colors_precomp=kw.get("colors_precomp"),
        scales=kw.get("scales"), rotations=kw.get("rotations"), transMat_precomp=kw.get("cov3D_precomp"),
        scale_modifier=scale_modifier)
    g = None
    if cot is not None and f.get("pre") is not None:
        g = oracle.rasterize_backward(f, cot[0], cot[1])
    return f, g


# names: leaf name in the public API -> key in the oracle's backward dict
ORACLE_GRAD_KEYS = dict(means3D="dL_dmeans3D", means2D="dL_dmeans2D", opacities="dL_dopacity", shs="dL_dsh",
                        colors_precomp="dL_dcolors", scales="dL_dscales", rotations="dL_drotations",
                        cov3D_precomp="dL_dtransMat")


def robust_rel_fro(a, b, drop_frac=0.002):
    """rel. Frobenius error after dropping the `drop_frac` rows (surfels) with the largest deviation: the
    CPU oracle and the GPU can take a different discrete decision (alpha >= 1/255, T < 1e-4, median) for
    a pixel of an ill-conditioned surfel, which moves that surfel's gradient by ~1e-3 rel."""
    a = torch.as_tensor(a).double()
    b = torch.as_tensor(b).double()
    a = a.reshape(a.shape[0], -1)
    b = b.reshape(b.shape[0], -1)
    d = (a - b).abs().max(1).values
    k = int(drop_frac * a.shape[0]) + 1
    keep = torch.ones(a.shape[0], dtype=torch.bool)
    keep[torch.topk(d, min(k, a.shape[0])).indices] = False
    return rel_fro(a[keep], b[keep])


def assert_grads_match_reference(new, ref, ref_again=None, what=""):
    """Gradient bar against the reference BUILD: |a-b| <= 1e-4 max(|a|,|b|) + 1e-7 scale element-wise.
    Both builds accumulate with unordered fp32 atomics, so the reference does not even reproduce itself
    element-wise, and a sum of cancelling terms taken in another order moves by more
    than 1e-4 of its (small) value: the budget is twice what two runs of the reference differ by + 2
    + 1e-5 of the entries, and the tensor as a whole must agree to 1e-5 rel."""
    for k in ref:
        if not k.startswith("grad_") or ref[k] is None:
            continue
        viol, _ = grad_violations(new[k].cpu(), ref[k].cpu())
        noise = 0
        if ref_again is not None:
            noise, _ = grad_violations(ref_again[k].cpu(), ref[k].cpu())
        assert viol <= 2 * noise + 2 + int(1e-5 * ref[k].numel()), (what, k, viol, noise)
        assert rel_fro(new[k].cpu(), ref[k].cpu()) < 1e-5, (what, k)


def compare_with_reference_build(scene, cam, variant="sh", bg=0.0, what="", scale_modifier=1.0, backward=True,
                                 images_bit_exact=False):
    """The whole bar against the reference BUILD (oracle/_ref) on one scene, zero budget on everything integer:
    radii, instance count, sorted (tile|depth) keys, depth-sorted surfel list, tile ranges and last-contributor counts
    bit-exact (read out of the reference's own scratch buffers); images <= 1e-5; gradients within
    assert_grads_match_reference.  Returns (ours, reference, our intermediates)."""
    import gaussmart_b200 as gb
    from gaussmart_b200 import synthetic
    import ref_loader
    import workspace_views
    ref = ref_loader.load_dsr()
    W, H = cam.image_width, cam.image_height
    cot = synthetic.make_cotangents(W, H, 777) if backward else None
    o_ref = run_cuda(ref, scene, cam, bg, cot, variant, scale_modifier, backward=backward)
    o_new = run_cuda(gb, scene, cam, bg, cot, variant, scale_modifier, backward=backward)
    assert torch.equal(o_ref["radii"], o_new["radii"]), what
    for k in ("color", "allmap"):
        if images_bit_exact:
            assert torch.equal(o_new[k], o_ref[k]), (what, k)
        assert fwd_close(o_new[k].cpu(), o_ref[k].cpu()).all(), (what, k)
    if backward:
        o_ref2 = run_cuda(ref, scene, cam, bg, cot, variant, scale_modifier)
        assert_grads_match_reference(o_new, o_ref, o_ref2, what)
        del o_ref2
    S = o_ref["settings"]
    e = torch.empty(0, device="cuda")
    kw = {k: v.cuda() for k, v in variant_inputs(scene, cam, variant, scale_modifier).items()}
    rr = ref._C.rasterize_gaussians(
        S.bg, scene.means3D.cuda(), kw.get("colors_precomp", e), scene.opacities.cuda(), kw.get("scales", e),
        kw.get("rotations", e), S.scale_modifier, kw.get("cov3D_precomp", e), S.viewmatrix, S.projmatrix, S.tanfovx,
        S.tanfovy, S.image_height, S.image_width, kw.get("shs", e), S.sh_degree, S.campos, False, False)
    R_ref, binB, imgB = rr[0], rr[5], rr[6]
    im = workspace_views.forward_with_intermediates(o_new["settings"], scene.means3D.cuda(), scene.opacities.cuda(), **kw)
    assert im["num_rendered"] == R_ref, what
    pl, keys = ref_loader.ref_binning_views(binB, R_ref)
    assert torch.equal(keys, im["keys"]), what
    assert torch.equal(pl, im["point_list"]), what
    acc, nc, rg = ref_loader.ref_image_views(imgB, W, H)
    assert torch.equal(rg, im["ranges"]), what
    assert torch.equal(nc[0], im["n_contrib"][0]), what
    # the median contributor of a pixel nothing contributed to is (uint32_t)(-1.0f) in the reference (forward.cu:437):
    # undefined, garbage in its sm_100a build; it is only read where last_contributor > 0
    has = nc[0] > 0
    assert torch.equal(nc[1][has], im["n_contrib"][1][has]), what
    assert fwd_close(im["final_T"].cpu(), acc.cpu()).all(), what
    return o_new, o_ref, im
