#!/usr/bin/env python
"""GPU-box tool: run the reference CUDA build (oracle/_ref) and gaussmart_b200 on the same
synthetic scene, compare every output and intermediate, time both, optionally dump goldens.

    python tools/compare_ref.py --config C1 [--P 10000] [--dump gpurun_out/golden] [--bg 0|1]
"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import gaussmart_b200 as gb  # noqa: E402
from gaussmart_b200 import synthetic  # noqa: E402
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import workspace_views as debug  # noqa: E402
import ref_loader  # noqa: E402


def relerr(a, b):
    a = a.double(); b = b.double()
    scale = max(a.abs().max().item(), b.abs().max().item(), 1e-30)
    tol = 1e-4 * torch.maximum(a.abs(), b.abs()) + 1e-7 * scale
    viol = ((a - b).abs() > tol).sum().item()
    return dict(max_abs=(a - b).abs().max().item(), scale=scale, viol=viol, n=a.numel(),
                rel_fro=((a - b).norm() / max(b.norm().item(), 1e-30)).item())


OURS_DEBUG = 0


def run_impl(mod, scene, cam, bg, cot, variant, scale_modifier=1.0, backward=True):
    dev = "cuda"
    S = mod.GaussianRasterizationSettings(**synthetic.settings_kwargs(
        synthetic.Camera(*[x.to(dev) if isinstance(x, torch.Tensor) else x for x in cam]), scene.sh_degree, bg,
        scale_modifier, OURS_DEBUG if mod is gb else False))
    rast = mod.GaussianRasterizer(S)
    means3D = scene.means3D.to(dev).requires_grad_(True)
    means2D = torch.zeros_like(means3D, requires_grad=True)
    opac = scene.opacities.to(dev).requires_grad_(True)
    kw = {}
    leaves = dict(means3D=means3D, means2D=means2D, opacities=opac)
    if variant in ("sh", "sh_cov"):
        kw["shs"] = scene.shs.to(dev).requires_grad_(True); leaves["shs"] = kw["shs"]
    else:
        kw["colors_precomp"] = torch.sigmoid(scene.shs[:, 0, :]).contiguous().to(dev).requires_grad_(True)
        leaves["colors_precomp"] = kw["colors_precomp"]
    if variant in ("sh", "colors"):
        kw["scales"] = scene.scales.to(dev).requires_grad_(True); leaves["scales"] = kw["scales"]
        kw["rotations"] = scene.rotations.to(dev).requires_grad_(True); leaves["rotations"] = kw["rotations"]
    else:
        kw["cov3D_precomp"] = precomp_T(scene, cam, scale_modifier).to(dev).requires_grad_(True)
        leaves["cov3D_precomp"] = kw["cov3D_precomp"]
    color, radii, allmap = rast(means3D=means3D, means2D=means2D, opacities=opac, **kw)
    out = dict(color=color.detach(), radii=radii.detach(), allmap=allmap.detach())
    if backward:
        loss = (color * cot[0].to(dev)).sum() + (allmap * cot[1].to(dev)).sum()
        loss.backward()
        for k, v in leaves.items():
            out["grad_" + k] = v.grad.detach() if v.grad is not None else None
    return out, S, leaves


def precomp_T(scene, cam, mod=1.0):
    """[P,9] rows Tu,Tv,Tw computed in float64 on CPU (gaussian_renderer/__init__.py:64-75 analogue)."""
    q = scene.rotations.double()
    w, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    R = torch.stack([
        torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)], -1),
        torch.stack([2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)], -1),
        torch.stack([2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)], -1)], 1)  # [P,3,3] rows
    s = scene.scales.double() * mod
    L0 = R[:, :, 0] * s[:, 0:1]
    L1 = R[:, :, 1] * s[:, 1:2]
    P = scene.means3D.shape[0]
    Mt = torch.zeros(P, 3, 4, dtype=torch.float64)
    Mt[:, 0, :3] = L0; Mt[:, 1, :3] = L1; Mt[:, 2, :3] = scene.means3D.double(); Mt[:, 2, 3] = 1
    W, H = cam.image_width, cam.image_height
    N = torch.tensor([[W / 2, 0, 0], [0, H / 2, 0], [0, 0, 0], [(W - 1) / 2, (H - 1) / 2, 1]], dtype=torch.float64)
    T = Mt @ cam.full_proj_transform.double() @ N  # [P,3,3]; T[:, :, j] = column j
    return torch.cat([T[:, :, 0], T[:, :, 1], T[:, :, 2]], dim=1).float().contiguous()


def time_it(fn, warm=3, iters=10):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(iters + 1)]
    ev[0].record()
    for i in range(iters):
        fn()
        ev[i + 1].record()
    torch.cuda.synchronize()
    ts = sorted(ev[i].elapsed_time(ev[i + 1]) for i in range(iters))
    return ts[len(ts) // 2]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="C1")
    ap.add_argument("--P", type=int, default=None)
    ap.add_argument("--variant", default="sh", choices=["sh", "colors", "sh_cov", "colors_cov"])
    ap.add_argument("--bg", type=float, default=0.0)
    ap.add_argument("--scale_modifier", type=float, default=1.0)
    ap.add_argument("--dump", default=None)
    ap.add_argument("--time", action="store_true")
    ap.add_argument("--view", type=int, default=0)
    ap.add_argument("--out", default=None)
    ap.add_argument("--ours-debug", type=int, default=0)
    a = ap.parse_args()
    global OURS_DEBUG
    OURS_DEBUG = a.ours_debug

    ref = ref_loader.load_dsr()
    scene, cams = synthetic.make_config(a.config, P=a.P)
    cam = cams[a.view]
    W, H = cam.image_width, cam.image_height
    bg = torch.full((3,), a.bg, device="cuda")
    cot = synthetic.make_cotangents(W, H, 777)
    res = dict(config=a.config, P=scene.means3D.shape[0], W=W, H=H, variant=a.variant, bg=a.bg)

    o_ref, S_ref, _ = run_impl(ref, scene, cam, bg, cot, a.variant, a.scale_modifier)
    o_new, S_new, _ = run_impl(gb, scene, cam, bg, cot, a.variant, a.scale_modifier)
    torch.cuda.synchronize()

    res["radii_mismatch"] = int((o_ref["radii"] != o_new["radii"]).sum().item())
    if res["radii_mismatch"]:
        bad = torch.nonzero(o_ref["radii"] != o_new["radii"]).reshape(-1)[:8]
        res["radii_bad"] = [(int(i), int(o_ref["radii"][i]), int(o_new["radii"][i])) for i in bad]
    res["P_vis"] = int((o_ref["radii"] > 0).sum().item())
    res["color_max_abs"] = (o_ref["color"] - o_new["color"]).abs().max().item()
    res["allmap_max_abs"] = [(o_ref["allmap"][c] - o_new["allmap"][c]).abs().max().item() for c in range(7)]
    for k in o_ref:
        if k.startswith("grad_") and o_ref[k] is not None:
            res[k] = relerr(o_new[k], o_ref[k])

    # intermediates: sorted keys, point list, ranges, n_contrib
    dev = "cuda"
    e = torch.empty(0, device=dev)
    kw = dict(shs=scene.shs.to(dev)) if a.variant.startswith("sh") else dict(
        colors_precomp=torch.sigmoid(scene.shs[:, 0, :]).contiguous().to(dev))
    if a.variant in ("sh", "colors"):
        kw.update(scales=scene.scales.to(dev), rotations=scene.rotations.to(dev))
    else:
        kw.update(cov3D_precomp=precomp_T(scene, cam, a.scale_modifier).to(dev))
    mine = debug.forward_with_intermediates(S_new, scene.means3D.to(dev), scene.opacities.to(dev), **kw)
    rr = ref._C.rasterize_gaussians(
        S_ref.bg, scene.means3D.to(dev), kw.get("colors_precomp", e), scene.opacities.to(dev), kw.get("scales", e),
        kw.get("rotations", e), S_ref.scale_modifier, kw.get("cov3D_precomp", e), S_ref.viewmatrix, S_ref.projmatrix,
        S_ref.tanfovx, S_ref.tanfovy, S_ref.image_height, S_ref.image_width, kw.get("shs", e), S_ref.sh_degree,
        S_ref.campos, False, False)
    R_ref, _, _, _, geomB, binB, imgB = rr
    res["R_ref"], res["R_new"] = int(R_ref), int(mine["num_rendered"])
    pl = keys = None
    if R_ref > 0:
        pl, keys = ref_loader.ref_binning_views(binB, R_ref)
    if R_ref == mine["num_rendered"] and R_ref > 0:
        res["keys_mismatch"] = int((keys != mine["keys"]).sum().item())
        res["point_list_mismatch"] = int((pl != mine["point_list"]).sum().item())
    acc, nc, rg = ref_loader.ref_image_views(imgB, W, H)
    res["ranges_mismatch"] = int((rg != mine["ranges"]).sum().item())
    res["n_contrib_mismatch"] = [int((nc[i] != mine["n_contrib"][i]).sum().item()) for i in range(2)]
    res["final_T_max_abs"] = (acc - mine["final_T"]).abs().max().item()
    for ch in range(2):
        bad = torch.nonzero(nc[ch] != mine["n_contrib"][ch]).reshape(-1)[:6]
        if bad.numel():
            res[f"n_contrib_bad{ch}"] = [(int(i) % W, int(i) // W, int(nc[ch][i]), int(mine["n_contrib"][ch][i]),
                                          float(acc[0][i]), float(mine["final_T"][0][i])) for i in bad]
    res["E_list_entries"] = int(nc[0].long().sum().item())

    if a.time:
        def mk(mod):
            out, S, leaves = run_impl(mod, scene, cam, bg, cot, a.variant, a.scale_modifier, backward=False)
            rast = mod.GaussianRasterizer(S)
            c0, c1 = cot[0].cuda(), cot[1].cuda()
            names = [k for k in leaves if k not in ("means3D", "means2D", "opacities")]

            def fwd():
                with torch.no_grad():
                    rast(means3D=leaves["means3D"], means2D=leaves["means2D"], opacities=leaves["opacities"],
                         **{k: leaves[k] for k in names})

            def fwdbwd():
                for v in leaves.values():
                    v.grad = None
                c, r, m = rast(means3D=leaves["means3D"], means2D=leaves["means2D"], opacities=leaves["opacities"],
                               **{k: leaves[k] for k in names})
                ((c * c0).sum() + (m * c1).sum()).backward()
            return fwd, fwdbwd
        f_ref, fb_ref = mk(ref)
        f_new, fb_new = mk(gb)
        res["ms_fwd_ref"] = time_it(f_ref); res["ms_fwd_new"] = time_it(f_new)
        res["ms_fwdbwd_ref"] = time_it(fb_ref); res["ms_fwdbwd_new"] = time_it(fb_new)

    if a.dump:
        os.makedirs(a.dump, exist_ok=True)
        tag = f"{a.config}_P{res['P']}_{a.variant}_bg{int(a.bg)}"
        np.savez_compressed(
            os.path.join(a.dump, f"ref_{tag}.npz"),
            color=o_ref["color"].cpu().numpy(), allmap=o_ref["allmap"].cpu().numpy(),
            radii=o_ref["radii"].cpu().numpy(), keys=keys.cpu().numpy() if R_ref > 0 else np.zeros(0, np.int64),
            point_list=pl.cpu().numpy() if R_ref > 0 else np.zeros(0, np.int32), ranges=rg.cpu().numpy(),
            n_contrib=nc.cpu().numpy(), final_T=acc.cpu().numpy(),
            **{k: v.cpu().numpy() for k, v in o_ref.items() if k.startswith("grad_") and v is not None})
    s = json.dumps(res)
    print(s)
    if a.out:
        with open(a.out, "a") as f:
            f.write(s + "\n")


if __name__ == "__main__":
    main()
