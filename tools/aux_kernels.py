#!/usr/bin/env python
"""Runs the auxiliary kernels of the path once each at C4 size, for an ncu capture (profiles/r02_aux_ncu.txt):
distCUDA2 (min/max, Morton, 32-bit radix sort, tree build, search), the SH expand of the view-parallel exchange,
densify_and_prune (plan / scan / map / gather), the tile scan at 3840x2160."""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import gaussmart_b200 as gb
from gaussmart_b200 import synthetic
from gaussmart_b200.rasterizer import sh_grad_expand

P = 5_000_000
scene = synthetic.make_scene(P, 1237, 3, synthetic.BASE_SEED + 4)
pts = scene.means3D.cuda()
for _ in range(2):
    d2 = gb.distCUDA2(pts)
g = torch.Generator(device="cuda").manual_seed(1)
V = 8
campos = torch.randn(V, 3, device="cuda", generator=g) * 5
rgb = torch.randn(V, P, 3, device="cuda", generator=g)
rgb[torch.rand(V, P, device="cuda", generator=g) < 0.45] = 0
d = torch.empty(P, 16, 3, device="cuda")
for _ in range(2):
    sh_grad_expand(pts, campos, rgb, d, 3)
# one 4K forward (tile scan over 32,400 tiles)
cam = synthetic.make_camera(3840, 2160, 0.3)
cam_d = synthetic.Camera(*[x.cuda() if isinstance(x, torch.Tensor) else x for x in cam])
S = gb.GaussianRasterizationSettings(**synthetic.settings_kwargs(cam_d, 3, torch.zeros(3).cuda()))
with torch.no_grad():
    for _ in range(2):
        gb.GaussianRasterizer(S)(means3D=pts, means2D=torch.zeros_like(pts), opacities=scene.opacities.cuda(), shs=scene.shs.cuda(),
                                 scales=scene.scales.cuda(), rotations=scene.rotations.cuda())
# densify_and_prune on a model of 2 M surfels
try:
    import ref_callers
    from argparse import ArgumentParser
    from gaussmart_b200.densify import densify_and_prune
    C = ref_callers.load("ours")
    Pm = 2_000_000
    r = lambda *s: torch.randn(*s, device="cuda", generator=g)
    gm = C.GaussianModel(3)
    gm._xyz, gm._features_dc = torch.nn.Parameter(r(Pm, 3) * 2.0), torch.nn.Parameter(r(Pm, 1, 3) * 0.5)
    gm._features_rest, gm._opacity = torch.nn.Parameter(r(Pm, 15, 3) * 0.05), torch.nn.Parameter(r(Pm, 1) * 2.5 - 1.0)
    gm._scaling = torch.nn.Parameter(r(Pm, 2) * 0.9 - 3.1)
    gm._rotation = torch.nn.Parameter(r(Pm, 4))
    gm._segments = torch.zeros(Pm, dtype=torch.long, device="cuda")
    gm.max_radii2D = torch.zeros(Pm, device="cuda")
    gm.spatial_lr_scale = 1.0
    gm.training_setup(C.arguments.OptimizationParams(ArgumentParser()))
    gm.denom = torch.randint(0, 40, (Pm, 1), device="cuda", generator=g).float()
    gm.xyz_gradient_accum = gm.denom * torch.rand(Pm, 1, device="cuda", generator=g) * 4e-4
    print("densify", densify_and_prune(gm, 0.0002, 0.05, 4.4, 20))
except Exception as e:
    print("densify skipped:", e)
torch.cuda.synchronize()
print("ok", float(d2.sum()), float(d.abs().sum()))
