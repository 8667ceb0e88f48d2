"""TEST INFRASTRUCTURE ONLY -- torch fp32 restatement of what GauSSmart does with the rasterizer's allmap on the training
path, the checker of gsm_surface_regularizer (csrc/regularizer.cu).  Only tests/, __graft_entry__.smoke() and bench.py's
reference arm may import this; the product (gaussmart_b200/) never does.

Follows gaussian_renderer/__init__.py:117-147 (post-processing inside render()), utils/point_utils.py:9-37
(depths_to_points, depth_to_normal) and train.py:131-140 (normal_loss, dist_loss).  Device-agnostic (the reference
hard-codes .cuda()).  Pinned: depth_to_normal against the reference function itself on seeded inputs
(tests/golden/regularizer_ref_*.npz, made by tests/golden/make_regularizer_golden.py); the render() glue lines cannot be
called in isolation (they sit inside render()), they are restated here line by line.
"""
import torch


def pixel_rays(view, device):
    # point_utils.py:9-22
    V, FP = view.world_view_transform.to(device), view.full_proj_transform.to(device)
    W, H = view.image_width, view.image_height
    cam2world = V.T.inverse()
    ndc_to_pixel = torch.tensor([[W / 2, 0, 0, W / 2], [0, H / 2, 0, H / 2], [0, 0, 0, 1]], device=device).to(V.dtype).T
    intrinsics = ((cam2world.T @ FP) @ ndc_to_pixel)[:3, :3].T
    gx, gy = torch.meshgrid(torch.arange(W, device=device).to(V.dtype), torch.arange(H, device=device).to(V.dtype), indexing="xy")
    pix = torch.stack([gx, gy, torch.ones_like(gx)], dim=-1).reshape(-1, 3)
    return pix @ intrinsics.inverse().T @ cam2world[:3, :3].T, cam2world[:3, 3]


def depth_to_normal(view, depth):
    # point_utils.py:23-37: back-project, central differences over rows / columns, cross, normalize; zero border
    dirs, origin = pixel_rays(view, depth.device)
    pts = (depth.reshape(-1, 1) * dirs + origin).reshape(*depth.shape[1:], 3)
    normal = torch.zeros_like(pts)
    d_rows = pts[2:, 1:-1] - pts[:-2, 1:-1]
    d_cols = pts[1:-1, 2:] - pts[1:-1, :-2]
    normal[1:-1, 1:-1, :] = torch.nn.functional.normalize(torch.cross(d_rows, d_cols, dim=-1), dim=-1)
    return normal


def surface_maps(allmap, view, depth_ratio, normal_fn=depth_to_normal):
    # gaussian_renderer/__init__.py:117-147
    alpha = allmap[1:2]
    rend_normal = (allmap[2:5].permute(1, 2, 0) @ (view.world_view_transform.to(allmap.device)[:3, :3].T)).permute(2, 0, 1)
    median = torch.nan_to_num(allmap[5:6], 0, 0)
    expected = torch.nan_to_num(allmap[0:1] / alpha, 0, 0)
    surf_depth = expected * (1 - depth_ratio) + depth_ratio * median
    surf_normal = normal_fn(view, surf_depth).permute(2, 0, 1) * alpha.detach()
    return dict(rend_alpha=alpha, rend_normal=rend_normal, rend_dist=allmap[6:7], surf_depth=surf_depth, surf_normal=surf_normal)


def regularizer_terms(allmap, view, depth_ratio, lambda_normal, lambda_dist, normal_fn=depth_to_normal):
    # train.py:131-140
    m = surface_maps(allmap, view, depth_ratio, normal_fn)
    normal_error = (1 - (m["rend_normal"] * m["surf_normal"]).sum(dim=0))[None]
    return lambda_normal * normal_error.mean(), lambda_dist * m["rend_dist"].mean(), m


def surface_regularizer(allmap, view, depth_ratio=0.0, lambda_normal=0.05, lambda_dist=0.0):
    n, d, _ = regularizer_terms(allmap, view, depth_ratio, lambda_normal, lambda_dist)
    return n + d
