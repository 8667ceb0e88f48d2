"""Seeded synthetic scenes and cameras for parity tests and bench.py.

Everything is generated in fp32 on the CPU with a seeded torch.Generator so that every
implementation (this repo, the reference CUDA build, the CPU oracle) sees identical bits.
Camera matrices are built exactly as scene/cameras.py:56-59 with
utils/graphics_utils.py:38-67 (getWorld2View2, getProjectionMatrix znear=0.01 zfar=100).
"""
import math
from typing import NamedTuple

import numpy as np
import torch

CONFIGS = {
    # name: (P, W, H, sh_degree, n_views)
    "C1": (10_000, 128, 128, 0, 1),
    "C2": (300_000, 800, 600, 3, 49),
    "C3": (2_000_000, 1957, 1091, 3, 1),
    "C4": (5_000_000, 1237, 822, 3, 1),
    "C5": (6_000_000, 1237, 822, 3, 8),
    "C5_4K": (6_000_000, 3840, 2160, 3, 8),
}
BASE_SEED = 20261104
SCENE_CENTRE = (0.0, 0.0, 4.0)


class Camera(NamedTuple):
    image_width: int
    image_height: int
    FoVx: float
    FoVy: float
    tanfovx: float
    tanfovy: float
    world_view_transform: torch.Tensor  # [4,4] as stored by the reference (transposed, row-vector use)
    full_proj_transform: torch.Tensor   # [4,4]
    camera_center: torch.Tensor         # [3]


class Scene(NamedTuple):
    means3D: torch.Tensor    # [P,3]
    scales: torch.Tensor     # [P,2] post-activation
    rotations: torch.Tensor  # [P,4] unit (w,x,y,z)
    opacities: torch.Tensor  # [P,1] post-activation
    shs: torch.Tensor        # [P,M,3]
    sh_degree: int


def _world2view2(R, t):
    Rt = np.zeros((4, 4))
    Rt[:3, :3] = R.transpose()
    Rt[:3, 3] = t
    Rt[3, 3] = 1.0
    C2W = np.linalg.inv(Rt)
    Rt = np.linalg.inv(C2W)
    return np.float32(Rt)


def _projection(znear, zfar, fovX, fovY):
    tanHalfFovY = math.tan(fovY / 2)
    tanHalfFovX = math.tan(fovX / 2)
    top = tanHalfFovY * znear
    bottom = -top
    right = tanHalfFovX * znear
    left = -right
    P = torch.zeros(4, 4)
    P[0, 0] = 2.0 * znear / (right - left)
    P[1, 1] = 2.0 * znear / (top - bottom)
    P[0, 2] = (right + left) / (right - left)
    P[1, 2] = (top + bottom) / (top - bottom)
    P[3, 2] = 1.0
    P[2, 2] = zfar / (zfar - znear)
    P[2, 3] = -(zfar * znear) / (zfar - znear)
    return P


def camera_pose(theta=0.0, radius=4.0):
    """(R, T) as scene/cameras.py:Camera takes them (R: camera-to-world rotation, T: world-to-camera translation, the
    COLMAP convention of the reference) of a camera on a circle of `radius` around the scene centre looking inward."""
    c = np.array(SCENE_CENTRE)
    C = c + radius * np.array([math.sin(theta), 0.0, -math.cos(theta)])
    z = (c - C) / np.linalg.norm(c - C)
    x = np.cross(np.array([0.0, 1.0, 0.0]), z)
    x /= np.linalg.norm(x)
    y = np.cross(z, x)
    R = np.stack([x, y, z], axis=1)
    t = -R.T @ C
    return R, t


def camera_fov(W, H, focal=None, fovx=None):
    if fovx is None:
        focal = 0.9 * W if focal is None else focal
        fovx = 2.0 * math.atan(W / (2.0 * focal))
    else:
        focal = W / (2.0 * math.tan(fovx / 2.0))
    return fovx, 2.0 * math.atan(H / (2.0 * focal))


def make_camera(W, H, theta=0.0, focal=None, radius=4.0, fovx=None):
    """Camera on a circle of `radius` around the scene centre looking inward; theta=0 is R=I, T=0."""
    fovx, fovy = camera_fov(W, H, focal, fovx)
    R, t = camera_pose(theta, radius)
    wvt = torch.tensor(_world2view2(R, t)).transpose(0, 1).contiguous()
    proj = _projection(0.01, 100.0, fovx, fovy).transpose(0, 1)
    full = (wvt.unsqueeze(0).bmm(proj.unsqueeze(0))).squeeze(0).contiguous()
    center = wvt.inverse()[3, :3].contiguous()
    return Camera(W, H, fovx, fovy, math.tan(fovx * 0.5), math.tan(fovy * 0.5), wvt, full, center)


def make_cameras(W, H, n_views):
    if n_views == 1:
        return [make_camera(W, H, 0.0)]
    return [make_camera(W, H, 2.0 * math.pi * v / n_views) for v in range(n_views)]


def make_scene(P, W, sh_degree, seed, max_sh_degree=None):
    """SURVEY.md 8(d) surfel distribution: 70% noisy sphere shell, 20% cube, 10% behind the near plane."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    max_sh_degree = sh_degree if max_sh_degree is None else max_sh_degree
    M = (max_sh_degree + 1) ** 2
    c = torch.tensor(SCENE_CENTRE)
    n_surf = int(0.7 * P)
    n_cube = int(0.2 * P)
    n_near = P - n_surf - n_cube
    nrm = torch.randn(n_surf, 3, generator=g)
    nrm = nrm / nrm.norm(dim=1, keepdim=True).clamp_min(1e-12)
    eps = torch.randn(n_surf, 1, generator=g)
    surf = c + 1.5 * nrm * (1.0 + 0.03 * eps)
    cube = c + (torch.rand(n_cube, 3, generator=g) * 2.0 - 1.0) * 4.0
    near = torch.rand(n_near, 3, generator=g)
    near = torch.stack([near[:, 0] * 8.0 - 4.0, near[:, 1] * 8.0 - 4.0, near[:, 2] * 4.2 - 4.0], dim=1)
    means = torch.cat([surf, cube, near], dim=0)
    perm = torch.randperm(P, generator=g)
    means = means[perm].contiguous().float()
    s0 = 1.2 * 4.0 / (0.9 * W)
    scales = torch.exp(math.log(s0) + 0.6 * torch.randn(P, 2, generator=g)).float()
    rot = torch.randn(P, 4, generator=g)
    rot = (rot / rot.norm(dim=1, keepdim=True).clamp_min(1e-12)).float()
    opac = torch.sigmoid(1.5 * torch.randn(P, 1, generator=g)).float()
    shs = torch.empty(P, M, 3)
    shs[:, 0, :] = 0.5 * torch.randn(P, 3, generator=g)
    if M > 1:
        shs[:, 1:, :] = 0.05 * torch.randn(P, M - 1, 3, generator=g)
    return Scene(means, scales, rot, opac, shs.float().contiguous(), sh_degree)


def make_config(name, P=None, seed_offset=0):
    Pn, W, H, deg, n_views = CONFIGS[name]
    P = Pn if P is None else P
    idx = list(CONFIGS).index(name) + 1
    scene = make_scene(P, W, deg, BASE_SEED + idx + seed_offset)
    return scene, make_cameras(W, H, n_views)


def make_cotangents(W, H, seed):
    g = torch.Generator(device="cpu").manual_seed(seed)
    n = W * H
    return (torch.randn(3, H, W, generator=g) / n).float(), (torch.randn(7, H, W, generator=g) / n).float()


def settings_kwargs(cam, sh_degree, bg, scale_modifier=1.0, debug=False):
    """Arguments of GaussianRasterizationSettings exactly as gaussian_renderer/__init__.py:37-51 passes them."""
    return dict(image_height=int(cam.image_height), image_width=int(cam.image_width), tanfovx=cam.tanfovx,
                tanfovy=cam.tanfovy, bg=bg, scale_modifier=scale_modifier, viewmatrix=cam.world_view_transform,
                projmatrix=cam.full_proj_transform, sh_degree=sh_degree, campos=cam.camera_center,
                prefiltered=False, debug=debug)
