"""CPU oracle: a restatement of the reference's 2DGS surfel rasterizer math in PyTorch on the CPU.

TEST INFRASTRUCTURE ONLY.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs may import this module; nothing under gaussmart_b200/ does.

Pinning: the reference ships no tests, golden vectors or fixtures for this path, and
it has no CPU implementation, so the oracle cannot be checked against the reference inside the
build container.  It is pinned instead against outputs of the UNMODIFIED reference CUDA build
(oracle/_ref, built by oracle/build_ref.py) captured on a B200 by tools/compare_ref.py and committed
under tests/golden/ (tests/test_oracle_golden.py).

Each function cites the reference lines it follows (paths relative to
submodules/diff-surfel-rasterization/cuda_rasterizer/).  Arithmetic is fp32; where the reference's
sm_100a build contracts a*b+c into an FMA (read from its PTX) the oracle uses `fma`, emulated as
one fp64 multiply-add rounded to fp32, so integer outputs (radii, tile rects, sort keys) agree
with the GPU up to rare double-rounding cases.
"""
import math

import numpy as np
import torch

F32 = torch.float32
F64 = torch.float64

NEAR_N = 0.2
FAR_N = 100.0
MIN_RADIUS = float(np.frombuffer(np.uint32(0x4007C3AC).tobytes(), dtype=np.float32)[0])  # fl(3 * 0.707106f)
FAR_RATIO = float(np.frombuffer(np.uint32(0x3F8041AB).tobytes(), dtype=np.float32)[0])   # fl(100/(100-0.2))
ALPHA_MAX = float(np.float32(0.99))
ALPHA_MIN = float(np.float32(1.0) / np.float32(255.0))
T_EPS = float(np.float32(0.0001))

SH_C0 = 0.28209479177387814
SH_C1 = 0.4886025119029199
SH_C2 = [1.0925484305920792, -1.0925484305920792, 0.31539156525252005, -1.0925484305920792, 0.5462742152960396]
SH_C3 = [-0.5900435899266435, 2.890611442640554, -0.4570457994644658, 0.3731763325901154, -0.4570457994644658,
         1.445305721320277, -0.5900435899266435]


def f32(x):
    return torch.as_tensor(x, dtype=F32)


def fma(a, b, c):
    """fp32 fused multiply-add: exact product (fits fp64), one add in fp64, round to fp32."""
    return (torch.as_tensor(a, dtype=F32).to(F64) * torch.as_tensor(b, dtype=F32).to(F64) +
            torch.as_tensor(c, dtype=F32).to(F64)).to(F32)


def get_higher_msb(n):
    """rasterizer_impl.cu:35-50."""
    msb = 16
    step = msb
    while step > 1:
        step //= 2
        if n >> msb:
            msb += step
        else:
            msb -= step
    if n >> msb:
        msb += 1
    return msb


# ------------------------------------------------------------------------------------------------
# forward preprocess  (forward.cu:20-251, auxiliary.h:67-77,185-292)
# ------------------------------------------------------------------------------------------------
def quat_to_rotmat(q):
    """auxiliary.h:213-235.  q: [P,4] (w,x,y,z).  Returns R as [P,3,3] with R[:, :, c] = column c."""
    n2 = fma(q[:, 2], q[:, 2], fma(q[:, 1], q[:, 1], fma(q[:, 0], q[:, 0], q[:, 3] * q[:, 3])))
    s = (1.0 / torch.sqrt(n2.to(F64))).to(F32)  # rsqrt.approx on the GPU (<= 2 ulp)
    w, x, y, z = q[:, 0] * s, q[:, 1] * s, q[:, 2] * s, q[:, 3] * s
    yy, zz = y * y, z * z
    wz, wy, wx = w * z, w * y, w * x
    xx_zz, xx_yy, yy_zz = fma(x, x, zz), fma(x, x, yy), yy + zz
    # off-diagonal sums are fused by ptxas in the reference build (SASS): x*y +- w*z -> fma(x, y, +-wz)
    a, b = fma(x, y, wz), fma(x, y, -wz)
    c, d = fma(x, z, -wy), fma(x, z, wy)
    e, f = fma(y, z, wx), fma(y, z, -wx)
    one = f32(1.0)
    c0 = torch.stack([one - (yy_zz + yy_zz), a + a, c + c], dim=1)
    c1 = torch.stack([b + b, one - (xx_zz + xx_zz), e + e], dim=1)
    c2 = torch.stack([d + d, f + f, one - (xx_yy + xx_yy)], dim=1)
    return torch.stack([c0, c1, c2], dim=2)


def dot3_ref(a0, a1, a2, b0, b1, b2):
    """a0*b0 + a1*b1 + a2*b2 as the reference build contracts it: fma(a2,b2, fma(a0,b0, a1*b1))."""
    return fma(a2, b2, fma(a0, b0, a1 * b1))


def compute_transmat(means3D, scales, mod, rotations, projmatrix, viewmatrix, W, H):
    """forward.cu:75-115.  Returns T [P,9] (Tu,Tv,Tw) and the view-space normal [P,3]."""
    R = quat_to_rotmat(rotations)
    sx, sy = f32(mod) * scales[:, 0], f32(mod) * scales[:, 1]
    L0 = R[:, :, 0] * sx[:, None]
    L1 = R[:, :, 1] * sy[:, None]
    L2 = R[:, :, 2]
    pm = projmatrix.reshape(-1)
    vm = viewmatrix.reshape(-1)
    px, py, pz = means3D[:, 0], means3D[:, 1], means3D[:, 2]
    A0, A1, A2 = [], [], []
    for c in range(4):
        A0.append(dot3_ref(L0[:, 0], L0[:, 1], L0[:, 2], pm[c], pm[4 + c], pm[8 + c]))
        A1.append(dot3_ref(L1[:, 0], L1[:, 1], L1[:, 2], pm[c], pm[4 + c], pm[8 + c]))
        A2.append(pm[12 + c] + dot3_ref(px, py, pz, pm[c], pm[4 + c], pm[8 + c]))
    hW, hWm1 = f32(float(W)) * 0.5, f32(float(W - 1)) * 0.5
    hH, hHm1 = f32(float(H)) * 0.5, f32(float(H - 1)) * 0.5
    Tu = [fma(hWm1, A[3], hW * A[0]) for A in (A0, A1, A2)]
    Tv = [fma(hHm1, A[3], hH * A[1]) for A in (A0, A1, A2)]
    Tw = [A0[3], A1[3], A2[3]]
    T = torch.stack(Tu + Tv + Tw, dim=1)
    normal = torch.stack([dot3_ref(L2[:, 0], L2[:, 1], L2[:, 2], vm[0 + k], vm[4 + k], vm[8 + k]) for k in range(3)], dim=1)
    return T, normal


def sh_to_rgb(deg, means3D, campos, shs):
    """forward.cu:20-71.  Returns (rgb [P,3] clamped at 0, clamped flags bool [P,3])."""
    d = means3D - campos[None, :]
    ln = torch.sqrt(fma(d[:, 2], d[:, 2], fma(d[:, 0], d[:, 0], d[:, 1] * d[:, 1])))
    x, y, z = d[:, 0] / ln, d[:, 1] / ln, d[:, 2] / ln
    sh = shs
    C = lambda v: f32(v)
    col = lambda t: t[:, None]
    r = C(SH_C0) * sh[:, 0]
    if deg > 0:
        r = fma(-col(y * C(SH_C1)), sh[:, 1], r)
        r = fma(col(z * C(SH_C1)), sh[:, 2], r)
        r = fma(-col(x * C(SH_C1)), sh[:, 3], r)
        if deg > 1:
            xx, yy, zz, xy, yz, xz = x * x, y * y, z * z, x * y, y * z, x * z
            zz2 = zz + zz
            xx_yy = xx - yy
            r = fma(col(xy * C(SH_C2[0])), sh[:, 4], r)
            r = fma(col(yz * C(SH_C2[1])), sh[:, 5], r)
            r = fma(col(((zz2 - xx) - yy) * C(SH_C2[2])), sh[:, 6], r)
            r = fma(col(xz * C(SH_C2[3])), sh[:, 7], r)
            r = fma(col(xx_yy * C(SH_C2[4])), sh[:, 8], r)
            if deg > 2:
                q4 = fma(zz, f32(4.0), -xx) - yy
                r = fma(col((y * C(SH_C3[0])) * fma(xx, f32(3.0), -yy)), sh[:, 9], r)
                r = fma(col(z * (xy * C(SH_C3[1]))), sh[:, 10], r)
                r = fma(col((y * C(SH_C3[2])) * q4), sh[:, 11], r)
                r = fma(col((z * C(SH_C3[3])) * fma(yy, f32(-3.0), fma(xx, f32(-3.0), zz2))), sh[:, 12], r)
                r = fma(col((x * C(SH_C3[4])) * q4), sh[:, 13], r)
                r = fma(col((z * C(SH_C3[5])) * xx_yy), sh[:, 14], r)
                r = fma(col((x * C(SH_C3[6])) * fma(yy, f32(-3.0), xx)), sh[:, 15], r)
    r = r + 0.5
    clamped = r < 0
    return torch.clamp_min(r, 0.0), clamped


def trunc_int(x):
    """C float->int conversion as the GPU does it (cvt.rzi.s32.f32: truncation, saturating, NaN->0)."""
    x = torch.nan_to_num(x.to(F64), nan=0.0, posinf=2147483647.0, neginf=-2147483648.0)
    return torch.clamp(torch.trunc(x), -2147483648.0, 2147483647.0).to(torch.int64)


def get_rect(cx, cy, radius_i, grid_x, grid_y):
    """auxiliary.h:67-77 with an integer radius (converted to float before the subtraction)."""
    rf = radius_i.to(F32)
    s = f32(0.0625)
    minx = torch.clamp(trunc_int((cx - rf) * s), 0, grid_x)
    miny = torch.clamp(trunc_int((cy - rf) * s), 0, grid_y)
    maxx = torch.clamp(trunc_int((((cx + rf) + 16.0) + (-1.0)) * s), 0, grid_x)
    maxy = torch.clamp(trunc_int((((cy + rf) + 16.0) + (-1.0)) * s), 0, grid_y)
    return minx, miny, maxx, maxy


def preprocess_forward(means3D, scales, rotations, opacities, shs, colors_precomp, transMat_precomp, viewmatrix,
                       projmatrix, campos, W, H, sh_degree, scale_modifier=1.0):
    """forward.cu:175-251.  All inputs CPU fp32.  Returns a dict of per-surfel arrays."""
    P = means3D.shape[0]
    vm = viewmatrix.reshape(-1)
    grid_x, grid_y = (W + 15) // 16, (H + 15) // 16
    px, py, pz = means3D[:, 0], means3D[:, 1], means3D[:, 2]
    # auxiliary.h:79-87 (association read from the reference PTX)
    pv = [vm[12 + k] + dot3_ref(px, py, pz, vm[0 + k], vm[4 + k], vm[8 + k]) for k in range(3)]
    valid = pv[2] > f32(NEAR_N)                                           # auxiliary.h:200
    if transMat_precomp is None:
        T, normal = compute_transmat(means3D, scales, scale_modifier, rotations, projmatrix, viewmatrix, W, H)
    else:
        T = transMat_precomp.clone()
        normal = torch.zeros(P, 3, dtype=F32)
        normal[:, 2] = 1.0
    # dual-visible flip (forward.cu:209-214)
    pn = dot3_ref(pv[0], pv[1], pv[2], normal[:, 0], normal[:, 1], normal[:, 2])
    valid &= pn != 0
    mult = torch.where(pn < 0, f32(1.0), f32(-1.0))
    normal = normal * mult[:, None]
    # AABB (forward.cu:119-145), cutoff 3
    Tu, Tv, Tw = T[:, 0:3], T[:, 3:6], T[:, 6:9]
    nine = f32(9.0)
    d = fma(-Tw[:, 2], Tw[:, 2], fma(Tw[:, 0] * Tw[:, 0], nine, (Tw[:, 1] * Tw[:, 1]) * nine))
    valid &= d != 0
    d_safe = torch.where(d != 0, d, f32(1.0))
    inv = f32(1.0) / d_safe
    f9 = inv * nine

    def fdot(a, b):  # <f, a*b> with f = (9/d, 9/d, -1/d), fused as in the reference SASS
        return fma(a[:, 2] * b[:, 2], -inv, fma(f9, a[:, 0] * b[:, 0], f9 * (a[:, 1] * b[:, 1])))

    cx, cy = fdot(Tu, Tw), fdot(Tv, Tw)
    su = fma(Tu[:, 2] * Tu[:, 2], inv, -fma(f9, Tu[:, 0] * Tu[:, 0], f9 * (Tu[:, 1] * Tu[:, 1])))
    sv = fma(Tv[:, 2] * Tv[:, 2], inv, -fma(f9, Tv[:, 0] * Tv[:, 0], f9 * (Tv[:, 1] * Tv[:, 1])))
    hx = torch.sqrt(torch.clamp_min(fma(cx, cx, su), f32(1e-4)))
    hy = torch.sqrt(torch.clamp_min(fma(cy, cy, sv), f32(1e-4)))
    radius = torch.ceil(torch.maximum(torch.maximum(hx, hy), f32(MIN_RADIUS)))
    radius_i = trunc_int(radius)
    minx, miny, maxx, maxy = get_rect(cx, cy, radius_i, grid_x, grid_y)
    area = (maxx - minx) * (maxy - miny)
    valid &= area != 0
    if colors_precomp is None:
        rgb, clamped = sh_to_rgb(sh_degree, means3D, campos, shs)
    else:
        rgb, clamped = colors_precomp.clone(), torch.zeros(P, 3, dtype=torch.bool)
    radii = torch.where(valid, radius_i, torch.zeros_like(radius_i)).to(torch.int32)
    tiles = torch.where(valid, area, torch.zeros_like(area))
    return dict(radii=radii, tiles_touched=tiles, centre=torch.stack([cx, cy], 1), depth=pv[2], T=T, normal=normal,
                opacity=opacities.reshape(-1), rgb=rgb, clamped=clamped, valid=valid,
                rect=torch.stack([minx, miny, maxx, maxy], 1), grid=(grid_x, grid_y))


# ------------------------------------------------------------------------------------------------
# binning  (rasterizer_impl.cu:70-138, 276-319)
# ------------------------------------------------------------------------------------------------
def bin_and_sort(pre):
    """duplicateWithKeys + stable sort + identifyTileRanges.  Returns keys (int64, sorted), point_list, ranges."""
    grid_x, grid_y = pre["grid"]
    vis = torch.nonzero(pre["valid"]).reshape(-1)
    rect = pre["rect"][vis]
    w = rect[:, 2] - rect[:, 0]
    n = w * (rect[:, 3] - rect[:, 1])
    owner = torch.repeat_interleave(torch.arange(vis.numel()), n)
    start = torch.cumsum(n, 0) - n
    t = torch.arange(int(n.sum())) - start[owner]
    y = rect[owner, 1] + t // w[owner]
    x = rect[owner, 0] + t % w[owner]
    tile = y * grid_x + x
    depth_bits = pre["depth"][vis].contiguous().view(torch.int32).to(torch.int64) & 0xFFFFFFFF
    keys = (tile << 32) | depth_bits[owner]
    values = vis[owner]
    # keys are emitted in surfel order; a stable sort reproduces cub's permutation. Depth > 0.2 so the
    # float bits order like unsigned ints and int64 comparison is safe (tile ids < 2^31).
    order = torch.argsort(keys, stable=True)
    keys_s, vals_s = keys[order], values[order]
    Nt = grid_x * grid_y
    ranges = torch.zeros(Nt, 2, dtype=torch.int64)
    R = keys_s.numel()
    if R > 0:
        tl = keys_s >> 32
        first = torch.ones(R, dtype=torch.bool)
        first[1:] = tl[1:] != tl[:-1]
        idx = torch.nonzero(first).reshape(-1)
        ranges[tl[idx], 0] = idx
        ends = torch.cat([idx[1:], torch.tensor([R])])
        ranges[tl[idx], 1] = ends
    return keys_s, vals_s.to(torch.int32), ranges.to(torch.int32)


# ------------------------------------------------------------------------------------------------
# compositing forward  (forward.cu:256-448)
# ------------------------------------------------------------------------------------------------
def _tile_pixels(grid_x, grid_y, W, H):
    t = torch.arange(256)
    tx = torch.arange(grid_x * grid_y) % grid_x
    ty = torch.arange(grid_x * grid_y) // grid_x
    px = tx[:, None] * 16 + (t & 15)[None, :]
    py = ty[:, None] * 16 + (t >> 4)[None, :]
    inside = (px < W) & (py < H)
    return px, py, inside


def _pair_geometry(Tu, Tv, Tw, cxy, pxf, pyf):
    """forward.cu:355-372 in the reference build's association.  Tu,Tv,Tw: [n,3]; cxy [n,2]; pxf,pyf [n,256]."""
    c = lambda a, i: a[:, i, None]
    kx, ky, kz = fma(pxf, c(Tw, 0), -c(Tu, 0)), fma(pxf, c(Tw, 1), -c(Tu, 1)), fma(pxf, c(Tw, 2), -c(Tu, 2))
    lx, ly, lz = fma(pyf, c(Tw, 0), -c(Tv, 0)), fma(pyf, c(Tw, 1), -c(Tv, 1)), fma(pyf, c(Tw, 2), -c(Tv, 2))
    ppx = fma(ky, lz, -(kz * ly))
    ppy = fma(kz, lx, -(kx * lz))
    ppz = fma(kx, ly, -(ky * lx))
    ok = ppz != 0
    ppz_s = torch.where(ok, ppz, f32(1.0))
    sx, sy = ppx / ppz_s, ppy / ppz_s
    rho3d = fma(sx, sx, sy * sy)
    dx, dy = c(cxy, 0) - pxf, c(cxy, 1) - pyf
    r2 = fma(dy, dy, dx * dx)
    rho2d = r2 + r2
    rho = torch.minimum(rho3d, rho2d)
    depth = c(Tw, 2) + fma(c(Tw, 0), sx, c(Tw, 1) * sy)
    ok = ok & ~(depth < f32(NEAR_N))
    power = rho * f32(-0.5)
    ok = ok & ~(power > 0)
    return dict(k=(kx, ky, kz), l=(lx, ly, lz), ppz=ppz_s, s=(sx, sy), rho3d=rho3d, rho2d=rho2d, d=(dx, dy),
                depth=depth, power=power, ok=ok)


def composite_forward(pre, point_list, ranges, bg, W, H):
    """Returns color [3,H,W], allmap [7,H,W], final_T [3,N], n_contrib [2,N] (int32)."""
    grid_x, grid_y = pre["grid"]
    Nt = grid_x * grid_y
    px, py, inside = _tile_pixels(grid_x, grid_y, W, H)
    lens = (ranges[:, 1] - ranges[:, 0]).to(torch.int64)
    order = torch.argsort(lens, descending=True, stable=True)
    lens_o, start_o = lens[order], ranges[order, 0].to(torch.int64)
    pxf, pyf = px[order].to(F32), py[order].to(F32)
    n_all = Nt
    shp = (n_all, 256)
    T = torch.ones(shp, dtype=F32)
    done = ~inside[order]
    contributor = torch.zeros(shp, dtype=torch.int64)
    last_contributor = torch.zeros(shp, dtype=torch.int64)
    C = torch.zeros((3,) + shp, dtype=F32)
    Nn = torch.zeros((3,) + shp, dtype=F32)
    D = torch.zeros(shp, dtype=F32); M1 = torch.zeros(shp, dtype=F32); M2 = torch.zeros(shp, dtype=F32)
    dist = torch.zeros(shp, dtype=F32); med_d = torch.zeros(shp, dtype=F32)
    med_c = torch.full(shp, -1.0, dtype=F32)
    Tm, nrm, opa, rgb, cxy = pre["T"], pre["normal"], pre["opacity"], pre["rgb"], pre["centre"]
    pl = point_list.to(torch.int64)
    maxlen = int(lens_o[0].item()) if Nt > 0 else 0
    neg_lens = -lens_o
    for k in range(maxlen):
        n = int(torch.searchsorted(neg_lens, torch.tensor(-k), right=False).item())  # tiles with len > k
        if n == 0:
            break
        sl = slice(0, n)
        if bool(done[sl].all()):
            # every pixel of every remaining tile has terminated
            break
        ids = pl[start_o[sl] + k]
        g = _pair_geometry(Tm[ids, 0:3], Tm[ids, 3:6], Tm[ids, 6:9], cxy[ids], pxf[sl], pyf[sl])
        live = ~done[sl]
        contributor[sl] += live.to(torch.int64)
        G = torch.exp(g["power"])
        alpha = torch.minimum(opa[ids][:, None] * G, f32(ALPHA_MAX))
        ok = g["ok"] & live & ~(alpha < f32(ALPHA_MIN))
        Tc = T[sl]
        test_T = Tc * (f32(1.0) - alpha)
        stop = ok & (test_T < f32(T_EPS))
        done[sl] |= stop
        ok = ok & ~stop
        w = torch.where(ok, Tc * alpha, f32(0.0))
        depth = g["depth"]
        A = f32(1.0) - Tc
        m = (f32(-NEAR_N) / depth + f32(1.0)) * f32(FAR_RATIO)
        m = torch.where(ok, m, f32(0.0))
        depth0 = torch.where(ok, depth, f32(0.0))
        mm = m * m
        dist[sl] = fma(w, fma(-M1[sl], m + m, fma(A, mm, M2[sl])), dist[sl])
        D[sl] = fma(depth0, w, D[sl])
        M1[sl] = fma(m, w, M1[sl])
        M2[sl] = fma(mm, w, M2[sl])
        med = ok & (Tc > 0.5)
        med_d[sl] = torch.where(med, depth, med_d[sl])
        med_c[sl] = torch.where(med, contributor[sl].to(F32), med_c[sl])
        for ch in range(3):
            Nn[ch][sl] = fma(w, nrm[ids, ch][:, None], Nn[ch][sl])
            C[ch][sl] = fma(w, rgb[ids, ch][:, None], C[ch][sl])
        T[sl] = torch.where(ok, test_T, Tc)
        last_contributor[sl] = torch.where(ok, contributor[sl], last_contributor[sl])

    # scatter back to image layout
    inv = torch.empty_like(order)
    inv[order] = torch.arange(Nt)

    def to_img(a):  # [Nt(order),256] -> [H,W]
        a = a[inv].reshape(grid_y, grid_x, 16, 16).permute(0, 2, 1, 3).reshape(grid_y * 16, grid_x * 16)
        return a[:H, :W].contiguous()

    Timg = to_img(T)
    color = torch.stack([fma(Timg, bg[ch], to_img(C[ch])) for ch in range(3)])
    allmap = torch.stack([to_img(D), f32(1.0) - Timg, to_img(Nn[0]), to_img(Nn[1]), to_img(Nn[2]), to_img(med_d),
                          to_img(dist)])
    final_T = torch.stack([Timg, to_img(M1), to_img(M2)]).reshape(3, -1)
    n_contrib = torch.stack([to_img(last_contributor), to_img(torch.clamp_min(med_c, 0).to(torch.int64))]).reshape(2, -1)
    return color, allmap, final_T, n_contrib.to(torch.int32)


# ------------------------------------------------------------------------------------------------
# compositing backward  (backward.cu:143-446) -- explicit restatement, NOT autograd of the forward
# ------------------------------------------------------------------------------------------------
def composite_backward(pre, point_list, ranges, bg, W, H, final_T, n_contrib, dL_dpix, dL_dothers):
    """Returns per-surfel accumulations (float64): dL_dT [P,9], dL_dcentre [P,2], dL_dcolor [P,3],
    dL_dnormal [P,3], dL_dopacity [P]."""
    grid_x, grid_y = pre["grid"]
    Nt = grid_x * grid_y
    P = pre["T"].shape[0]
    px, py, inside = _tile_pixels(grid_x, grid_y, W, H)
    lens = (ranges[:, 1] - ranges[:, 0]).to(torch.int64)
    order = torch.argsort(lens, descending=True, stable=True)
    lens_o, start_o = lens[order], ranges[order, 0].to(torch.int64)
    pxf, pyf = px[order].to(F32), py[order].to(F32)
    ins = inside[order]

    def from_img(img):  # [H,W] -> [Nt(order),256], zero outside
        pad = torch.zeros(grid_y * 16, grid_x * 16, dtype=img.dtype)
        pad[:H, :W] = img
        a = pad.reshape(grid_y, 16, grid_x, 16).permute(0, 2, 1, 3).reshape(Nt, 256)
        return a[order]

    T_final = from_img(final_T[0].reshape(H, W))
    final_D = from_img(final_T[1].reshape(H, W))
    final_D2 = from_img(final_T[2].reshape(H, W))
    final_A = f32(1.0) - T_final
    last_c = from_img(n_contrib[0].reshape(H, W).to(torch.int64))
    med_c = from_img(n_contrib[1].reshape(H, W).to(torch.int64))
    dpix = [from_img(dL_dpix[ch]) for ch in range(3)]
    d_depth, d_acc = from_img(dL_dothers[0]), from_img(dL_dothers[1])
    d_nrm = [from_img(dL_dothers[2 + ch]) for ch in range(3)]
    d_med, d_reg = from_img(dL_dothers[5]), from_img(dL_dothers[6])
    bg_dot = bg[0] * dpix[0] + bg[1] * dpix[1] + bg[2] * dpix[2]

    shp = (Nt, 256)
    z = lambda: torch.zeros(shp, dtype=F32)
    T = T_final.clone()
    accum_rec = [z(), z(), z()]; last_color = [z(), z(), z()]
    last_alpha = z(); last_depth = z(); last_normal = [z(), z(), z()]
    accum_depth_rec = z(); accum_alpha_rec = z(); accum_normal_rec = [z(), z(), z()]
    last_dL_dT = z()

    g_T = torch.zeros(P, 9, dtype=F64); g_c = torch.zeros(P, 2, dtype=F64); g_col = torch.zeros(P, 3, dtype=F64)
    g_n = torch.zeros(P, 3, dtype=F64); g_o = torch.zeros(P, dtype=F64)
    Tm, nrm, opa, rgb, cxy = pre["T"], pre["normal"], pre["opacity"], pre["rgb"], pre["centre"]
    pl = point_list.to(torch.int64)
    maxlen = int(lens_o[0].item()) if Nt > 0 else 0
    neg_lens = -lens_o
    one = f32(1.0)
    for j in range(maxlen):
        n = int(torch.searchsorted(neg_lens, torch.tensor(-j), right=False).item())
        if n == 0:
            break
        sl = slice(0, n)
        pos = lens_o[sl] - 1 - j                      # == reference `contributor` after decrement
        act = ins[sl] & (pos[:, None] < last_c[sl])
        if not bool(act.any()):
            continue
        ids = pl[start_o[sl] + pos]
        Tu, Tv, Tw = Tm[ids, 0:3], Tm[ids, 3:6], Tm[ids, 6:9]
        g = _pair_geometry(Tu, Tv, Tw, cxy[ids], pxf[sl], pyf[sl])
        G = torch.exp(g["power"])
        o = opa[ids][:, None]
        alpha = torch.minimum(o * G, f32(ALPHA_MAX))
        act = act & g["ok"] & ~(alpha < f32(ALPHA_MIN))
        if not bool(act.any()):
            continue
        sx, sy = g["s"]; c_d = g["depth"]; ppz = g["ppz"]
        kx, ky, kz = g["k"]; lx, ly, lz = g["l"]; dx, dy = g["d"]
        alpha = torch.where(act, alpha, f32(0.0))     # inactive pairs leave every state untouched
        Tn = torch.where(act, T[sl] / (one - alpha), T[sl])
        T[sl] = Tn
        w = alpha * Tn
        la = last_alpha[sl]
        dL_dalpha = torch.zeros((n, 256), dtype=F32)
        for ch in range(3):
            c = rgb[ids, ch][:, None].expand(n, 256)
            ar = torch.where(act, la * last_color[ch][sl] + (one - la) * accum_rec[ch][sl], accum_rec[ch][sl])
            accum_rec[ch][sl] = ar
            last_color[ch][sl] = torch.where(act, c, last_color[ch][sl])
            dL_dalpha = dL_dalpha + (c - ar) * dpix[ch][sl]
            g_col[:, ch].index_add_(0, ids, torch.where(act, w * dpix[ch][sl], f32(0.0)).to(F64).sum(1))
        m_d = f32(FAR_RATIO) * (one - f32(NEAR_N) / c_d)
        dmd_dd = f32(FAR_N * NEAR_N) / ((f32(FAR_N - NEAR_N) * c_d) * c_d)
        dL_dz = torch.where(pos[:, None] == med_c[sl] - 1, d_med[sl], f32(0.0))
        dL_dweight = (final_D2[sl] + m_d * m_d * final_A[sl] - f32(2.0) * m_d * final_D[sl]) * d_reg[sl]
        dL_dalpha = dL_dalpha + (dL_dweight - last_dL_dT[sl])
        last_dL_dT[sl] = torch.where(act, dL_dweight * alpha + (one - alpha) * last_dL_dT[sl], last_dL_dT[sl])
        dL_dmd = f32(2.0) * (Tn * alpha) * (m_d * final_A[sl] - final_D[sl]) * d_reg[sl]
        dL_dz = dL_dz + dL_dmd * dmd_dd
        adr = torch.where(act, la * last_depth[sl] + (one - la) * accum_depth_rec[sl], accum_depth_rec[sl])
        accum_depth_rec[sl] = adr
        last_depth[sl] = torch.where(act, c_d, last_depth[sl])
        dL_dalpha = dL_dalpha + (c_d - adr) * d_depth[sl]
        aar = torch.where(act, la + (one - la) * accum_alpha_rec[sl], accum_alpha_rec[sl])
        accum_alpha_rec[sl] = aar
        dL_dalpha = dL_dalpha + (one - aar) * d_acc[sl]
        for ch in range(3):
            nv = nrm[ids, ch][:, None].expand(n, 256)
            anr = torch.where(act, la * last_normal[ch][sl] + (one - la) * accum_normal_rec[ch][sl],
                              accum_normal_rec[ch][sl])
            accum_normal_rec[ch][sl] = anr
            last_normal[ch][sl] = torch.where(act, nv, last_normal[ch][sl])
            dL_dalpha = dL_dalpha + (nv - anr) * d_nrm[ch][sl]
            g_n[:, ch].index_add_(0, ids, torch.where(act, alpha * Tn * d_nrm[ch][sl], f32(0.0)).to(F64).sum(1))
        dL_dalpha = dL_dalpha * Tn
        last_alpha[sl] = torch.where(act, alpha, la)
        dL_dalpha = dL_dalpha + (-T_final[sl] / (one - alpha)) * bg_dot[sl]
        dL_dG = o * dL_dalpha
        dL_dz = dL_dz + alpha * Tn * d_depth[sl]

        brA = act & (g["rho3d"] <= g["rho2d"])
        brB = act & ~(g["rho3d"] <= g["rho2d"])
        dL_dsx = dL_dG * -G * sx + dL_dz * Tw[:, 0, None]
        dL_dsy = dL_dG * -G * sy + dL_dz * Tw[:, 1, None]
        dsx_pz, dsy_pz = dL_dsx / ppz, dL_dsy / ppz
        dpx_, dpy_, dpz_ = dsx_pz, dsy_pz, -(dsx_pz * sx + dsy_pz * sy)
        dkx, dky, dkz = ly * dpz_ - lz * dpy_, lz * dpx_ - lx * dpz_, lx * dpy_ - ly * dpx_
        dlx, dly, dlz = dpy_ * kz - dpz_ * ky, dpz_ * kx - dpx_ * kz, dpx_ * ky - dpy_ * kx
        zero = f32(0.0)
        contrib_T = [
            torch.where(brA, -dkx, zero), torch.where(brA, -dky, zero), torch.where(brA, -dkz, zero),
            torch.where(brA, -dlx, zero), torch.where(brA, -dly, zero), torch.where(brA, -dlz, zero),
            torch.where(brA, pxf[sl] * dkx + pyf[sl] * dlx + dL_dz * sx, torch.where(brB, sx * dL_dz, zero)),
            torch.where(brA, pxf[sl] * dky + pyf[sl] * dly + dL_dz * sy, torch.where(brB, sy * dL_dz, zero)),
            torch.where(brA, pxf[sl] * dkz + pyf[sl] * dlz + dL_dz, torch.where(brB, dL_dz, zero)),
        ]
        for q, v in enumerate(contrib_T):
            g_T[:, q].index_add_(0, ids, v.to(F64).sum(1))
        dG_ddelx = -G * f32(2.0) * dx
        dG_ddely = -G * f32(2.0) * dy
        g_c[:, 0].index_add_(0, ids, torch.where(brB, dL_dG * dG_ddelx, zero).to(F64).sum(1))
        g_c[:, 1].index_add_(0, ids, torch.where(brB, dL_dG * dG_ddely, zero).to(F64).sum(1))
        g_o.index_add_(0, ids, torch.where(act, G * dL_dalpha, zero).to(F64).sum(1))
    return dict(dL_dT=g_T, dL_dcentre=g_c, dL_dcolor=g_col, dL_dnormal=g_n, dL_dopacity=g_o)


# ------------------------------------------------------------------------------------------------
# backward preprocess  (backward.cu:20-139, 449-636; auxiliary.h:128-138, 238-282)
# ------------------------------------------------------------------------------------------------
def bwd_image_size(W, H, tanfovx, tanfovy):
    """backward.cu:613-614 with focal from rasterizer_impl.cu:387-388, all in fp32."""
    tx, ty = np.float32(tanfovx), np.float32(tanfovy)
    focal_x = np.float32(W) / (np.float32(2.0) * tx)
    focal_y = np.float32(H) / (np.float32(2.0) * ty)
    Wb = int(np.float32(np.float32(focal_x * tx) * np.float32(2.0)))
    Hb = int(np.float32(np.float32(focal_y * ty) * np.float32(2.0)))
    return Wb, Hb


def sh_backward(deg, M, means3D, campos, shs, clamped, dL_dcolor):
    """backward.cu:20-139.  Returns dL_dsh [P,M,3] and the extra dL_dmean3D [P,3]."""
    P = means3D.shape[0]
    d = (means3D - campos[None, :]).to(F32)
    ln = torch.sqrt((d * d).sum(1))
    x, y, z = d[:, 0] / ln, d[:, 1] / ln, d[:, 2] / ln
    g = dL_dcolor.clone().to(F32)
    g[clamped] = 0.0
    out = torch.zeros(P, M, 3, dtype=F32)
    col = lambda t: t[:, None]
    sh = shs
    dx = torch.zeros(P, 3, dtype=F32); dy = torch.zeros(P, 3, dtype=F32); dz = torch.zeros(P, 3, dtype=F32)
    out[:, 0] = SH_C0 * g
    if deg > 0:
        out[:, 1] = col(-SH_C1 * y) * g
        out[:, 2] = col(SH_C1 * z) * g
        out[:, 3] = col(-SH_C1 * x) * g
        dx = -SH_C1 * sh[:, 3]
        dy = -SH_C1 * sh[:, 1]
        dz = SH_C1 * sh[:, 2]
        if deg > 1:
            xx, yy, zz, xy, yz, xz = x * x, y * y, z * z, x * y, y * z, x * z
            out[:, 4] = col(SH_C2[0] * xy) * g
            out[:, 5] = col(SH_C2[1] * yz) * g
            out[:, 6] = col(SH_C2[2] * (2.0 * zz - xx - yy)) * g
            out[:, 7] = col(SH_C2[3] * xz) * g
            out[:, 8] = col(SH_C2[4] * (xx - yy)) * g
            dx = dx + col(SH_C2[0] * y) * sh[:, 4] + col(SH_C2[2] * 2.0 * -x) * sh[:, 6] + col(SH_C2[3] * z) * sh[:, 7] + col(SH_C2[4] * 2.0 * x) * sh[:, 8]
            dy = dy + col(SH_C2[0] * x) * sh[:, 4] + col(SH_C2[1] * z) * sh[:, 5] + col(SH_C2[2] * 2.0 * -y) * sh[:, 6] + col(SH_C2[4] * 2.0 * -y) * sh[:, 8]
            dz = dz + col(SH_C2[1] * y) * sh[:, 5] + col(SH_C2[2] * 2.0 * 2.0 * z) * sh[:, 6] + col(SH_C2[3] * x) * sh[:, 7]
            if deg > 2:
                out[:, 9] = col(SH_C3[0] * y * (3.0 * xx - yy)) * g
                out[:, 10] = col(SH_C3[1] * xy * z) * g
                out[:, 11] = col(SH_C3[2] * y * (4.0 * zz - xx - yy)) * g
                out[:, 12] = col(SH_C3[3] * z * (2.0 * zz - 3.0 * xx - 3.0 * yy)) * g
                out[:, 13] = col(SH_C3[4] * x * (4.0 * zz - xx - yy)) * g
                out[:, 14] = col(SH_C3[5] * z * (xx - yy)) * g
                out[:, 15] = col(SH_C3[6] * x * (xx - 3.0 * yy)) * g
                dx = dx + (col(SH_C3[0] * 3.0 * 2.0 * xy) * sh[:, 9] + col(SH_C3[1] * yz) * sh[:, 10] +
                           col(SH_C3[2] * -2.0 * xy) * sh[:, 11] + col(SH_C3[3] * -3.0 * 2.0 * xz) * sh[:, 12] +
                           col(SH_C3[4] * (-3.0 * xx + 4.0 * zz - yy)) * sh[:, 13] + col(SH_C3[5] * 2.0 * xz) * sh[:, 14] +
                           col(SH_C3[6] * 3.0 * (xx - yy)) * sh[:, 15])
                dy = dy + (col(SH_C3[0] * 3.0 * (xx - yy)) * sh[:, 9] + col(SH_C3[1] * xz) * sh[:, 10] +
                           col(SH_C3[2] * (-3.0 * yy + 4.0 * zz - xx)) * sh[:, 11] + col(SH_C3[3] * -3.0 * 2.0 * yz) * sh[:, 12] +
                           col(SH_C3[4] * -2.0 * xy) * sh[:, 13] + col(SH_C3[5] * -2.0 * yz) * sh[:, 14] +
                           col(SH_C3[6] * -3.0 * 2.0 * xy) * sh[:, 15])
                dz = dz + (col(SH_C3[1] * xy) * sh[:, 10] + col(SH_C3[2] * 4.0 * 2.0 * yz) * sh[:, 11] +
                           col(SH_C3[3] * 3.0 * (2.0 * zz - xx - yy)) * sh[:, 12] + col(SH_C3[4] * 4.0 * 2.0 * xz) * sh[:, 13] +
                           col(SH_C3[5] * (xx - yy)) * sh[:, 14])
    dL_ddir = torch.stack([(dx * g).sum(1), (dy * g).sum(1), (dz * g).sum(1)], 1)
    # dnormvdv (auxiliary.h:128-138)
    v, dv = d, dL_ddir
    sum2 = (v * v).sum(1)
    inv32 = 1.0 / torch.sqrt(sum2 * sum2 * sum2)
    dm = torch.stack([
        ((sum2 - v[:, 0] * v[:, 0]) * dv[:, 0] - v[:, 1] * v[:, 0] * dv[:, 1] - v[:, 2] * v[:, 0] * dv[:, 2]) * inv32,
        (-v[:, 0] * v[:, 1] * dv[:, 0] + (sum2 - v[:, 1] * v[:, 1]) * dv[:, 1] - v[:, 2] * v[:, 1] * dv[:, 2]) * inv32,
        (-v[:, 0] * v[:, 2] * dv[:, 0] - v[:, 1] * v[:, 2] * dv[:, 1] + (sum2 - v[:, 2] * v[:, 2]) * dv[:, 2]) * inv32], 1)
    return out, dm


def preprocess_backward(pre, acc, means3D, scales, rotations, shs, transMat_precomp, viewmatrix, projmatrix, campos,
                        W, H, tanfovx, tanfovy, sh_degree):
    """backward.cu:449-636 incl. the deviations of SURVEY.md A.5.  `acc` is composite_backward's dict.
    Returns the 8 public gradients (fp32)."""
    P = means3D.shape[0]
    vis = pre["radii"] > 0
    Wb, Hb = bwd_image_size(W, H, tanfovx, tanfovy)
    vm = viewmatrix.reshape(-1).to(F32)
    pm = projmatrix.reshape(-1).to(F32)
    dT = acc["dL_dT"].to(F32)
    gc = acc["dL_dcentre"].to(F32)
    dcolor = acc["dL_dcolor"].to(F32)
    dnormal = acc["dL_dnormal"].to(F32)
    precomp = transMat_precomp is not None
    if precomp:
        T = transMat_precomp.to(F32)
    else:
        # recompute R, L (scale modifier forced to 1, backward.cu:487) and T = M^T (world2ndc ndc2pix)
        q = rotations
        s = 1.0 / torch.sqrt((q * q).sum(1))
        w, x, y, z = q[:, 0] * s, q[:, 1] * s, q[:, 2] * s, q[:, 3] * s
        R0 = torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y + w * z), 2 * (x * z - w * y)], 1)
        R1 = torch.stack([2 * (x * y - w * z), 1 - 2 * (x * x + z * z), 2 * (y * z + w * x)], 1)
        R2 = torch.stack([2 * (x * z + w * y), 2 * (y * z - w * x), 1 - 2 * (x * x + y * y)], 1)
        L0, L1 = R0 * scales[:, 0:1], R1 * scales[:, 1:2]
        hW, hWm1, hH, hHm1 = Wb / 2.0, (Wb - 1) / 2.0, Hb / 2.0, (Hb - 1) / 2.0
        Pm = torch.zeros(3, 4, dtype=F32)
        for r in range(4):
            Pm[0, r] = pm[4 * r + 0] * hW + pm[4 * r + 3] * hWm1
            Pm[1, r] = pm[4 * r + 1] * hH + pm[4 * r + 3] * hHm1
            Pm[2, r] = pm[4 * r + 3]
        rows = []
        for k in range(3):
            rows += [L0 @ Pm[k, :3], L1 @ Pm[k, :3], means3D @ Pm[k, :3] + Pm[k, 3]]
        T = torch.stack(rows, 1)
        normal = torch.stack([vm[0 + k] * R2[:, 0] + vm[4 + k] * R2[:, 1] + vm[8 + k] * R2[:, 2] for k in range(3)], 1)
    T0, T1, T2 = T[:, 0:3], T[:, 3:6], T[:, 6:9]
    dT0, dT1, dT2 = dT[:, 0:3].clone(), dT[:, 3:6].clone(), dT[:, 6:9].clone()
    chained = (gc[:, 0] != 0) | (gc[:, 1] != 0)
    tv = torch.tensor([9.0, 9.0, -1.0], dtype=F32)
    d = (tv * T2 * T2).sum(1)
    d = torch.where(chained, d, f32(1.0))
    fv = tv[None, :] * (1.0 / d)[:, None]
    gx, gy = gc[:, 0:1], gc[:, 1:2]
    a0, a1 = gx * fv * T2, gy * fv * T2
    a3 = gx * fv * T0 + gy * fv * T1
    dL_df = gx * T0 * T2 + gy * T1 * T2
    dL_dd = (dL_df * fv).sum(1) * (-1.0 / d)
    a3 = a3 + dL_dd[:, None] * (tv * T2 * 2.0)
    m = chained[:, None].to(F32)
    dT0, dT1, dT2 = dT0 + m * a0, dT1 + m * a1, dT2 + m * a3

    z3 = torch.zeros(P, 3, dtype=F32)
    dmeans3D = z3.clone(); dscales = torch.zeros(P, 2, dtype=F32); drot = torch.zeros(P, 4, dtype=F32)
    dtransMat = dT.clone()
    if precomp:
        dtransMat = torch.where(chained[:, None], torch.cat([dT0, dT1, dT2], 1), dT)
    else:
        dTk = [dT0, dT1, dT2]
        dM = []
        for c in range(3):
            dM.append(torch.stack([Pm[0, r] * dTk[0][:, c] + Pm[1, r] * dTk[1][:, c] + Pm[2, r] * dTk[2][:, c]
                                   for r in range(3)], 1))
        dtn = torch.stack([vm[4 * r + 0] * dnormal[:, 0] + vm[4 * r + 1] * dnormal[:, 1] + vm[4 * r + 2] * dnormal[:, 2]
                           for r in range(3)], 1)
        pv = torch.stack([vm[0 + k] * means3D[:, 0] + vm[4 + k] * means3D[:, 1] + vm[8 + k] * means3D[:, 2] + vm[12 + k]
                          for k in range(3)], 1)
        cosv = -(pv * normal).sum(1)
        mult = torch.where(cosv > 0, f32(1.0), f32(-1.0))
        dtn = dtn * mult[:, None]
        dRS0, dRS1, dRS2 = dM[0], dM[1], dtn
        vR0, vR1, vR2 = dRS0 * scales[:, 0:1], dRS1 * scales[:, 1:2], dRS2
        X, Y, Z = 0, 1, 2
        drot = torch.stack([
            2 * (x * (vR1[:, Z] - vR2[:, Y]) + y * (vR2[:, X] - vR0[:, Z]) + z * (vR0[:, Y] - vR1[:, X])),
            2 * (-2 * x * (vR1[:, Y] + vR2[:, Z]) + y * (vR0[:, Y] + vR1[:, X]) + z * (vR0[:, Z] + vR2[:, X]) + w * (vR1[:, Z] - vR2[:, Y])),
            2 * (x * (vR0[:, Y] + vR1[:, X]) - 2 * y * (vR0[:, X] + vR2[:, Z]) + z * (vR1[:, Z] + vR2[:, Y]) + w * (vR2[:, X] - vR0[:, Z])),
            2 * (x * (vR0[:, Z] + vR2[:, X]) + y * (vR1[:, Z] + vR2[:, Y]) - 2 * z * (vR0[:, X] + vR1[:, Y]) + w * (vR0[:, Y] - vR1[:, X]))], 1)
        dscales = torch.stack([(dRS0 * R0).sum(1), (dRS1 * R1).sum(1)], 1)
        dmeans3D = dM[2]
    M = shs.shape[1] if shs is not None else 0
    dsh = torch.zeros(P, M, 3, dtype=F32)
    if shs is not None:
        dsh, add = sh_backward(sh_degree, M, means3D, campos, shs, pre["clamped"], dcolor)
        dmeans3D = dmeans3D + add
    depth = pre["T"][:, 8]
    a2 = dtransMat[:, 2] if precomp else dT[:, 2]
    a5 = dtransMat[:, 5] if precomp else dT[:, 5]
    dmeans2D = torch.stack([a2 * depth * 0.5 * float(Wb), a5 * depth * 0.5 * float(Hb), torch.zeros(P)], 1).to(F32)
    v = vis[:, None].to(F32)
    return dict(dL_dmeans3D=dmeans3D * v, dL_dmeans2D=dmeans2D * v, dL_dsh=dsh * v[:, :, None],
                dL_dcolors=dcolor * v, dL_dopacity=(acc["dL_dopacity"].to(F32) * vis.to(F32))[:, None],
                dL_dscales=dscales * v, dL_drotations=drot * v, dL_dtransMat=dtransMat * v)


# ------------------------------------------------------------------------------------------------
# top level: what `_C.rasterize_gaussians` / `_C.rasterize_gaussians_backward` compute
# ------------------------------------------------------------------------------------------------
def rasterize_forward(means3D, opacities, viewmatrix, projmatrix, campos, W, H, tanfovx, tanfovy, bg, sh_degree=0,
                      shs=None, colors_precomp=None, scales=None, rotations=None, transMat_precomp=None,
                      scale_modifier=1.0):
    t = lambda a: None if a is None else a.detach().to("cpu", F32).contiguous()
    means3D, opacities, shs, colors_precomp, scales, rotations, transMat_precomp, viewmatrix, projmatrix, campos, bg = map(
        t, (means3D, opacities, shs, colors_precomp, scales, rotations, transMat_precomp, viewmatrix, projmatrix, campos, bg))
    P = means3D.shape[0]
    if P == 0:  # rasterize_points.cu:99 -- zero images, no launch
        return dict(color=torch.zeros(3, H, W), allmap=torch.zeros(7, H, W), radii=torch.zeros(0, dtype=torch.int32),
                    num_rendered=0, pre=None)
    pre = preprocess_forward(means3D, scales, rotations, opacities, shs, colors_precomp, transMat_precomp, viewmatrix,
                             projmatrix, campos, W, H, sh_degree, scale_modifier)
    keys, point_list, ranges = bin_and_sort(pre)
    color, allmap, final_T, n_contrib = composite_forward(pre, point_list, ranges, bg, W, H)
    return dict(color=color, allmap=allmap, radii=pre["radii"], num_rendered=int(keys.numel()), keys=keys,
                point_list=point_list, ranges=ranges, final_T=final_T, n_contrib=n_contrib, pre=pre,
                inputs=dict(means3D=means3D, opacities=opacities, shs=shs, colors_precomp=colors_precomp, scales=scales,
                            rotations=rotations, transMat_precomp=transMat_precomp, viewmatrix=viewmatrix,
                            projmatrix=projmatrix, campos=campos, bg=bg, W=W, H=H, tanfovx=tanfovx, tanfovy=tanfovy,
                            sh_degree=sh_degree))


def rasterize_backward(fwd, dL_dcolor, dL_dallmap):
    i = fwd["inputs"]
    W, H = i["W"], i["H"]
    dL_dcolor = dL_dcolor.detach().to("cpu", F32)
    dL_dallmap = dL_dallmap.detach().to("cpu", F32)
    acc = composite_backward(fwd["pre"], fwd["point_list"], fwd["ranges"], i["bg"], W, H, fwd["final_T"],
                             fwd["n_contrib"], dL_dcolor, dL_dallmap)
    return preprocess_backward(fwd["pre"], acc, i["means3D"], i["scales"], i["rotations"], i["shs"],
                               i["transMat_precomp"], i["viewmatrix"], i["projmatrix"], i["campos"], W, H,
                               i["tanfovx"], i["tanfovy"], i["sh_degree"])


def mark_visible(means3D, viewmatrix):
    """rasterizer_impl.cu:54-66."""
    vm = viewmatrix.reshape(-1).to(F32)
    p = means3D.to(F32)
    vz = vm[14] + dot3_ref(p[:, 0], p[:, 1], p[:, 2], vm[2], vm[6], vm[10])
    return vz > f32(NEAR_N)


# ------------------------------------------------------------------------------------------------
# distCUDA2  (simple_knn.cu:132-184): exact mean of the 3 smallest squared distances to other points
# ------------------------------------------------------------------------------------------------
def dist2_ref(dx, dy, dz):
    """simple_knn.cu:135-136 as contracted by the reference build: fma(dz,dz, fma(dx,dx, dy*dy))."""
    return fma(dz, dz, fma(dx, dx, dy * dy))


def knn_mean_dist2(points, chunk=1024, candidates=None):
    """Brute force for small P (chunked), or exact re-ranking of k-d tree candidates for large P."""
    pts = points.detach().to("cpu", F32).contiguous()
    P = pts.shape[0]
    out = torch.empty(P, dtype=F32)
    FLT_MAX = float(np.finfo(np.float32).max)
    if candidates is None and P > 20000:
        from scipy.spatial import cKDTree
        k = min(P, 12)
        _, nn = cKDTree(pts.double().numpy()).query(pts.double().numpy(), k=k)
        nn = torch.from_numpy(nn.astype(np.int64))
        d = pts[nn] - pts[:, None, :]
        d2 = dist2_ref(d[..., 0], d[..., 1], d[..., 2])
        d2[nn == torch.arange(P)[:, None]] = FLT_MAX
        best = torch.sort(d2, dim=1).values[:, :3]
        return ((best[:, 0] + best[:, 1]) + best[:, 2]) / f32(3.0)
    for s in range(0, P, chunk):
        q = pts[s:s + chunk]
        d = pts[None, :, :] - q[:, None, :]
        d2 = dist2_ref(d[..., 0], d[..., 1], d[..., 2])
        d2[torch.arange(q.shape[0]), torch.arange(s, s + q.shape[0])] = FLT_MAX
        if P < 4:
            d2 = torch.cat([d2, torch.full((q.shape[0], 3), FLT_MAX)], 1)
        best = torch.sort(d2, dim=1).values[:, :3]
        out[s:s + chunk] = ((best[:, 0] + best[:, 1]) + best[:, 2]) / f32(3.0)
    return out
