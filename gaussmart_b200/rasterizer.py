"""Host-side mirror of the reference's `diff_surfel_rasterization` package
(submodules/diff-surfel-rasterization/diff_surfel_rasterization/__init__.py:21-222) on top of the
C ABI of libgaussmart_b200.so.

Same names, argument order, return signature, validation messages and autograd contract as the
reference, so gaussian_renderer/__init__.py:render() and scene/gaussian_model.py run unchanged:

    GaussianRasterizationSettings(image_height, image_width, tanfovx, tanfovy, bg, scale_modifier,
                                  viewmatrix, projmatrix, sh_degree, campos, prefiltered, debug)
    GaussianRasterizer(raster_settings)(means3D, means2D, opacities, shs=None, colors_precomp=None,
                                        scales=None, rotations=None, cov3D_precomp=None)
        -> (color[3,H,W], radii[P] int32, allmap[7,H,W])

PyTorch is used for device memory, streams and autograd plumbing only; every kernel is in the
shared library.  There is no CPU path: CPU tensors raise, as in rasterize_points.cu:27-28.
"""
import ctypes
import os
from typing import NamedTuple

import torch
import torch.nn as nn

from . import _lib


def cpu_deep_copy_tuple(input_tuple):
    copied_tensors = [item.cpu().clone() if isinstance(item, torch.Tensor) else item for item in input_tuple]
    return tuple(copied_tensors)


def _ptr(t):
    """Device pointer of a tensor, NULL for None / empty (the reference's empty tensors select the
    in-kernel path through null pointers, dsr.py:198-208)."""
    if t is None or t.numel() == 0:
        return None
    return ctypes.c_void_p(t.data_ptr())


def _check_cuda(name, t):
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor")


def _f32c(name, t):
    if t.dtype != torch.float32:
        raise RuntimeError(f"{name} must be float32, got {t.dtype}")
    return t.contiguous()


# ---- sizing of the binning workspace without a per-forward host stall -------------------------------------------------
# The reference reads R = num_rendered back with a blocking cudaMemcpy in the middle of every forward
# (rasterizer_impl.cu:282) to size its binning buffer.  Here the workspace is sized from a grow-only high-water mark of
# R per device, the whole forward is enqueued at once (gsm_forward's asynchronous mode), and R comes back through a
# pinned status slot behind the scan.  WHEN the host looks at it is the policy:
#   "event" (default) after everything is enqueued, wait for the slot's event -- it sits behind the SCAN, so the host
#           resumes while scatter / sort / compositing are still queued or running and the device never idles; on
#           overflow (R > capacity: the device rendered an empty frame) the forward is issued again with a larger
#           workspace before anything is returned.  Exactly the reference's semantics.
#   "lazy"  never wait in the forward: the status is read at backward entry, on the next forward of the device, or when
#           `num_rendered` is converted to an int.  An overflow (R more than doubled against everything seen before)
#           can then only be REPORTED -- the frame was rendered empty -- so it raises; for loops that own the whole
#           iteration (view_parallel) and replay it.  Opt in with set_sync_policy("lazy") / GAUSSMART_B200_SYNC=lazy.
#   "sync"  the reference's behaviour: gsm_forward's synchronous mode (one stream synchronisation per forward).
_binning_capacity = {}
_sync_policy = os.environ.get("GAUSSMART_B200_SYNC", "event")
_HEADROOM = {"event": 1.25, "sync": 1.25, "lazy": 2.0}


def set_sync_policy(policy):
    """'event' | 'lazy' | 'sync' (see above); returns the previous policy."""
    global _sync_policy
    if policy not in _HEADROOM:
        raise ValueError(f"unknown sync policy {policy!r}")
    prev, _sync_policy = _sync_policy, policy
    return prev


class BinningOverflow(RuntimeError):
    pass


class _StatusSlots:
    """Pool of gsm_status slots (16 B pinned + one event each), recycled."""
    def __init__(self):
        self.free = []

    def take(self):
        if self.free:
            return self.free.pop()
        h = ctypes.c_void_p()
        _lib.check(_lib.load().gsm_status_create(ctypes.byref(h)), "status_create")
        return h

    def give(self, h):
        self.free.append(h)


_slots = _StatusSlots()
_pending = {}  # device index -> list of unresolved PendingForward (lazy policy)


class PendingForward:
    """Status of one asynchronously issued forward: resolves to R = num_rendered."""
    def __init__(self, slot, key, capacity, prefiltered):
        self.slot, self.key, self.capacity, self.prefiltered = slot, key, capacity, prefiltered
        self.value = None
        self.overflow = False

    def ready(self):
        return self.value is not None or _lib.load().gsm_status_query(self.slot) == 1

    def resolve(self):
        if self.value is not None:
            return self.value
        R, ov, pf = ctypes.c_longlong(0), ctypes.c_int(0), ctypes.c_int(0)
        _lib.check(_lib.load().gsm_status_wait(self.slot, ctypes.byref(R), ctypes.byref(ov), ctypes.byref(pf)), "status_wait")
        _slots.give(self.slot)
        self.slot = None
        self.value, self.overflow = int(R.value), bool(ov.value)
        lst = _pending.get(self.key[0])
        if lst and self in lst:
            lst.remove(self)
        # grow-only high-water mark: the workspace keeps the policy's headroom over the largest R seen
        want = int(self.value * _HEADROOM[_sync_policy]) + 1024
        if want > _binning_capacity.get(self.key, 0):
            _binning_capacity[self.key] = want
        if pf:
            raise RuntimeError("gaussmart_b200 forward failed: Point is filtered although prefiltered is set. "
                               "This shouldn't happen!")
        return self.value

    def check(self):
        """resolve() + raise if the frame overflowed the workspace (it was rendered empty)."""
        self.resolve()
        if self.overflow:
            raise BinningOverflow(
                f"gaussmart_b200: a forward issued under the 'lazy' sync policy produced {self.value} (surfel, tile) "
                f"instances for a binning workspace of {self.capacity}: that frame was rendered EMPTY. The workspace "
                "has been enlarged; render the frame again, or use set_sync_policy('event').")
        return self.value

    def __int__(self):
        return self.check()

    __index__ = __int__

    def __eq__(self, other):
        return int(self) == other

    def __hash__(self):
        return id(self)

    def __repr__(self):
        return f"PendingForward({self.value if self.value is not None else 'pending'})"


def _drain_pending(dev_index, block=False):
    """Resolve the device's outstanding lazy forwards that have passed their scan (all of them with block=True)."""
    for pf in list(_pending.get(dev_index, ())):
        if block or pf.ready():
            pf.check()


class _Workspaces(NamedTuple):
    geometry: torch.Tensor
    image: torch.Tensor
    binning: torch.Tensor
    capacity: int


def _forward_impl(bg, means3D, colors_precomp, opacities, scales, rotations, scale_modifier, cov3Ds_precomp,
                  viewmatrix, projmatrix, tanfovx, tanfovy, image_height, image_width, sh, sh_degree, campos,
                  prefiltered, debug, work_counters=None, forward_only=False, sh_rest=None, raw_params=False):
    """Equivalent of RasterizeGaussiansCUDA (rasterize_points.cu:39-134).
    work_counters: optional zeroed int64[4] CUDA tensor receiving work counts (bench only).
    forward_only: inference render (no backward will follow): sets debug bit 2 of the C ABI.
    sh_rest / raw_params: the raw-parameter entry (rasterize_gaussians_raw): sh = features_dc [P,1,3] with
    sh_rest = features_rest [P,M-1,3]; scales / rotations / opacities as stored, activated in-kernel."""
    lib = _lib.load()
    if means3D.ndim != 2 or means3D.shape[1] != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")
    for name, t in (("background", bg), ("means3D", means3D), ("colors", colors_precomp), ("opacity", opacities),
                    ("scales", scales), ("rotations", rotations), ("transMat_precomp", cov3Ds_precomp),
                    ("viewmatrix", viewmatrix), ("projmatrix", projmatrix), ("sh", sh), ("campos", campos)):
        _check_cuda(name, t)
    P = means3D.shape[0]
    H, W = int(image_height), int(image_width)
    dev = means3D.device
    bg = _f32c("background", bg)
    means3D = _f32c("means3D", means3D)
    colors_precomp = _f32c("colors", colors_precomp)
    opacities = _f32c("opacity", opacities)
    scales = _f32c("scales", scales)
    rotations = _f32c("rotations", rotations)
    cov3Ds_precomp = _f32c("transMat_precomp", cov3Ds_precomp)
    viewmatrix = _f32c("viewmatrix", viewmatrix)
    projmatrix = _f32c("projmatrix", projmatrix)
    sh = _f32c("sh", sh)
    campos = _f32c("campos", campos)
    M = sh.shape[1] if sh.numel() != 0 else 0
    if sh_rest is not None:
        _check_cuda("sh_rest", sh_rest)
        sh_rest = _f32c("sh_rest", sh_rest)
        if M != 1 or sh_rest.shape[0] != P:
            raise RuntimeError("split SH storage needs sh = features_dc [P,1,3] and sh_rest = features_rest [P,M-1,3]")
        M = 1 + sh_rest.shape[1]

    with torch.cuda.device(dev):
        out_color = torch.empty((3, H, W), dtype=torch.float32, device=dev)
        out_others = torch.empty((7, H, W), dtype=torch.float32, device=dev)
        radii = torch.empty((P,), dtype=torch.int32, device=dev)
        geometry = torch.empty((lib.gsm_geometry_bytes(P),), dtype=torch.uint8, device=dev)
        image = torch.empty((lib.gsm_image_bytes(W, H),), dtype=torch.uint8, device=dev)
        stream = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        num_rendered = ctypes.c_int(0)

        a = _lib.ForwardArgs()
        a.P, a.D, a.M, a.width, a.height = P, int(sh_degree), M, W, H
        a.tan_fovx, a.tan_fovy, a.scale_modifier = float(tanfovx), float(tanfovy), float(scale_modifier)
        a.prefiltered, a.debug = int(bool(prefiltered)), int(debug) | (4 if forward_only else 0)
        a.background, a.means3D, a.shs, a.colors_precomp = _ptr(bg), _ptr(means3D), _ptr(sh), _ptr(colors_precomp)
        a.shs_rest, a.raw_params = (_ptr(sh_rest) if sh_rest is not None and sh_rest.numel() else None), int(bool(raw_params))
        a.opacities, a.scales, a.rotations = _ptr(opacities), _ptr(scales), _ptr(rotations)
        a.transMat_precomp = _ptr(cov3Ds_precomp)
        a.viewmatrix, a.projmatrix, a.campos = _ptr(viewmatrix), _ptr(projmatrix), _ptr(campos)
        a.out_color, a.out_others, a.radii = _ptr(out_color), _ptr(out_others), _ptr(radii)
        a.geometry, a.geometry_bytes = _ptr(geometry), geometry.numel()
        a.image, a.image_bytes = _ptr(image), image.numel()
        a.num_rendered_host = ctypes.pointer(num_rendered)
        a.work_counters = _ptr(work_counters)

        key = (dev.index, )
        policy = _sync_policy
        if policy == "lazy":
            _drain_pending(dev.index)
        capacity = _binning_capacity.get(key, 0)
        pending = None
        while True:
            binning = torch.empty((lib.gsm_binning_bytes(capacity, W, H),), dtype=torch.uint8, device=dev)
            a.binning, a.binning_bytes, a.binning_capacity = _ptr(binning), binning.numel(), capacity
            if policy == "sync" or capacity == 0 or P == 0:
                # (first forward on a device: nothing to size from yet)
                a.status = None
                code = lib.gsm_forward(ctypes.byref(a), stream)
                if code == _lib.GSM_ERR_BINNING_OVERFLOW:
                    capacity = int(num_rendered.value * _HEADROOM[policy]) + 1024
                    _binning_capacity[key] = capacity
                    continue
                _lib.check(code, "forward")
                result = num_rendered.value
                break
            slot = _slots.take()
            a.status = slot
            code = lib.gsm_forward(ctypes.byref(a), stream)
            if code != _lib.GSM_OK:
                _slots.give(slot)
                _lib.check(code, "forward")
            pending = PendingForward(slot, key, capacity, bool(prefiltered))
            if policy == "lazy":
                _pending.setdefault(dev.index, []).append(pending)
                result = pending
                break
            result = pending.resolve()       # "event": behind the scan, with the rest of the forward already queued
            if pending.overflow:             # rendered empty: again, with the enlarged workspace
                capacity = _binning_capacity[key]
                continue
            break
    # the capacity the forward ran with travels with the buffer (backward needs it to find the strip masks)
    binning.gsm_capacity = capacity
    binning.gsm_handover = not forward_only
    return result, out_color, out_others, radii, geometry, binning, image


def _backward_impl(bg, means3D, radii, colors_precomp, scales, rotations, scale_modifier, cov3Ds_precomp, viewmatrix,
                   projmatrix, tanfovx, tanfovy, grad_out_color, grad_depth, sh, sh_degree, campos, geomBuffer,
                   num_rendered, binningBuffer, imgBuffer, debug, out=None, accumulate=False, binning_capacity=None,
                   rgb_out=None, stats=None, sh_rest=None, raw_opacities=None, rgb_peers=None):
    """Equivalent of RasterizeGaussiansBackwardCUDA (rasterize_points.cu:136-233).

    sh_rest / raw_opacities: the raw-parameter entry (must match the forward): with sh_rest the SH gradient comes back as
    the pair (dL_dfeatures_dc [P,1,3], dL_dfeatures_rest [P,M-1,3]) in place of dL_dsh; with raw_opacities the gradients
    of scales / rotations / opacities are w.r.t. the RAW parameters.

    `out` (optional) maps means3D / shs / opacities / scales / rotations to preallocated tensors the
    kernel writes into (or adds to, with `accumulate`) -- the view-parallel gradient bucket.
    `stats` (optional) = (grad_norm[P], visible[P], max_radii[P]): densification statistics accumulated by the kernel.
    `rgb_out` (optional, [P,3]) selects the factored SH gradient: the kernel writes dL/dRGB there and no
    dL_dsh rows (gsm_backward_args.dL_drgb); `sh_grad_expand` rebuilds them.
    `rgb_peers` (optional, list of raw device addresses, one per rank of the NVLink domain): the factored SH gradient
    PUSHED by the kernel into every rank's receive slot over peer memory (gsm_backward_args.dL_drgb_peers)."""
    lib = _lib.load()
    if isinstance(num_rendered, PendingForward):
        num_rendered.check()   # lazy policy: the forward's status is looked at here (its scan finished long ago)
    if getattr(binningBuffer, "gsm_handover", True) is False:
        raise RuntimeError("gaussmart_b200 backward: the forward was an inference render (forward_only: no "
                           "forward->backward hand-over was written); its buffers cannot be back-propagated")
    P = means3D.shape[0]
    H, W = grad_out_color.shape[1], grad_out_color.shape[2]
    dev = means3D.device
    M = sh.shape[1] if sh.numel() != 0 else 0
    split = sh_rest is not None and sh_rest.numel() != 0
    if split:
        M = 1 + sh_rest.shape[1]
        if rgb_out is not None or rgb_peers:
            raise RuntimeError("split SH storage and the factored SH gradient are not combined")
    grad_out_color = _f32c("dL_dout_color", grad_out_color)
    grad_depth = _f32c("dL_dout_others", grad_depth)
    # autograd saves the tensors the CALLER passed: a strided view must be packed here as the forward packed it
    # (rasterize_points.cu:112-130 calls .contiguous() on everything, forward and backward)
    bg, means3D, colors_precomp = _f32c("background", bg), _f32c("means3D", means3D), _f32c("colors", colors_precomp)
    scales, rotations, cov3Ds_precomp = _f32c("scales", scales), _f32c("rotations", rotations), _f32c("transMat_precomp", cov3Ds_precomp)
    viewmatrix, projmatrix, campos = _f32c("viewmatrix", viewmatrix), _f32c("projmatrix", projmatrix), _f32c("campos", campos)
    sh = _f32c("sh", sh)
    if split:
        sh_rest = _f32c("sh_rest", sh_rest)
    with torch.cuda.device(dev):
        f32 = dict(dtype=torch.float32, device=dev)
        out = out or {}

        def _out(name, shape):
            t = out.get(name)
            if t is None:
                if accumulate:
                    raise RuntimeError(f"accumulate=True needs a preallocated output for {name}")
                return torch.empty(shape, **f32)
            if tuple(t.shape) != tuple(shape) or t.dtype != torch.float32 or not t.is_contiguous() or t.device != dev:
                raise RuntimeError(f"output tensor for {name} must be contiguous float32 {tuple(shape)} on {dev}")
            return t

        # gradients of inputs that were not given are neither allocated nor written (NULL selects that
        # in the kernel); autograd returns None for them, as the reference's front end does (dsr.py:144-154)
        dL_dmeans3D = _out("means3D", (P, 3))
        dL_dmeans2D = torch.empty((P, 3), **f32)
        dL_dcolors = _out("colors_precomp", (P, 3)) if colors_precomp.numel() != 0 else None
        dL_dopacity = _out("opacities", (P, 1))
        dL_dtransMat = _out("cov3D_precomp", (P, 9)) if cov3Ds_precomp.numel() != 0 else None
        if rgb_out is not None and (M == 0 or tuple(rgb_out.shape) != (P, 3) or rgb_out.dtype != torch.float32
                                    or not rgb_out.is_contiguous() or rgb_out.device != dev):
            raise RuntimeError("rgb_out must be a contiguous float32 [P,3] tensor on the surfels' device (and shs given)")
        dL_dsh = _out("shs", (P, 1 if split else M, 3)) if (M != 0 and rgb_out is None and not rgb_peers) else None
        dL_dsh_rest = _out("shs_rest", (P, M - 1, 3)) if split else None
        dL_dscales = _out("scales", (P, 2)) if scales.numel() != 0 else None
        dL_drotations = _out("rotations", (P, 4)) if rotations.numel() != 0 else None
        if P != 0:
            gws = torch.empty((lib.gsm_grad_workspace_bytes(P),), dtype=torch.uint8, device=dev)
            a = _lib.BackwardArgs()
            a.P, a.D, a.M, a.width, a.height = P, int(sh_degree), M, W, H
            a.tan_fovx, a.tan_fovy, a.scale_modifier, a.debug = float(tanfovx), float(tanfovy), float(scale_modifier), int(debug)
            a.background, a.means3D, a.shs, a.colors_precomp = _ptr(bg), _ptr(means3D), _ptr(sh), _ptr(colors_precomp)
            a.scales, a.rotations, a.transMat_precomp = _ptr(scales), _ptr(rotations), _ptr(cov3Ds_precomp)
            a.viewmatrix, a.projmatrix, a.campos, a.radii = _ptr(viewmatrix), _ptr(projmatrix), _ptr(campos), _ptr(radii)
            a.geometry, a.image, a.binning = _ptr(geomBuffer), _ptr(imgBuffer), _ptr(binningBuffer)
            if binning_capacity is None:
                binning_capacity = getattr(binningBuffer, "gsm_capacity", None)
            if binning_capacity is None:
                raise RuntimeError("backward needs the binning capacity of the forward call")
            a.binning_capacity = int(binning_capacity)
            a.dL_dout_color, a.dL_dout_others = _ptr(grad_out_color), _ptr(grad_depth)
            a.grad_workspace, a.grad_workspace_bytes = _ptr(gws), gws.numel()
            a.dL_dmeans2D, a.dL_dcolors, a.dL_dopacity = _ptr(dL_dmeans2D), _ptr(dL_dcolors), _ptr(dL_dopacity)
            a.dL_dmeans3D, a.dL_dtransMat, a.dL_dsh = _ptr(dL_dmeans3D), _ptr(dL_dtransMat), _ptr(dL_dsh)
            a.dL_dscales, a.dL_drotations = _ptr(dL_dscales), _ptr(dL_drotations)
            a.accumulate = int(bool(accumulate))
            a.dL_drgb = _ptr(rgb_out)
            if rgb_peers:
                if M == 0 or len(rgb_peers) > 8:
                    raise RuntimeError("rgb_peers needs shs and at most 8 ranks")
                a.n_rgb_peers = len(rgb_peers)
                for k, addr in enumerate(rgb_peers):
                    a.dL_drgb_peers[k] = int(addr)
            if split:
                a.shs_rest, a.dL_dsh_rest = _ptr(sh_rest), _ptr(dL_dsh_rest)
            if raw_opacities is not None:
                a.opacities, a.raw_params = _ptr(_f32c("opacity", raw_opacities)), 1
            if stats is not None:  # (grad_norm[P] f32, visible[P] f32, max_radii[P] i32): fused densification statistics
                gn, vis, mr = stats
                for t, dt in ((gn, torch.float32), (vis, torch.float32), (mr, torch.int32)):
                    if tuple(t.shape) != (P,) or t.dtype != dt or not t.is_contiguous() or t.device != dev:
                        raise RuntimeError("stats must be contiguous [P] tensors (float32, float32, int32) on the surfels' device")
                a.densify_grad_norm, a.densify_visible, a.max_radii = _ptr(gn), _ptr(vis), _ptr(mr)
            stream = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
            _lib.check(lib.gsm_backward(ctypes.byref(a), stream), "backward")
    if split:
        dL_dsh = (dL_dsh, dL_dsh_rest)
    return dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dtransMat, dL_dsh, dL_dscales, dL_drotations


def rasterize_gaussians(means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                        raster_settings):
    inputs = (means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp)
    if not (torch.is_grad_enabled() and any(isinstance(t, torch.Tensor) and t.requires_grad for t in inputs)):
        # inference (render.py, torch.no_grad): autograd would record nothing, so the forward runs without the
        # forward->backward hand-over
        rs = raster_settings
        _, color, depth, radii, _, _, _ = _forward_impl(
            rs.bg, means3D, colors_precomp, opacities, scales, rotations, rs.scale_modifier, cov3Ds_precomp, rs.viewmatrix,
            rs.projmatrix, rs.tanfovx, rs.tanfovy, rs.image_height, rs.image_width, sh, rs.sh_degree, rs.campos,
            rs.prefiltered, rs.debug, forward_only=True)
        return color, radii, depth
    return _RasterizeGaussians.apply(means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                                     cov3Ds_precomp, raster_settings)


def rasterize_gaussians_raw(xyz, means2D, features_dc, features_rest, opacity_raw, scaling_raw, rotation_raw,
                            raster_settings):
    """Raw-parameter entry: what render() computes as
        rasterizer(means3D=pc.get_xyz, means2D=..., shs=pc.get_features, opacities=pc.get_opacity,
                   scales=pc.get_scaling, rotations=pc.get_rotation)             (gaussian_renderer/__init__.py:60-106)
    from the STORED parameters `_xyz, _features_dc, _features_rest, _opacity, _scaling, _rotation`: the getters of
    scene/gaussian_model.py:104-123 (exp, F.normalize, sigmoid, torch.cat) run inside the preprocess kernels, forward
    and backward, so no activated copy, no concatenated SH tensor and none of their autograd nodes exist.
    Returns (color, radii, allmap) like GaussianRasterizer.forward; gradients flow to the raw parameters."""
    return _RasterizeGaussiansRaw.apply(xyz, means2D, features_dc, features_rest, opacity_raw, scaling_raw, rotation_raw,
                                        raster_settings)


class _RasterizeGaussiansRaw(torch.autograd.Function):
    @staticmethod
    def forward(ctx, xyz, means2D, features_dc, features_rest, opacity_raw, scaling_raw, rotation_raw, rs):
        e = torch.empty(0, device=xyz.device)
        needs_grad = any(ctx.needs_input_grad)
        num_rendered, color, depth, radii, geomBuffer, binningBuffer, imgBuffer = _forward_impl(
            rs.bg, xyz, e, opacity_raw, scaling_raw, rotation_raw, rs.scale_modifier, e, rs.viewmatrix, rs.projmatrix,
            rs.tanfovx, rs.tanfovy, rs.image_height, rs.image_width, features_dc, rs.sh_degree, rs.campos, rs.prefiltered,
            rs.debug, forward_only=not needs_grad, sh_rest=features_rest, raw_params=True)
        ctx.raster_settings = rs
        ctx.num_rendered = num_rendered
        ctx.binning_capacity = binningBuffer.gsm_capacity
        ctx.save_for_backward(xyz, features_dc, features_rest, opacity_raw, scaling_raw, rotation_raw, radii, geomBuffer,
                              binningBuffer, imgBuffer)
        ctx.mark_non_differentiable(radii)
        return color, radii, depth

    @staticmethod
    def backward(ctx, grad_out_color, grad_radii, grad_depth):
        rs = ctx.raster_settings
        (xyz, features_dc, features_rest, opacity_raw, scaling_raw, rotation_raw, radii, geomBuffer, binningBuffer,
         imgBuffer) = ctx.saved_tensors
        H, W = rs.image_height, rs.image_width
        if grad_out_color is None:
            grad_out_color = torch.zeros((3, H, W), dtype=torch.float32, device=xyz.device)
        if grad_depth is None:
            grad_depth = torch.zeros((7, H, W), dtype=torch.float32, device=xyz.device)
        e = torch.empty(0, device=xyz.device)
        (dL_dmeans2D, _, dL_dopacity, dL_dmeans3D, _, dL_dsh, dL_dscales, dL_drotations) = _backward_impl(
            rs.bg, xyz, radii, e, scaling_raw, rotation_raw, rs.scale_modifier, e, rs.viewmatrix, rs.projmatrix, rs.tanfovx,
            rs.tanfovy, grad_out_color, grad_depth, features_dc, rs.sh_degree, rs.campos, geomBuffer, ctx.num_rendered,
            binningBuffer, imgBuffer, rs.debug, binning_capacity=ctx.binning_capacity, sh_rest=features_rest,
            raw_opacities=opacity_raw)
        if isinstance(dL_dsh, tuple):
            d_dc, d_rest = dL_dsh
        else:
            d_dc, d_rest = dL_dsh, torch.zeros_like(features_rest)
        return dL_dmeans3D, dL_dmeans2D, d_dc, d_rest, dL_dopacity.view_as(opacity_raw), dL_dscales, dL_drotations, None


class _RasterizeGaussians(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                raster_settings):
        args = (raster_settings.bg, means3D, colors_precomp, opacities, scales, rotations,
                raster_settings.scale_modifier, cov3Ds_precomp, raster_settings.viewmatrix,
                raster_settings.projmatrix, raster_settings.tanfovx, raster_settings.tanfovy,
                raster_settings.image_height, raster_settings.image_width, sh, raster_settings.sh_degree,
                raster_settings.campos, raster_settings.prefiltered, raster_settings.debug)
        if raster_settings.debug:
            cpu_args = cpu_deep_copy_tuple(args)  # copy them before they can be corrupted
            try:
                num_rendered, color, depth, radii, geomBuffer, binningBuffer, imgBuffer = _forward_impl(*args)
            except Exception as ex:
                torch.save(cpu_args, "snapshot_fw.dump")
                print("\nAn error occured in forward. Please forward snapshot_fw.dump for debugging.")
                raise ex
        else:
            num_rendered, color, depth, radii, geomBuffer, binningBuffer, imgBuffer = _forward_impl(*args)

        ctx.raster_settings = raster_settings
        ctx.num_rendered = num_rendered
        ctx.opacity_shape = opacities.shape
        # backward needs the persistent prefix of the binning workspace (surfel list, block masks, per-entry pixel
        # masks: 37 of its 53 bytes per instance); the buffer is kept whole rather than copied
        ctx.binning_capacity = binningBuffer.gsm_capacity
        point_list = binningBuffer
        ctx.save_for_backward(colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, geomBuffer,
                              point_list, imgBuffer)
        ctx.mark_non_differentiable(radii)
        return color, radii, depth

    @staticmethod
    def backward(ctx, grad_out_color, grad_radii, grad_depth):
        num_rendered = ctx.num_rendered
        raster_settings = ctx.raster_settings
        (colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, geomBuffer, binningBuffer,
         imgBuffer) = ctx.saved_tensors
        H, W = raster_settings.image_height, raster_settings.image_width
        if grad_out_color is None:
            grad_out_color = torch.zeros((3, H, W), dtype=torch.float32, device=means3D.device)
        if grad_depth is None:
            grad_depth = torch.zeros((7, H, W), dtype=torch.float32, device=means3D.device)
        args = (raster_settings.bg, means3D, radii, colors_precomp, scales, rotations, raster_settings.scale_modifier,
                cov3Ds_precomp, raster_settings.viewmatrix, raster_settings.projmatrix, raster_settings.tanfovx,
                raster_settings.tanfovy, grad_out_color, grad_depth, sh, raster_settings.sh_degree,
                raster_settings.campos, geomBuffer, num_rendered, binningBuffer, imgBuffer, raster_settings.debug)
        kw = dict(binning_capacity=ctx.binning_capacity)
        if raster_settings.debug:
            cpu_args = cpu_deep_copy_tuple(args)
            try:
                out = _backward_impl(*args, **kw)
            except Exception as ex:
                torch.save(cpu_args, "snapshot_bw.dump")
                print("\nAn error occured in backward. Writing snapshot_bw.dump for debugging.\n")
                raise ex
        else:
            out = _backward_impl(*args, **kw)
        (grad_means2D, grad_colors_precomp, grad_opacities, grad_means3D, grad_cov3Ds_precomp, grad_sh, grad_scales,
         grad_rotations) = out
        grads = (
            grad_means3D,
            grad_means2D,
            grad_sh if sh.numel() != 0 else None,
            grad_colors_precomp if colors_precomp.numel() != 0 else None,
            grad_opacities.reshape(ctx.opacity_shape),
            grad_scales if scales.numel() != 0 else None,
            grad_rotations if rotations.numel() != 0 else None,
            grad_cov3Ds_precomp if cov3Ds_precomp.numel() != 0 else None,
            None,
        )
        return grads


def sh_grad_expand(means3D, campos_all, rgb_all, dL_dsh, sh_degree, accumulate=False, view_stride=None):
    """dL_dsh[P,M,3] (= or +=) sum over views v of basis(normalize(means3D - campos_all[v])) (x) rgb_all[v]
    (gsm_sh_grad_expand): the receiving side of the factored view-parallel exchange.
    view_stride: floats between consecutive views in `rgb_all` when it is a flat padded buffer (>= 3 P)."""
    lib = _lib.load()
    P, M = dL_dsh.shape[0], dL_dsh.shape[1]
    n_views = campos_all.shape[0]
    dev = dL_dsh.device
    for name, t, shape in (("means3D", means3D, (P, 3)), ("campos_all", campos_all, (n_views, 3)), ("dL_dsh", dL_dsh, (P, M, 3))):
        _check_cuda(name, t)
        if tuple(t.shape) != shape or t.dtype != torch.float32 or not t.is_contiguous():
            raise RuntimeError(f"{name} must be contiguous float32 {shape}")
    _check_cuda("rgb_all", rgb_all)
    if view_stride is None:
        view_stride = 3 * P
        if tuple(rgb_all.shape) != (n_views, P, 3):
            raise RuntimeError(f"rgb_all must be contiguous float32 {(n_views, P, 3)}")
    if rgb_all.dtype != torch.float32 or not rgb_all.is_contiguous() or view_stride < 3 * P or \
            rgb_all.numel() < (n_views - 1) * view_stride + 3 * P:
        raise RuntimeError("rgb_all must be contiguous float32 holding n_views slots of view_stride floats")
    with torch.cuda.device(dev):
        stream = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        _lib.check(lib.gsm_sh_grad_expand_strided(P, int(sh_degree), M, n_views, _ptr(means3D), _ptr(campos_all), _ptr(rgb_all),
                                                  int(view_stride), _ptr(dL_dsh), int(bool(accumulate)), stream), "sh_grad_expand")
    return dL_dsh


def rasterize_into_bucket(params, camera, bucket, accumulate, bg=None, scale_modifier=1.0, rgb_out=None,
                          sh_degree=None, fused_stats=False, rgb_peers=None):
    """Forward + (deferred) backward of one view with the surfel gradients written by the backward
    preprocess kernel straight into `bucket.views` (gaussmart_b200.view_parallel.GradBucket) -- no
    autograd, no pack kernel before the collective.

    params : dict means3D[P,3], shs[P,M,3], opacities[P,1], scales[P,2], rotations[P,4] (CUDA, fp32)
    camera : object with image_width/height, tanfovx/y, world_view_transform, full_proj_transform,
             camera_center (scene/cameras.py:17-59 attribute names), tensors on the same device
    Returns (color, radii, allmap, backward) with backward(dL_dcolor, dL_dallmap) -> dL_dmeans2D[P,3].
    """
    dev = params["means3D"].device
    e = torch.empty(0, device=dev)
    if bg is None:
        bg = torch.zeros(3, dtype=torch.float32, device=dev)
    sh = params["shs"]
    if sh_degree is None:  # active degree (gaussian_model.py:oneupSHdegree); default: all allocated coefficients
        sh_degree = int(round(sh.shape[1] ** 0.5)) - 1
    H, W = int(camera.image_height), int(camera.image_width)
    with torch.no_grad():
        R, color, allmap, radii, geom, binning, image = _forward_impl(
            bg, params["means3D"], e, params["opacities"], params["scales"], params["rotations"], scale_modifier, e,
            camera.world_view_transform, camera.full_proj_transform, camera.tanfovx, camera.tanfovy, H, W, sh,
            sh_degree, camera.camera_center, False, False)

    def backward(dL_dcolor, dL_dallmap):
        with torch.no_grad():
            g = _backward_impl(bg, params["means3D"].contiguous(), radii, e, params["scales"].contiguous(),
                               params["rotations"].contiguous(), scale_modifier, e,
                               camera.world_view_transform.contiguous(), camera.full_proj_transform.contiguous(),
                               camera.tanfovx, camera.tanfovy, dL_dcolor, dL_dallmap, sh.contiguous(), sh_degree,
                               camera.camera_center.contiguous(), geom, R, binning, image, False,
                               out=bucket.views, accumulate=accumulate, rgb_out=rgb_out, rgb_peers=rgb_peers,
                               stats=(bucket.stats[0], bucket.stats[1], bucket.max_radii) if fused_stats else None)
        return g[0]

    return color, radii, allmap, backward


class GaussianRasterizationSettings(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    debug: bool


def mark_visible(positions, viewmatrix, projmatrix):
    """Equivalent of markVisible (rasterize_points.cu:235-254): bool[P], view-space z > 0.2."""
    lib = _lib.load()
    _check_cuda("means3D", positions)
    P = positions.shape[0]
    dev = positions.device
    positions = _f32c("means3D", positions)
    viewmatrix = _f32c("viewmatrix", viewmatrix)
    projmatrix = _f32c("projmatrix", projmatrix)
    present = torch.zeros((P,), dtype=torch.bool, device=dev)
    if P != 0:
        with torch.cuda.device(dev):
            stream = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
            _lib.check(lib.gsm_mark_visible(P, _ptr(positions), _ptr(viewmatrix), _ptr(projmatrix), _ptr(present),
                                            stream), "mark_visible")
    return present


class GaussianRasterizer(nn.Module):
    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def markVisible(self, positions):
        # Mark visible points (based on frustum culling for camera) with a boolean
        with torch.no_grad():
            raster_settings = self.raster_settings
            visible = mark_visible(positions, raster_settings.viewmatrix, raster_settings.projmatrix)
        return visible

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
                cov3D_precomp=None):
        raster_settings = self.raster_settings

        if (shs is None and colors_precomp is None) or (shs is not None and colors_precomp is not None):
            raise Exception('Please provide excatly one of either SHs or precomputed colors!')

        if ((scales is None or rotations is None) and cov3D_precomp is None) or \
                ((scales is not None or rotations is not None) and cov3D_precomp is not None):
            raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')

        dev = means3D.device
        if shs is None:
            shs = torch.empty(0, device=dev)
        if colors_precomp is None:
            colors_precomp = torch.empty(0, device=dev)
        if scales is None:
            scales = torch.empty(0, device=dev)
        if rotations is None:
            rotations = torch.empty(0, device=dev)
        if cov3D_precomp is None:
            cov3D_precomp = torch.empty(0, device=dev)

        return rasterize_gaussians(means3D, means2D, shs, colors_precomp, opacities, scales, rotations,
                                   cov3D_precomp, raster_settings)
