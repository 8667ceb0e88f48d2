"""View-parallel multi-GPU step (new work -- the reference is single-GPU, its only
multi-GPU mechanism is whole-job fan-out, scripts/nerf_eval.py:26,57-65).

One process per GPU.  Surfel parameters are replicated; rank r rasterises cameras {r, r+N, ...} of
the step's batch; the per-surfel gradients of all local views are accumulated by the backward
preprocess kernel DIRECTLY into one flat fp32 bucket (no pack kernel), which is then sum-all-reduced
with NCCL over NVLink/NVSwitch.  Densification statistics follow train.py:199-203 / gm.py:551-553
semantics: the per-view norm of the view-space gradient is summed over views (reduce the norms, not
the raw gradients), visibility counts are summed, radii are max-reduced.

The data path has exactly one exchange step (the gradient all-reduce); rendering needs none.
"""
from typing import Callable, Dict, List, NamedTuple, Optional, Sequence
from torch._C._autograd import DeviceType

import os

import torch
import torch.distributed as dist

# order of the segments inside the flat bucket: the Adam parameter groups of gm.py:282-289
# (xyz 3 | f_dc+f_rest M*3 | opacity 1 | scaling 2 | rotation 4) = 58 floats/surfel at M = 16
SEGMENTS = (("means3D", lambda M: (3,)), ("shs", lambda M: (M, 3)), ("opacities", lambda M: (1,)),
            ("scales", lambda M: (2,)), ("rotations", lambda M: (4,)))


def shard_views(n_views: int, rank: int, world_size: int) -> List[int]:
    """Views of a step's batch owned by `rank`: {rank, rank + N, ...}."""
    if world_size <= 0 or not (0 <= rank < world_size):
        raise ValueError(f"bad rank/world_size {rank}/{world_size}")
    return list(range(rank, n_views, world_size))


def floats_per_surfel(M: int) -> int:
    return 3 + 3 * M + 1 + 2 + 4


class GradBucket:
    """Flat fp32 gradient bucket [P * floats_per_surfel(M)] with one contiguous segment per parameter
    tensor; `views[name]` are tensors aliasing the segments that the backward writes into."""

    def __init__(self, P: int, M: int, device, with_stats: bool = True):
        self.P, self.M = P, M
        # every segment starts on a 128-byte boundary (the kernels use 128-bit accesses)
        sizes = []
        for name, shape_of in SEGMENTS:
            shape = (P,) + shape_of(M)
            cnt = 1
            for s in shape:
                cnt *= s
            sizes.append((name, shape, cnt))
        n = sum((cnt + 31) // 32 * 32 for _, _, cnt in sizes)
        self.flat = torch.zeros(n, dtype=torch.float32, device=device)
        self.views: Dict[str, torch.Tensor] = {}
        off = 0
        for name, shape, cnt in sizes:
            self.views[name] = self.flat[off:off + cnt].view(shape)
            off += (cnt + 31) // 32 * 32
        assert off == n
        # densification statistics (gm.py:551-553): [0] sum of per-view ||dL_dmean2D.xy||, [1] visibility count
        self.stats = torch.zeros(2, P, dtype=torch.float32, device=device) if with_stats else None
        self.max_radii = torch.zeros(P, dtype=torch.int32, device=device) if with_stats else None

    def zero_(self):
        self.flat.zero_()
        if self.stats is not None:
            self.stats.zero_()
            self.max_radii.zero_()

    def nbytes(self) -> int:
        return self.flat.numel() * 4


def add_densification_stats(bucket: GradBucket, dL_dmeans2D: torch.Tensor, radii: torch.Tensor):
    """gm.py:551-553 + train.py:199-203 for one view: norm of the view-space gradient of visible surfels."""
    vis = radii > 0
    bucket.stats[0] += torch.where(vis, dL_dmeans2D[:, :2].norm(dim=-1), torch.zeros((), device=radii.device))
    bucket.stats[1] += vis.to(torch.float32)
    torch.maximum(bucket.max_radii, radii, out=bucket.max_radii)


def allreduce_bucket(bucket: GradBucket, group=None, async_op: bool = False):
    """The one exchange step of the path.  Returns the list of work handles when async_op."""
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return []
    works = [dist.all_reduce(bucket.flat, op=dist.ReduceOp.SUM, group=group, async_op=async_op)]
    if bucket.stats is not None:
        works.append(dist.all_reduce(bucket.stats, op=dist.ReduceOp.SUM, group=group, async_op=async_op))
        works.append(dist.all_reduce(bucket.max_radii, op=dist.ReduceOp.MAX, group=group, async_op=async_op))
    return [w for w in works if w is not None] if async_op else []


class ViewResult(NamedTuple):
    color: torch.Tensor
    radii: torch.Tensor
    allmap: torch.Tensor


def view_parallel_step(params: Dict[str, torch.Tensor], cameras: Sequence, bucket: GradBucket,
                       cotangent_fn: Callable, raster_fn: Optional[Callable] = None, group=None,
                       rank: Optional[int] = None, world_size: Optional[int] = None,
                       reduce: bool = True) -> List[ViewResult]:
    """One data-parallel step over `cameras` (the step's global batch).

    params       : dict means3D, shs, opacities, scales, rotations (replicated, identical on all ranks)
    cotangent_fn : (view_index, color, allmap) -> (dL_dcolor[3,H,W], dL_dallmap[7,H,W])
    raster_fn    : (params, camera, bucket, accumulate) -> (color, radii, allmap, backward(dc, da) -> dL_dmeans2D);
                   defaults to the CUDA rasterizer writing straight into the bucket.
    Returns this rank's ViewResults.  After the call bucket.flat holds the gradient summed over ALL views.
    """
    if rank is None:
        rank = dist.get_rank(group) if dist.is_initialized() else 0
    if world_size is None:
        world_size = dist.get_world_size(group) if dist.is_initialized() else 1
    if raster_fn is None:
        from .rasterizer import rasterize_into_bucket
        raster_fn = rasterize_into_bucket
    mine = shard_views(len(cameras), rank, world_size)
    results = []
    if not mine:
        bucket.zero_()  # a rank without a view still contributes zeros to the sum
    for j, v in enumerate(mine):
        color, radii, allmap, backward = raster_fn(params, cameras[v], bucket, j > 0)
        dc, da = cotangent_fn(v, color, allmap)
        dL_dmeans2D = backward(dc, da)
        if bucket.stats is not None:
            if j == 0:
                bucket.stats.zero_()
                bucket.max_radii.zero_()
            add_densification_stats(bucket, dL_dmeans2D, radii)
        results.append(ViewResult(color, radii, allmap))
    if reduce:
        allreduce_bucket(bucket, group)
    return results


# ------------------------------------------------------------------------------------------------
# Factored exchange: dL_dsh travels as dL_dRGB
# ------------------------------------------------------------------------------------------------
# dL_dsh of ONE view is the rank-1 product basis(normalize(mean - campos_view)) (x) dL_dRGB
# (backward.cu:20-139): 3M floats that are a function of 3 floats and of data every rank already holds
# (surfel positions are replicated, camera centres are 12 bytes per view).  So instead of all-reducing the
# 232 B/surfel bucket, ranks all-reduce the 40 B/surfel of the other gradients, all-gather dL_dRGB
# (12 B/surfel/view) and rebuild the SH gradient locally (gsm_sh_grad_expand): 4x fewer bytes on the wire at
# 8 views, same sum.
SMALL_SEGMENTS = (("means3D", (3,)), ("opacities", (1,)), ("scales", (2,)), ("rotations", (4,)))


class FactoredExchange:
    """Buffers of the factored exchange for P surfels with M SH coefficients on `world_size` ranks, each
    rasterising at most `max_local_views` views per step."""

    def __init__(self, P: int, M: int, device, world_size: int, max_local_views: int = 1, with_stats: bool = True):
        self.P, self.M, self.world_size, self.max_local_views = P, M, world_size, max_local_views
        sizes = [(name, (P,) + shape, P * shape[0]) for name, shape in SMALL_SEGMENTS]
        P_pad = (P + 31) // 32 * 32
        n = sum((cnt + 31) // 32 * 32 for _, _, cnt in sizes) + (2 * P_pad if with_stats else 0)
        self.flat = torch.zeros(n, dtype=torch.float32, device=device)   # the all-reduced part
        self.views: Dict[str, torch.Tensor] = {}
        off = 0
        for name, shape, cnt in sizes:
            self.views[name] = self.flat[off:off + cnt].view(shape)
            off += (cnt + 31) // 32 * 32
        self._stats_off = off  # the densification statistics ride in the same all-reduce (both are sums)
        self.rgb_send = torch.zeros(max_local_views, P, 3, dtype=torch.float32, device=device)
        self.rgb_recv = torch.zeros(world_size, max_local_views, P, 3, dtype=torch.float32, device=device)
        self.campos = torch.zeros(world_size * max_local_views, 3, dtype=torch.float32, device=device)
        self.dL_dsh = torch.zeros(P, M, 3, dtype=torch.float32, device=device)  # rebuilt locally, identical on all ranks
        self.stats = self.flat[off:off + 2 * P_pad].view(2, P_pad)[:, :P] if with_stats else None
        self.max_radii = torch.zeros(P, dtype=torch.int32, device=device) if with_stats else None

    def wire_bytes(self) -> int:
        """Bytes a rank contributes per step: all-reduced small bucket + its all-gathered dL_dRGB slices."""
        return self.flat.numel() * 4 + self.rgb_send.numel() * 4

    def gradients(self) -> Dict[str, torch.Tensor]:
        g = dict(self.views)
        g["shs"] = self.dL_dsh
        return g


def exchange_factored(ex: FactoredExchange, group=None, after_gather: Optional[Callable] = None):
    """The exchange step of the factored path: one all-gather (dL_dRGB) + one all-reduce (the rest, statistics
    included) + the max of the radii.  `after_gather()` (the local SH expand) is launched as soon as the gathered
    dL_dRGB is visible to the compute stream, so that it runs WHILE the all-reduce is still on the wire: the
    collectives are issued asynchronously and only waited for where their result is needed."""
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        ex.rgb_recv[0].copy_(ex.rgb_send)
        if after_gather is not None:
            after_gather()
        return
    if dist.get_backend(group) == "nccl":
        wg = dist.all_gather_into_tensor(ex.rgb_recv.view(-1), ex.rgb_send.view(-1), group=group, async_op=True)
    else:
        wg = dist.all_gather(list(ex.rgb_recv.unbind(0)), ex.rgb_send, group=group, async_op=True)
    wr = dist.all_reduce(ex.flat, op=dist.ReduceOp.SUM, group=group, async_op=True)   # gradients + statistics sums
    wm = dist.all_reduce(ex.max_radii, op=dist.ReduceOp.MAX, group=group, async_op=True) if ex.max_radii is not None else None
    dbg = os.environ.get("GSM_VP_PROBE", "")   # timing probes only (tools/): which part of the exchange is exposed
    wg.wait()
    if after_gather is not None and "skip_expand" not in dbg:
        after_gather()
    wr.wait()
    if wm is not None:
        wm.wait()


def view_parallel_step_factored(params: Dict[str, torch.Tensor], cameras: Sequence, ex: FactoredExchange,
                                cotangent_fn: Callable, raster_fn: Optional[Callable] = None,
                                expand_fn: Optional[Callable] = None, group=None, rank: Optional[int] = None,
                                world_size: Optional[int] = None, sh_degree: Optional[int] = None,
                                fused_stats: bool = False) -> List[ViewResult]:
    """view_parallel_step with the factored exchange.  After the call ex.gradients() holds, on every rank, the
    gradient of all five parameter tensors summed over ALL views of `cameras`.

    raster_fn : (params, camera, ex, accumulate, rgb_out=) -> (color, radii, allmap, backward) -- writes the small
                segments into ex.views and dL_dRGB of the view into rgb_out
    expand_fn : (means3D, campos_all[V,3], rgb_all[V,P,3], dL_dsh[P,M,3], sh_degree) -> None
    fused_stats : raster_fn's backward accumulates the densification statistics itself (the CUDA rasterizer does,
                with fused_stats=True: gsm_backward_args.densify_*); otherwise they are added here with torch ops
    sh_degree : ACTIVE SH degree the views were rasterised with (default: the degree the M allocated coefficients hold)
    """
    if rank is None:
        rank = dist.get_rank(group) if dist.is_initialized() else 0
    if world_size is None:
        world_size = dist.get_world_size(group) if dist.is_initialized() else 1
    if world_size != ex.world_size:
        raise ValueError(f"exchange buffers were sized for {ex.world_size} ranks, not {world_size}")
    if raster_fn is None:
        from .rasterizer import rasterize_into_bucket
        raster_fn = rasterize_into_bucket
    if expand_fn is None:
        from .rasterizer import sh_grad_expand
        expand_fn = sh_grad_expand
    n_views = len(cameras)
    if (n_views + world_size - 1) // world_size > ex.max_local_views:
        raise ValueError(f"{n_views} views on {world_size} ranks need max_local_views >= "
                         f"{(n_views + world_size - 1) // world_size}")
    mine = shard_views(n_views, rank, world_size)
    results = []
    if not mine:
        ex.flat.zero_()
    if ex.stats is not None:
        ex.flat[ex._stats_off:].zero_()
        ex.max_radii.zero_()
    for j, v in enumerate(mine):
        color, radii, allmap, backward = raster_fn(params, cameras[v], ex, j > 0, rgb_out=ex.rgb_send[j])
        dc, da = cotangent_fn(v, color, allmap)
        dL_dmeans2D = backward(dc, da)
        if ex.stats is not None and not fused_stats:
            add_densification_stats(ex, dL_dmeans2D, radii)
        results.append(ViewResult(color, radii, allmap))
    if len(mine) < ex.max_local_views:
        ex.rgb_send[len(mine):].zero_()  # unused slots contribute nothing
    # camera centre of slot (r, j) = view r + j*N; empty slots keep zeros (their dL_dRGB is zero)
    slots = ex.campos.view(world_size, ex.max_local_views, 3)
    for v in range(n_views):  # device-side copies of 12 bytes: no host synchronisation
        c = getattr(cameras[v], "camera_center", None)
        if c is not None:
            slots[v % world_size, v // world_size].copy_(c, non_blocking=True)
    if sh_degree is None:
        sh_degree = int(round(params["shs"].shape[1] ** 0.5)) - 1
    exchange_factored(ex, group, after_gather=lambda: expand_fn(params["means3D"], ex.campos,
                                                                ex.rgb_recv.view(-1, ex.P, 3), ex.dL_dsh, sh_degree))
    return results


# ------------------------------------------------------------------------------------------------
# Peer exchange: the factored exchange without NCCL launches, over NVLink peer memory
# ------------------------------------------------------------------------------------------------
# Round 1 ran the factored exchange as three NCCL collectives issued after the backward preprocess kernel (K6) had
# returned: all-gather of dL_dRGB, all-reduce SUM of the other gradients + densification sums, all-reduce MAX of the
# radii -- fully exposed, 1.6 ms of a 6.0 ms step at 8 GPUs.  Here the buffers live in SYMMETRIC memory
# (torch.distributed._symmetric_memory: every rank's copy is mapped into every process, NVLink 5 / NVSwitch) and
#   * K6 itself pushes its dL_dRGB rows into every rank's receive slot while it runs (gsm_backward_args.dL_drgb_peers:
#     coalesced 128-bit peer stores) -- the all-gather costs no launch and hides inside the kernel;
#   * the rest is ONE hand-written two-shot kernel (gsm_peer_allreduce, csrc/exchange.cu): rank r reduces shard r of all
#     ranks' buffers (SUM for the gradient / statistics floats, MAX for the radii, in one pass) and writes it back to
#     all ranks; the local SH expand is issued next to it on a side stream;
#   * two signal-pad barriers bracket the pair.
# Measured on 8 B200s (C4, 260 MB reduced + 420 MB of dL/dRGB rows per rank and step; profiles/r02_exchange_probes.md):
# the pushes cost K6 +0.19 ms; the two-shot kernel takes 0.69 ms (NVLink-bound: NCCL's rate for this size) whatever
# its launch shape; while it runs the expand next to it makes almost no progress (a spin kernel in its place overlaps
# perfectly), so the pair costs 1.08 ms against 0.69 + 0.46 one after the other.  Two alternatives are kept behind
# GSM_VP_REDUCE and were NOT faster: "multimem" (in-switch NVLS reduction: 0.58 ms + 0.33 ms of posted writes landing in
# the closing barrier) and "push" (scatter into staging + local sum + broadcast, stores only: 0.25 + 0.22 ms at 2 GPUs
# against 0.39 ms).
def peer_exchange_available(group=None) -> bool:
    """True when torch's symmetric memory can serve this process group (NCCL backend, CUDA, one NVLink domain)."""
    try:
        import torch.distributed._symmetric_memory as symm_mem  # noqa: F401
    except Exception:
        return False
    return (dist.is_initialized() and dist.get_backend(group) == "nccl" and 1 < dist.get_world_size(group) <= 8
            and os.environ.get("GSM_VP_EXCHANGE", "peer") != "nccl")


class PeerExchange:
    """Buffers of the peer exchange (same public fields as FactoredExchange: views, stats, max_radii, dL_dsh, campos,
    gradients()).  Raises if symmetric memory cannot be set up: callers fall back to FactoredExchange."""

    def __init__(self, P: int, M: int, device, group=None, max_local_views: int = 1, with_stats: bool = True):
        import torch.distributed._symmetric_memory as symm_mem
        self.group = group if group is not None else dist.group.WORLD
        self.world_size, self.rank = dist.get_world_size(self.group), dist.get_rank(self.group)
        self.P, self.M, self.max_local_views = P, M, max_local_views
        W, L = self.world_size, max_local_views
        sizes = [(name, (P,) + shape, P * shape[0]) for name, shape in SMALL_SEGMENTS]
        P_pad = (P + 31) // 32 * 32
        self.n_sum = sum((cnt + 31) // 32 * 32 for _, _, cnt in sizes) + 2 * P_pad
        self.n_max = P_pad
        self.flat = symm_mem.empty(self.n_sum + self.n_max, dtype=torch.float32, device=device)
        self.flat_hdl = symm_mem.rendezvous(self.flat, self.group)
        self.flat.zero_()
        self.views: Dict[str, torch.Tensor] = {}
        off = 0
        for name, shape, cnt in sizes:
            self.views[name] = self.flat[off:off + cnt].view(shape)
            off += (cnt + 31) // 32 * 32
        self._stats_off = off
        self.stats = self.flat[off:off + 2 * P_pad].view(2, P_pad)[:, :P] if with_stats else None
        self.max_radii = self.flat[self.n_sum:self.n_sum + P].view(torch.int32) if with_stats else None
        # staging area of the push-form all-reduce: this rank's shard of every rank's partial buffer
        from . import _lib
        self.stage = symm_mem.empty(int(_lib.load().gsm_peer_staging_floats(W, self.n_sum, self.n_max)), dtype=torch.float32, device=device)
        self.stage_hdl = symm_mem.rendezvous(self.stage, self.group)
        self._stage_ptrs = [int(p) for p in self.stage_hdl.buffer_ptrs]
        self.slot_stride = (3 * P + 31) // 32 * 32       # floats; every [P,3] slot starts 128-byte aligned
        self.rgb_recv = symm_mem.empty(W * L * self.slot_stride, dtype=torch.float32, device=device)
        self.rgb_hdl = symm_mem.rendezvous(self.rgb_recv, self.group)
        self.rgb_recv.zero_()
        self.campos = torch.zeros(W * L, 3, dtype=torch.float32, device=device)
        self.dL_dsh = torch.zeros(P, M, 3, dtype=torch.float32, device=device)
        self.side = torch.cuda.Stream(device=device)
        self._flat_ptrs = [int(p) for p in self.flat_hdl.buffer_ptrs]
        # NVSwitch multicast mapping of the reduced buffer (NVLS), when the fabric offers it
        self._flat_mc = 0
        try:
            if os.environ.get("GSM_VP_REDUCE", "peer") == "multimem" and self.flat_hdl.has_multicast_support(DeviceType.CUDA, device.index if hasattr(device, "index") else 0):
                self._flat_mc = int(self.flat_hdl.multicast_ptr or 0)
        except Exception:
            self._flat_mc = 0
        self._rgb_ptrs = [int(p) for p in self.rgb_hdl.buffer_ptrs]
        torch.cuda.synchronize(device)
        dist.barrier(self.group)                         # every rank's buffers are zeroed before anyone pushes

    def slot_index(self, rank: int, local_view: int) -> int:
        return rank * self.max_local_views + local_view

    def rgb_peer_addresses(self, local_view: int) -> List[int]:
        """Address of this rank's slot `local_view` in every rank's receive buffer."""
        off = self.slot_index(self.rank, local_view) * self.slot_stride * 4
        return [p + off for p in self._rgb_ptrs]

    def wire_bytes(self) -> int:
        return (self.n_sum + self.n_max) * 4 + self.max_local_views * self.P * 12

    def gradients(self) -> Dict[str, torch.Tensor]:
        g = dict(self.views)
        g["shs"] = self.dL_dsh
        return g


def view_parallel_step_peer(params: Dict[str, torch.Tensor], cameras: Sequence, ex: PeerExchange, cotangent_fn: Callable,
                            raster_fn: Optional[Callable] = None, expand_fn: Optional[Callable] = None,
                            sh_degree: Optional[int] = None, barrier_timeout_ms: int = 60_000) -> List[ViewResult]:
    """view_parallel_step_factored over peer memory.  After the call ex.gradients(), ex.stats and ex.max_radii hold, on
    every rank, the sums (max for the radii) over ALL views of `cameras`, bitwise identical on all ranks.

    raster_fn : (params, camera, ex, accumulate, rgb_peers=) -> (color, radii, allmap, backward); the backward must
                accumulate the densification statistics itself (rasterize_into_bucket(..., fused_stats=True))."""
    import ctypes
    from . import _lib
    if raster_fn is None:
        from .rasterizer import rasterize_into_bucket
        raster_fn = lambda p_, c_, b_, a_, rgb_peers=None: rasterize_into_bucket(p_, c_, b_, a_, rgb_peers=rgb_peers,  # noqa: E731
                                                                                 fused_stats=True)
    if expand_fn is None:
        from .rasterizer import sh_grad_expand
        expand_fn = sh_grad_expand
    W, rank, L = ex.world_size, ex.rank, ex.max_local_views
    n_views = len(cameras)
    if (n_views + W - 1) // W > L:
        raise ValueError(f"{n_views} views on {W} ranks need max_local_views >= {(n_views + W - 1) // W}")
    mine = shard_views(n_views, rank, W)
    results = []
    # this rank's partial sums start from zero (K6 overwrites the gradient segments with the first view and adds the
    # statistics); nobody else touches this buffer until the first barrier below
    if not mine:
        ex.flat.zero_()
    else:
        ex.flat[ex._stats_off:].zero_()
    for j, v in enumerate(mine):
        color, radii, allmap, backward = raster_fn(params, cameras[v], ex, j > 0, rgb_peers=ex.rgb_peer_addresses(j))
        dc, da = cotangent_fn(v, color, allmap)
        backward(dc, da)
        results.append(ViewResult(color, radii, allmap))
    for j in range(len(mine), L):  # unused slots of this rank contribute nothing: zeros pushed to every rank
        for k in range(W):
            ex.rgb_hdl.get_buffer(k, (ex.slot_stride,), torch.float32, ex.slot_index(rank, j) * ex.slot_stride).zero_()
    slots = ex.campos.view(W, L, 3)
    for v in range(n_views):
        c = getattr(cameras[v], "camera_center", None)
        if c is not None:
            slots[v % W, v // W].copy_(c, non_blocking=True)
    if sh_degree is None:
        sh_degree = int(round(params["shs"].shape[1] ** 0.5)) - 1
    dev = ex.flat.device
    main = torch.cuda.current_stream(dev)
    probe = getattr(ex, "probe", None)   # optional timing probe (bench / tools): list of (name, start event, end event)

    def mark(stream):
        if probe is None:
            return None
        e = torch.cuda.Event(enable_timing=True)
        e.record(stream)
        return e
    t0 = mark(main)
    lib = _lib.load()
    ptrs = (ctypes.c_void_p * W)(*ex._flat_ptrs)
    push = os.environ.get("GSM_VP_REDUCE", "peer") == "push" and not ex._flat_mc
    if push:   # shard s of this rank's partial sums -> rank s's staging area (posted stores)
        sptrs = (ctypes.c_void_p * W)(*ex._stage_ptrs)
        _lib.check(lib.gsm_peer_scatter(ctypes.c_void_p(ex.flat.data_ptr()), sptrs, W, rank, ex.n_sum, ex.n_max,
                                        ctypes.c_void_p(main.cuda_stream)), "peer_scatter")
    # barrier 1: every rank's K6 (and scatter) has finished -> all receive slots and all staged shards are complete
    ex.flat_hdl.barrier(channel=0, timeout_ms=barrier_timeout_ms)
    t1 = mark(main)

    def reduce_on(stream):
        if os.environ.get("GSM_VP_PROBE", "") == "sleep":   # timing probe: a one-CTA spin kernel in place of the reduction
            with torch.cuda.stream(stream):
                torch.cuda._sleep(int(0.4e-3 * 1.9e9))
            return
        if push:          # local sum of the staged shards, result pushed to every rank
            _lib.check(lib.gsm_peer_sum_broadcast(ctypes.c_void_p(ex.stage.data_ptr()), ptrs, W, rank, ex.n_sum, ex.n_max,
                                                  ctypes.c_void_p(stream.cuda_stream)), "peer_sum_broadcast")
        elif ex._flat_mc:   # in-switch reduction (NVLS)
            _lib.check(lib.gsm_peer_allreduce_multimem(ctypes.c_void_p(ex._flat_mc), W, rank, ex.n_sum, ex.n_max,
                                                       ctypes.c_void_p(stream.cuda_stream)), "peer_allreduce_multimem")
        else:
            _lib.check(lib.gsm_peer_allreduce(ptrs, W, rank, ex.n_sum, ex.n_max, ctypes.c_void_p(stream.cuda_stream)), "peer_allreduce")
    ov = os.environ.get("GSM_VP_OVERLAP", "1")
    if ov == "1":
        ex.side.wait_stream(main)
        with torch.cuda.stream(ex.side):      # the local SH expand runs next to the two-shot reduction
            expand_fn(params["means3D"], ex.campos, ex.rgb_recv, ex.dL_dsh, sh_degree, view_stride=ex.slot_stride)
            t_side = mark(ex.side)
        reduce_on(main)
        t2 = mark(main)
        main.wait_stream(ex.side)
    elif ov == "2":   # the reduction on the side stream, issued first; the expand on the main stream
        ex.side.wait_stream(main)
        with torch.cuda.stream(ex.side):
            reduce_on(ex.side)
            t2 = mark(ex.side)
        expand_fn(params["means3D"], ex.campos, ex.rgb_recv, ex.dL_dsh, sh_degree, view_stride=ex.slot_stride)
        t_side = mark(main)
        main.wait_stream(ex.side)
    else:  # one after the other on the main stream (the remote shards first: peers wait for them, not for the expand)
        reduce_on(main)
        t2 = mark(main)
        expand_fn(params["means3D"], ex.campos, ex.rgb_recv, ex.dL_dsh, sh_degree, view_stride=ex.slot_stride)
        t_side = mark(main)
    t3 = mark(main)
    # barrier 2: every shard has been written everywhere, every rank is done reading its receive slots
    ex.flat_hdl.barrier(channel=1, timeout_ms=barrier_timeout_ms)
    t4 = mark(main)
    if probe is not None:
        probe.append((("barrier1", t0, t1), ("allreduce", t1, t2), ("expand_end_after_barrier1", t1, t_side),
                      ("join", t2, t3), ("barrier2", t3, t4), ("exchange_total", t0, t4)))
    return results
