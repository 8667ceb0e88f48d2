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
