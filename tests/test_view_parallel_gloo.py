"""CPU, world_size 2 over gloo: the host logic of the view-parallel step -- view sharding,
bucket layout, accumulation over a rank's local views, the gradient all-reduce and the densification
statistics -- with a stub standing in for the CUDA rasterizer (kernels cannot run here)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from gaussmart_b200 import view_parallel as vp


def test_shard_views_partition():
    for n_views in (1, 7, 8, 49):
        for world in (1, 2, 4, 8):
            got = sorted(v for r in range(world) for v in vp.shard_views(n_views, r, world))
            assert got == list(range(n_views))
    assert vp.shard_views(8, 3, 8) == [3]
    assert vp.shard_views(8, 1, 4) == [1, 5]
    with pytest.raises(ValueError):
        vp.shard_views(8, 4, 4)


def test_bucket_layout():
    P, M = 32, 16
    b = vp.GradBucket(P, M, "cpu")
    assert vp.floats_per_surfel(M) == 58 and b.flat.numel() == 58 * P  # 232 B/surfel
    assert all(v.data_ptr() % 128 == b.flat.data_ptr() % 128 for v in b.views.values())
    assert vp.GradBucket(33, M, "cpu").flat.numel() >= 58 * 33  # odd sizes: segments padded to 128 B
    assert b.views["shs"].shape == (P, 16, 3) and b.views["rotations"].shape == (P, 4)
    # segments alias the flat buffer, are disjoint and cover it
    for i, (name, _) in enumerate(vp.SEGMENTS):
        b.views[name].fill_(float(i + 1))
    assert b.flat.min().item() == 1.0 and b.flat.sum().item() == P * (3 * 1 + 48 * 2 + 1 * 3 + 2 * 4 + 4 * 5)
    assert all(v.is_contiguous() for v in b.views.values())


def _stub_raster(params, camera, bucket, accumulate):
    """Deterministic fake: 'renders' view v and produces gradients that depend on v only."""
    v = camera
    P = params["means3D"].shape[0]
    radii = ((torch.arange(P) + v) % 3).to(torch.int32)  # visible where > 0
    color = torch.full((3, 4, 4), float(v))
    allmap = torch.zeros(7, 4, 4)

    def backward(dc, da):
        scale = dc.sum().item()
        for name, t in bucket.views.items():
            g = torch.full_like(t, scale * (v + 1))
            if accumulate:
                t += g
            else:
                t.copy_(g)
        d2 = torch.zeros(P, 3)
        d2[:, 0] = 3.0 * (v + 1)
        d2[:, 1] = 4.0 * (v + 1)   # norm = 5 (v+1)
        return d2
    return color, radii, allmap, backward


def _worker(rank, world, port, n_views, P, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        params = dict(means3D=torch.zeros(P, 3), shs=torch.zeros(P, 4, 3), opacities=torch.zeros(P, 1),
                      scales=torch.zeros(P, 2), rotations=torch.zeros(P, 4))
        bucket = vp.GradBucket(P, 4, "cpu")
        bucket.flat.fill_(123.0)  # stale content must not leak into the result
        res = vp.view_parallel_step(params, list(range(n_views)), bucket,
                                    cotangent_fn=lambda v, c, a: (torch.ones(3, 4, 4) / 48.0, torch.zeros(7, 4, 4)),
                                    raster_fn=_stub_raster)
        q.put((rank, [int(r.color[0, 0, 0].item()) for r in res],
               torch.cat([v.reshape(-1) for v in bucket.views.values()]), bucket.stats.clone(),
               bucket.max_radii.clone()))
    finally:
        dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.parametrize("n_views", [2, 5, 1])
def test_view_parallel_step_world2(n_views):
    world, P = 2, 6
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_views, P, q)) for r in range(world)]
    for p in procs:
        p.start()
    out = sorted((q.get(timeout=120) for _ in range(world)), key=lambda t: t[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    # each rank rendered its own views
    assert out[0][1] == list(range(0, n_views, 2)) and out[1][1] == list(range(1, n_views, 2))
    # every rank holds the SAME bucket = sum over ALL views of that view's gradient (v+1)
    expect = float(sum(v + 1 for v in range(n_views)))
    for _, _, flat, stats, max_radii in out:
        assert torch.allclose(flat, torch.full_like(flat, expect))
        # densification stats: norms are reduced per view (5 (v+1) where visible), visibility counted
        idx = torch.arange(P)
        vis = torch.stack([((idx + v) % 3) > 0 for v in range(n_views)]).float()
        norms = torch.tensor([5.0 * (v + 1) for v in range(n_views)])[:, None] * vis
        assert torch.allclose(stats[0], norms.sum(0)) and torch.allclose(stats[1], vis.sum(0))
        assert torch.equal(max_radii, torch.stack([((idx + v) % 3) for v in range(n_views)]).max(0).values.int())
    assert torch.equal(out[0][2], out[1][2])
