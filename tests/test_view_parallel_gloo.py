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
        # numpy: pickled by value (tensors travel as shared-memory handles that die with this process)
        q.put((rank, [int(r.color[0, 0, 0].item()) for r in res],
               torch.cat([v.reshape(-1) for v in bucket.views.values()]).numpy().copy(), bucket.stats.numpy().copy(),
               bucket.max_radii.numpy().copy()))
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
        flat, stats, max_radii = torch.from_numpy(flat), torch.from_numpy(stats), torch.from_numpy(max_radii)
        assert torch.allclose(flat, torch.full_like(flat, expect))
        # densification stats: norms are reduced per view (5 (v+1) where visible), visibility counted
        idx = torch.arange(P)
        vis = torch.stack([((idx + v) % 3) > 0 for v in range(n_views)]).float()
        norms = torch.tensor([5.0 * (v + 1) for v in range(n_views)])[:, None] * vis
        assert torch.allclose(stats[0], norms.sum(0)) and torch.allclose(stats[1], vis.sum(0))
        assert torch.equal(max_radii, torch.stack([((idx + v) % 3) for v in range(n_views)]).max(0).values.int())
    assert (out[0][2] == out[1][2]).all()


# ---- factored exchange (dL_dsh travels as dL_dRGB): host logic on gloo, stub rasterizer + torch expand ----
class _Cam:
    def __init__(self, v):
        self.v = v
        self.camera_center = torch.tensor([float(v), 2.0 * v, -1.0 - v])


def _stub_raster_factored(params, camera, ex, accumulate, rgb_out=None):
    v = camera.v
    P = params["means3D"].shape[0]
    radii = ((torch.arange(P) + v) % 3).to(torch.int32)

    def backward(dc, da):
        for name, t in ex.views.items():
            g = torch.full_like(t, float(v + 1))
            if accumulate:
                t += g
            else:
                t.copy_(g)
        rgb_out.copy_(torch.arange(P, dtype=torch.float32)[:, None] * torch.tensor([1.0, 10.0, 100.0]) * (v + 1))
        d2 = torch.zeros(P, 3)
        d2[:, 0] = 3.0 * (v + 1)
        d2[:, 1] = 4.0 * (v + 1)
        return d2
    return torch.full((3, 4, 4), float(v)), radii, torch.zeros(7, 4, 4), backward


def _torch_expand(means3D, campos_all, rgb_all, dL_dsh, sh_degree, accumulate=False):
    """basis stand-in: coefficient k of view slot s is (k + 1) * (1 + campos[s].x)"""
    M = dL_dsh.shape[1]
    k = torch.arange(1, M + 1, dtype=torch.float32)
    out = torch.einsum("s,k,spc->pkc", 1.0 + campos_all[:, 0], k, rgb_all)
    dL_dsh.copy_(out)


def _worker_factored(rank, world, port, n_views, P, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        M = 4
        params = dict(means3D=torch.zeros(P, 3), shs=torch.zeros(P, M, 3), opacities=torch.zeros(P, 1),
                      scales=torch.zeros(P, 2), rotations=torch.zeros(P, 4))
        ex = vp.FactoredExchange(P, M, "cpu", world, max_local_views=(n_views + world - 1) // world)
        ex.flat.fill_(55.0)
        ex.rgb_send.fill_(7.0)   # stale content must not leak
        ex.dL_dsh.fill_(9.0)
        res = vp.view_parallel_step_factored(params, [_Cam(v) for v in range(n_views)], ex,
                                             cotangent_fn=lambda v, c, a: (torch.ones(3, 4, 4), torch.zeros(7, 4, 4)),
                                             raster_fn=_stub_raster_factored, expand_fn=_torch_expand)
        g = ex.gradients()
        # numpy: pickled by value (tensors travel as shared-memory handles that die with this process)
        q.put((rank, [int(r.color[0, 0, 0].item()) for r in res], {k: v.numpy().copy() for k, v in g.items()},
               ex.stats.numpy().copy(), ex.max_radii.numpy().copy()))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n_views", [2, 5, 1])
def test_view_parallel_step_factored_world2(n_views):
    world, P, M = 2, 6, 4
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker_factored, args=(r, world, port, n_views, P, q)) for r in range(world)]
    for p in procs:
        p.start()
    out = sorted((q.get(timeout=120) for _ in range(world)), key=lambda t: t[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert out[0][1] == list(range(0, n_views, 2)) and out[1][1] == list(range(1, n_views, 2))
    expect_small = float(sum(v + 1 for v in range(n_views)))
    base = torch.arange(P, dtype=torch.float32)[:, None] * torch.tensor([1.0, 10.0, 100.0])
    k = torch.arange(1, M + 1, dtype=torch.float32)
    expect_sh = sum((1.0 + v) * k[None, :, None] * (base * (v + 1))[:, None, :] for v in range(n_views))
    for _, _, g, stats, max_radii in out:
        g = {k: torch.from_numpy(v) for k, v in g.items()}
        stats = torch.from_numpy(stats)
        for name in ("means3D", "opacities", "scales", "rotations"):
            assert torch.allclose(g[name], torch.full_like(g[name], expect_small))
        assert torch.allclose(g["shs"], expect_sh)   # every rank rebuilt the SAME SH gradient from all views
        idx = torch.arange(P)
        vis = torch.stack([((idx + v) % 3) > 0 for v in range(n_views)]).float()
        assert torch.allclose(stats[1], vis.sum(0))


def test_factored_exchange_sizes():
    ex = vp.FactoredExchange(1000, 16, "cpu", 8, 1)
    assert ex.flat.numel() >= 10 * 1000 and ex.rgb_recv.shape == (8, 1, 1000, 3)
    dense = vp.GradBucket(1000, 16, "cpu").nbytes()
    assert ex.wire_bytes() * 3 < dense   # 60 B (statistics included) vs 232 B per surfel contributed per rank
    with pytest.raises(ValueError):
        vp.view_parallel_step_factored({"means3D": torch.zeros(1000, 3), "shs": torch.zeros(1000, 16, 3)},
                                       [_Cam(v) for v in range(9)], ex, None, raster_fn=_stub_raster_factored,
                                       expand_fn=_torch_expand, rank=0, world_size=8)
