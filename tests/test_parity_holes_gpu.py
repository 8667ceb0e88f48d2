"""GPU: parity cases round 1 left open (VERDICT r01 "What's weak" 1-2), all against the reference BUILD (oracle/_ref)
with a zero budget on every integer output:
  (a) BASELINE.json configs[2] and configs[4] at their REAL sizes: C3 = 2 M surfels at 1957x1091 (8,487 tiles, 46 sort
      bits), C5 = 6 M surfels at 1237x822;
  (b) one 3840x2160 forward of the 6 M scene (32,400 tiles, 47 key bits) incl. keys and ranges;
  (c) crowded tiles: segments of more than 4,096 and more than 9,216 instances, i.e. the 144 KB shared-memory class and
      the global-memory ping-pong class of tile_sort_kernel (csrc/binning.cu) against cub::DeviceRadixSort's output in
      the reference's binningBuffer (rasterizer_impl.cu:304-309);
  (d) fronto-parallel planar clouds: thousands of identical view depths per tile, so the whole (tile, depth) order is
      decided by the tie rule (ascending surfel index = the reference's stable sort of its emission order);
  (e) every small edge-case variant of test_parity_gpu.py::SMALL (SH degrees 0/1/2, ragged W x H, sh_cov, colors_cov,
      bg = 1, scale_modifier = 1.7) that round 1 only held to the CPU oracle with a mismatch budget.
"""
import pytest
import torch

from gaussmart_b200 import synthetic
from helpers import compare_with_reference_build
import ref_loader

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not ref_loader.available(), reason="oracle/_ref (reference CUDA build) not present")]


# ---------------------------------------------------------------------------------------------------- (a), (b)
@pytest.mark.parametrize("cfg,view", [("C3", 0), ("C5", 3)])
def test_full_size_configs_match_reference_build(cfg, view):
    scene, cams = synthetic.make_config(cfg)
    P, W, H, _, _ = synthetic.CONFIGS[cfg]
    assert scene.means3D.shape[0] == P and (cams[0].image_width, cams[0].image_height) == (W, H)
    _, _, im = compare_with_reference_build(scene, cams[view % len(cams)], "sh", 0.0, cfg)
    assert im["num_rendered"] > P                                   # a real workload, not an empty frame
    torch.cuda.empty_cache()


def test_4k_forward_matches_reference_build():
    scene, cams = synthetic.make_config("C5_4K")
    cam = cams[1]
    assert (cam.image_width, cam.image_height) == (3840, 2160)
    _, _, im = compare_with_reference_build(scene, cam, "sh", 0.0, "C5_4K", backward=False)
    assert im["ranges"].shape[0] == 32_400
    assert int(im["keys"].max().item()) >> 32 >= 1 << 14            # tile ids need 15 bits: 47 key bits live
    torch.cuda.empty_cache()


# ---------------------------------------------------------------------------------------------------- (c)
@pytest.mark.parametrize("P,lo,hi", [(14_000, 4096, 9216), (300_000, 9216, 1 << 30)])
def test_crowded_tiles_exercise_the_large_sort_classes(P, lo, hi):
    W = H = 64
    scene = synthetic.make_scene(P, W, 1, 60_000 + P)
    scene = scene._replace(scales=scene.scales * 3.0)
    cam = synthetic.make_camera(W, H, 0.2)
    _, _, im = compare_with_reference_build(scene, cam, "sh", 0.0, f"crowded{P}")
    n = (im["ranges"][:, 1] - im["ranges"][:, 0]).long()
    assert ((n > lo) & (n <= hi)).any(), n.tolist()                 # the class under test really ran


# ---------------------------------------------------------------------------------------------------- (d)
@pytest.mark.parametrize("P,W,H,levels", [(60_000, 128, 128, 1), (400_000, 96, 96, 5), (30_000, 160, 120, 3)])
def test_planar_cloud_orders_ties_by_surfel_index(P, W, H, levels):
    """Camera at theta = 0 has R = I, T = 0, so view depth == world z exactly: `levels` distinct depth keys in all."""
    g = torch.Generator().manual_seed(P + levels)
    scene = synthetic.make_scene(P, W, 0, 70_000 + P)
    m = scene.means3D.clone()
    m[:, 0] = (torch.rand(P, generator=g) - 0.5) * 3.0
    m[:, 1] = (torch.rand(P, generator=g) - 0.5) * 3.0
    m[:, 2] = 3.0 + 0.5 * torch.randint(0, levels, (P,), generator=g).float()
    scene = scene._replace(means3D=m.contiguous(), opacities=scene.opacities * 0.05)   # long lists before T < 1e-4
    cam = synthetic.make_camera(W, H, 0.0)
    _, _, im = compare_with_reference_build(scene, cam, "sh", 0.0, f"planar{P}")
    keys = im["keys"]
    depth_bits = torch.unique(keys & 0xffffffff)
    assert depth_bits.numel() == levels
    same = keys[1:] == keys[:-1]
    assert same.float().mean().item() > 0.99                        # the order is the tie rule's
    pl = im["point_list"].long()
    assert (pl[1:][same] > pl[:-1][same]).all()


# ---------------------------------------------------------------------------------------------------- (e)
SMALL = [  # name, P, W, H, sh_degree, variant, bg, scale_modifier
    ("deg0", 3000, 96, 80, 0, "sh", 0.0, 1.0),
    ("deg1", 3000, 96, 80, 1, "sh", 0.0, 1.0),
    ("deg2_bg1", 3000, 96, 80, 2, "sh", 1.0, 1.0),
    ("deg3_ragged", 3000, 101, 67, 3, "sh", 0.0, 1.0),
    ("colors_bg1", 3000, 96, 80, 0, "colors", 1.0, 1.0),
    ("sh_cov", 3000, 96, 80, 3, "sh_cov", 0.0, 1.0),
    ("colors_cov", 3000, 96, 80, 0, "colors_cov", 0.0, 1.0),
    ("modifier_1p7", 3000, 96, 80, 1, "sh", 0.0, 1.7),
    ("modifier_1p7_cov", 3000, 101, 67, 2, "sh_cov", 1.0, 1.7),
    ("one_pixel_row", 2000, 97, 1, 1, "sh", 0.0, 1.0),             # H = 1: a single ragged tile row
]


@pytest.mark.parametrize("name,P,W,H,deg,variant,bg,mod", SMALL, ids=[c[0] for c in SMALL])
def test_small_variants_match_reference_build(name, P, W, H, deg, variant, bg, mod):
    scene = synthetic.make_scene(P, W, deg, 1234 + P + deg)
    scene = scene._replace(scales=scene.scales * 3.0)   # lists long enough to hit the T < 1e-4 stop
    cam = synthetic.make_camera(W, H, 0.3)
    # scale_modifier != 1: the reference's backward ignores it (backward.cu:487 forces 1.0) and so does ours -- the
    # gradients of a modified render are compared as well
    compare_with_reference_build(scene, cam, variant, bg, name, scale_modifier=mod)


# ---------------------------------------------------------------------------------------------------- binning paths
def test_binning_path_for_more_than_2_24_surfels_equals_the_packed_one():
    """Up to 2^24 surfels the block masks travel through the tile sort in the top byte of the pair values; beyond, the
    sort gathers each instance's box and slabs itself.  debug bit 3 selects the second path at any P: identical bits."""
    import gaussmart_b200 as gb
    from helpers import run_cuda, variant_inputs
    import workspace_views
    W, H = 200, 150
    scene = synthetic.make_scene(40_000, W, 2, 808)
    scene = scene._replace(scales=scene.scales * 2.5)
    cam = synthetic.make_camera(W, H, 0.4)
    cot = synthetic.make_cotangents(W, H, 3)
    a = run_cuda(gb, scene, cam, 0.0, cot, "sh")
    b = run_cuda(gb, scene, cam, 0.0, cot, "sh", debug=8)
    assert torch.equal(a["radii"], b["radii"]) and torch.equal(a["color"], b["color"]) and torch.equal(a["allmap"], b["allmap"])
    kw = {k: v.cuda() for k, v in variant_inputs(scene, cam, "sh").items()}
    ia = workspace_views.forward_with_intermediates(a["settings"], scene.means3D.cuda(), scene.opacities.cuda(), **kw)
    ib = workspace_views.forward_with_intermediates(b["settings"], scene.means3D.cuda(), scene.opacities.cuda(), **kw)
    for k in ("keys", "point_list", "ranges", "n_contrib"):
        assert torch.equal(ia[k], ib[k]), k
    for k in ("grad_means3D", "grad_shs", "grad_scales", "grad_rotations", "grad_opacities"):
        d = (a[k] - b[k]).abs().max().item()
        assert d <= 1e-4 * a[k].abs().max().item() + 1e-12, (k, d)
