import sys, torch
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests')
from gaussmart_b200 import synthetic, _lib
from gaussmart_b200.rasterizer import _forward_impl
P, W, H, deg, _ = synthetic.CONFIGS["C4"]
scene = synthetic.make_scene(P, W, deg, synthetic.BASE_SEED + 4)
cam = synthetic.make_cameras(W, H, 8)[0]
d = lambda x: x.cuda() if isinstance(x, torch.Tensor) else x
e = torch.empty(0, device="cuda")
out = _forward_impl(torch.zeros(3, device="cuda"), d(scene.means3D), e, d(scene.opacities), d(scene.scales), d(scene.rotations),
                    1.0, e, d(cam.world_view_transform), d(cam.full_proj_transform), cam.tanfovx, cam.tanfovy, H, W,
                    d(scene.shs), deg, d(cam.camera_center), False, False)
R, binning, image = out[0], out[5], out[6]
lib = _lib.load()
ko = lib.gsm_binning_sorted_keys_offset(binning.gsm_capacity)
base = (binning.data_ptr() + 127) // 128 * 128 - binning.data_ptr()
keys = binning[base + ko: base + ko + 8 * R].view(torch.int64)
depth = (keys & 0xffffffff)
tile = keys >> 32
Nt = ((W + 15) // 16) * ((H + 15) // 16)
cnt = torch.bincount(tile, minlength=Nt)
# per tile OR of xor with first
first = torch.zeros(Nt, dtype=torch.int64, device="cuda"); 
idx = torch.cumsum(cnt, 0) - cnt
valid = cnt > 0
first[valid] = depth[idx[valid]]
x = depth ^ first[tile]
# OR-reduce per tile via bit planes
nb = torch.zeros(Nt, dtype=torch.int64, device="cuda")
for b in range(32):
    has = torch.zeros(Nt, dtype=torch.int64, device="cuda").scatter_add_(0, tile, (x >> b) & 1) > 0
    nb = torch.where(has, torch.full_like(nb, b + 1), nb)
print("R", R, "tiles", Nt, "mean n", cnt.float().mean().item(), "max n", cnt.max().item())
print("n classes: <=4096:", (cnt <= 4096).sum().item(), " <=9216:", ((cnt > 4096) & (cnt <= 9216)).sum().item(), " >9216:", (cnt > 9216).sum().item())
print("elements in classes:", cnt[cnt <= 4096].sum().item(), cnt[(cnt > 4096) & (cnt <= 9216)].sum().item(), cnt[cnt > 9216].sum().item())
for k in range(20, 33): print("nbits", k, "tiles", (nb == k).sum().item(), "elems", cnt[nb == k].sum().item())
