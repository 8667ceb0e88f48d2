"""Test helper (not part of the product package): views into the workspaces left by gsm_forward, for parity tests (sorted keys, tile ranges,
final_T, n_contrib).  Layout offsets come from the library itself."""
import torch

from gaussmart_b200 import _lib
from gaussmart_b200.rasterizer import _forward_impl


def forward_with_intermediates(settings, means3D, opacities, shs=None, colors_precomp=None, scales=None,
                               rotations=None, cov3D_precomp=None):
    """Runs the forward and returns a dict with outputs and bit-exact intermediates."""
    lib = _lib.load()
    dev = means3D.device
    e = lambda t: torch.empty(0, device=dev) if t is None else t
    R, color, others, radii, geom, binning, image = _forward_impl(
        settings.bg, means3D, e(colors_precomp), opacities, e(scales), e(rotations), settings.scale_modifier,
        e(cov3D_precomp), settings.viewmatrix, settings.projmatrix, settings.tanfovx, settings.tanfovy,
        settings.image_height, settings.image_width, e(shs), settings.sh_degree, settings.campos,
        settings.prefiltered, settings.debug)
    W, H = settings.image_width, settings.image_height
    N = W * H
    Nt = ((W + 15) // 16) * ((H + 15) // 16)
    cap = binning.gsm_capacity  # the capacity the forward ran with travels with the buffer
    ko = lib.gsm_binning_sorted_keys_offset(cap)
    out = dict(num_rendered=R, color=color, allmap=others, radii=radii)
    if R > 0:
        out["point_list"] = binning[:4 * R].view(torch.int32)
        out["keys"] = binning[ko:ko + 8 * R].view(torch.int64)
    else:
        out["point_list"] = torch.empty(0, dtype=torch.int32, device=dev)
        out["keys"] = torch.empty(0, dtype=torch.int64, device=dev)
    ro = lib.gsm_image_ranges_offset(W, H)
    no = lib.gsm_image_ncontrib_offset(W, H)
    out["ranges"] = image[ro:ro + 8 * Nt].view(torch.int32).view(Nt, 2)
    out["n_contrib"] = image[no:no + 8 * N].view(torch.int32).view(2, N)
    out["final_T"] = image[0:12 * N].view(torch.float32).view(3, N)
    return out
