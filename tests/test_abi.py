"""CPU: the C-ABI shared library loads and exports every symbol include/gaussmart_b200.h declares;
host-only sizing arithmetic is sane.  No compute call is made (there is no GPU here)."""
import ctypes
import os
import re

import pytest

from gaussmart_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "gaussmart_b200.h")


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(gsm_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_something():
    syms = declared_symbols()
    assert "gsm_forward" in syms and "gsm_backward" in syms and "gsm_knn_dist2" in syms
    assert "gsm_mark_visible" in syms


def test_library_exports_every_declared_symbol():
    assert os.path.exists(_lib.LIB_PATH), "run `python -m gaussmart_b200.build`"
    lib = ctypes.CDLL(_lib.LIB_PATH)
    missing = [s for s in declared_symbols() if not hasattr(lib, s)]
    assert not missing, missing


def test_python_binding_lists_every_declared_symbol():
    assert sorted(_lib.EXPORTS) == declared_symbols()


def test_abi_version_and_constants():
    lib = _lib.load()
    assert lib.gsm_abi_version() == _lib.ABI_VERSION
    src = open(HEADER).read()
    assert int(re.search(r"#define GSM_ABI_VERSION (\d+)", src).group(1)) == _lib.ABI_VERSION
    assert int(re.search(r"#define GSM_SURFEL_RECORD_BYTES (\d+)", src).group(1)) == _lib.GSM_SURFEL_RECORD_BYTES
    assert int(re.search(r"#define GSM_ERR_BINNING_OVERFLOW (\d+)", src).group(1)) == _lib.GSM_ERR_BINNING_OVERFLOW


def test_struct_sizes_match_c_layout():
    # 10 x 4-byte scalars, then pointers/size_t (8-byte aligned)
    assert ctypes.sizeof(_lib.ForwardArgs) % 8 == 0
    assert _lib.ForwardArgs.background.offset == 40
    assert _lib.BackwardArgs.background.offset == 40  # 9 scalars = 36 -> padded to 40


@pytest.mark.parametrize("P", [0, 1, 10_000, 5_000_000])
def test_workspace_sizes(P):
    lib = _lib.load()
    g = lib.gsm_geometry_bytes(P)
    assert g >= P * _lib.GSM_SURFEL_RECORD_BYTES + 8 * P
    assert g < P * 128 + (1 << 20)
    assert lib.gsm_grad_workspace_bytes(P) >= P * _lib.GSM_GRAD_RECORD_FLOATS * 4
    assert lib.gsm_knn_workspace_bytes(P) > 0


def test_image_and_binning_sizes():
    lib = _lib.load()
    W, H = 1237, 822
    N = W * H
    Nt = ((W + 15) // 16) * ((H + 15) // 16)
    assert lib.gsm_image_bytes(W, H) >= 20 * N + 8 * Nt
    assert lib.gsm_image_ncontrib_offset(W, H) >= 12 * N
    assert lib.gsm_image_ranges_offset(W, H) >= 20 * N
    R = 10_000_000
    assert lib.gsm_binning_bytes(R, W, H) >= 21 * R
    assert 5 * R <= lib.gsm_binning_persistent_bytes(R) < lib.gsm_binning_bytes(R, W, H)
    assert lib.gsm_binning_sorted_keys_offset(R) >= 4 * R
    assert lib.gsm_binning_bytes(0, W, H) < (1 << 20)


def test_invalid_arguments_are_reported_without_touching_the_gpu():
    lib = _lib.load()
    assert lib.gsm_forward(None, None) == _lib.GSM_ERR_INVALID_ARGUMENT
    assert b"null args" in lib.gsm_last_error()
    a = _lib.ForwardArgs()
    a.P, a.width, a.height = -1, 16, 16
    assert lib.gsm_forward(ctypes.byref(a), None) == _lib.GSM_ERR_INVALID_ARGUMENT
    assert lib.gsm_backward(None, None) == _lib.GSM_ERR_INVALID_ARGUMENT
    assert lib.gsm_knn_dist2(-1, None, None, None, 0, None) == _lib.GSM_ERR_INVALID_ARGUMENT
    assert lib.gsm_knn_dist2(0, None, None, None, 0, None) == _lib.GSM_OK


def test_stage_names():
    lib = _lib.load()
    names = [lib.gsm_stage_name(i).decode() for i in range(lib.gsm_stage_count())]
    assert names[0] == "preprocess_fwd" and "render_bwd" in names and "tile_sort" in names


def test_static_scratch_loss_validates_its_arguments_before_any_cuda_call():
    """gsm_photometric_loss_static checks sizes and pointers first: these calls never reach the device."""
    lib = _lib.load()
    limit = lib.gsm_photometric_loss_static_max_elements()
    src = open(HEADER).read()
    assert limit == int(re.search(r"#define GSM_PHOTOMETRIC_LOSS_STATIC_MAX_ELEMENTS (\d+)", src).group(1))
    dummy = ctypes.create_string_buffer(64)
    ptr = ctypes.cast(dummy, ctypes.c_void_p)
    too_big = limit // (3 * 2048) + 1
    assert lib.gsm_photometric_loss_static(ptr, ptr, 3, too_big, 2048, 0.2, 1.0, None, ptr, None) == _lib.GSM_ERR_WORKSPACE_TOO_SMALL
    assert str(limit) in lib.gsm_last_error().decode()
    assert lib.gsm_photometric_loss_static(ptr, ptr, 0, 8, 8, 0.2, 1.0, None, ptr, None) == _lib.GSM_ERR_INVALID_ARGUMENT
    assert lib.gsm_photometric_loss_static(None, ptr, 3, 8, 8, 0.2, 1.0, None, ptr, None) == _lib.GSM_ERR_INVALID_ARGUMENT
    assert lib.gsm_photometric_loss_static(ptr, ptr, 3, 8, 8, 0.2, 1.0, None, None, None) == _lib.GSM_ERR_INVALID_ARGUMENT
