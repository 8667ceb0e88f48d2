"""Loads the UNMODIFIED reference CUDA build from oracle/_ref under alias module names
(`ref_dsr`, `ref_knn`) so it can live in the same process as gaussmart_b200.
Test / baseline infrastructure only."""
import importlib
import importlib.util
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "oracle", "_ref")


def available():
    return (os.path.exists(os.path.join(REF, "diff_surfel_rasterization", "_C.so")) and
            os.path.exists(os.path.join(REF, "simple_knn", "_C.so")))


def _load_pkg(alias, path):
    if alias in sys.modules:
        return sys.modules[alias]
    spec = importlib.util.spec_from_file_location(alias, os.path.join(path, "__init__.py"),
                                                  submodule_search_locations=[path])
    m = importlib.util.module_from_spec(spec)
    sys.modules[alias] = m
    spec.loader.exec_module(m)
    return m


def load_dsr():
    import torch  # noqa: F401  (libtorch must be loaded before the extension)
    return _load_pkg("ref_dsr", os.path.join(REF, "diff_surfel_rasterization"))


def load_knn():
    import torch  # noqa: F401
    _load_pkg("ref_knn", os.path.join(REF, "simple_knn"))
    return importlib.import_module("ref_knn._C")


def align128(x):
    return (x + 127) // 128 * 128


def ref_binning_views(binning, R):
    """(point_list u32[R], sorted keys u64[R]) out of the reference's binningBuffer
    (rasterizer_impl.cu:181-194)."""
    import torch
    o_pl = 0
    o_plu = align128(o_pl + 4 * R)
    o_keys = align128(o_plu + 4 * R)
    pl = binning[o_pl:o_pl + 4 * R].view(torch.int32)
    keys = binning[o_keys:o_keys + 8 * R].view(torch.int64)
    return pl, keys


def ref_image_views(img, W, H):
    """(final_T f32[3,N], n_contrib i32[2,N], ranges i32[Nt,2]) out of the reference's imgBuffer
    (rasterizer_impl.cu:172-179)."""
    import torch
    N = W * H
    Nt = ((W + 15) // 16) * ((H + 15) // 16)
    o_acc = 0
    o_nc = align128(o_acc + 12 * N)
    o_rg = align128(o_nc + 8 * N)
    acc = img[o_acc:o_acc + 12 * N].view(torch.float32).view(3, N)
    nc = img[o_nc:o_nc + 8 * N].view(torch.int32).view(2, N)
    rg = img[o_rg:o_rg + 8 * Nt].view(torch.int32).view(Nt, 2)
    return acc, nc, rg
