"""CPU: the host-side mirror of the reference's Python surface: names, field order,
validation messages (dsr.py:192-196), CPU tensors rejected (rasterize_points.cu:27-28), drop-in
package names, and the rule that the product never imports the oracle."""
import os
import re
import sys

import pytest
import torch

import gaussmart_b200 as gb

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _settings(**over):
    kw = dict(image_height=16, image_width=16, tanfovx=0.5, tanfovy=0.5, bg=torch.zeros(3), scale_modifier=1.0,
              viewmatrix=torch.eye(4), projmatrix=torch.eye(4), sh_degree=0, campos=torch.zeros(3),
              prefiltered=False, debug=False)
    kw.update(over)
    return gb.GaussianRasterizationSettings(**kw)


def test_settings_field_order_is_the_references():
    # dsr.py:158-170 -- callers may construct it positionally
    assert gb.GaussianRasterizationSettings._fields == (
        "image_height", "image_width", "tanfovx", "tanfovy", "bg", "scale_modifier", "viewmatrix", "projmatrix",
        "sh_degree", "campos", "prefiltered", "debug")


def test_exactly_one_colour_source():
    r = gb.GaussianRasterizer(_settings())
    x = torch.zeros(4, 3)
    with pytest.raises(Exception, match="Please provide excatly one of either SHs or precomputed colors!"):
        r(means3D=x, means2D=x, opacities=torch.zeros(4, 1), scales=torch.ones(4, 2), rotations=torch.ones(4, 4))
    with pytest.raises(Exception, match="Please provide excatly one of either SHs or precomputed colors!"):
        r(means3D=x, means2D=x, opacities=torch.zeros(4, 1), shs=torch.zeros(4, 1, 3), colors_precomp=x,
          scales=torch.ones(4, 2), rotations=torch.ones(4, 4))


def test_exactly_one_geometry_source():
    r = gb.GaussianRasterizer(_settings())
    x = torch.zeros(4, 3)
    msg = "Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!"
    with pytest.raises(Exception, match=msg):
        r(means3D=x, means2D=x, opacities=torch.zeros(4, 1), colors_precomp=x)
    with pytest.raises(Exception, match=msg):
        r(means3D=x, means2D=x, opacities=torch.zeros(4, 1), colors_precomp=x, scales=torch.ones(4, 2))
    with pytest.raises(Exception, match=msg):
        r(means3D=x, means2D=x, opacities=torch.zeros(4, 1), colors_precomp=x, scales=torch.ones(4, 2),
          rotations=torch.ones(4, 4), cov3D_precomp=torch.zeros(4, 9))


def test_cpu_tensors_are_rejected_not_silently_computed():
    """There is no CPU path: a CPU tensor raises, as the reference's CHECK_CUDA does."""
    r = gb.GaussianRasterizer(_settings())
    x = torch.zeros(4, 3)
    with pytest.raises(RuntimeError, match="must be a CUDA tensor"):
        r(means3D=x, means2D=x, opacities=torch.zeros(4, 1), colors_precomp=x, scales=torch.ones(4, 2),
          rotations=torch.ones(4, 4))
    with pytest.raises(RuntimeError, match="must be a CUDA tensor"):
        gb.distCUDA2(torch.zeros(8, 3))
    with pytest.raises(RuntimeError, match="must be a CUDA tensor"):
        r.markVisible(x)
    from gaussmart_b200.optim import add_densification_stats
    with pytest.raises(RuntimeError, match="must be a CUDA tensor"):
        add_densification_stats(torch.zeros(4), torch.zeros(4, 1), torch.zeros(4, 1), torch.ones(4, dtype=torch.int32), x)


def test_means3d_shape_message():
    r = gb.GaussianRasterizer(_settings())
    x = torch.zeros(4, 2)
    with pytest.raises(RuntimeError, match=r"means3D must have dimensions \(num_points, 3\)"):
        r(means3D=x, means2D=x, opacities=torch.zeros(4, 1), colors_precomp=torch.zeros(4, 3),
          scales=torch.ones(4, 2), rotations=torch.ones(4, 4))


def test_drop_in_package_names_resolve():
    gb.install()
    import diff_surfel_rasterization as dsr
    from simple_knn._C import distCUDA2
    assert dsr.GaussianRasterizer is gb.GaussianRasterizer
    assert dsr.GaussianRasterizationSettings is gb.GaussianRasterizationSettings
    assert distCUDA2 is gb.distCUDA2
    assert hasattr(dsr, "rasterize_gaussians") and hasattr(dsr, "_RasterizeGaussians")


def test_product_never_imports_the_oracle():
    pat = re.compile(r"^\s*(from|import)\s+.*\b(oracle|surfel_oracle)\b|oracle[/\\.]surfel_oracle|oracle/_ref", re.M)
    pkg = os.path.join(ROOT, "gaussmart_b200")
    offenders = []
    for d, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                if pat.search(open(os.path.join(d, f), errors="replace").read()):
                    offenders.append(os.path.join(d, f))
    assert not offenders, offenders


def test_missing_library_fails_loudly(monkeypatch):
    from gaussmart_b200 import _lib
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", "/nonexistent/libgaussmart_b200.so")
    with pytest.raises(ImportError, match="no CPU or PyTorch fallback"):
        _lib.load()


def test_no_triton_or_compile_in_product():
    pkg = os.path.join(ROOT, "gaussmart_b200")
    for d, _, files in os.walk(pkg):
        for f in files:
            if f.endswith(".py"):
                src = open(os.path.join(d, f)).read()
                assert "import triton" not in src and "torch.compile" not in src, f
