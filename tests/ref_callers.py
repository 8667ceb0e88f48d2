"""Loads the reference's UNCHANGED Python callers of the hot path -- gaussian_renderer.render(), scene.gaussian_model.
GaussianModel, scene.cameras.Camera, utils/*, arguments (the install copy oracle/build_ref.py leaves under
oracle/_ref/gaussmart/) -- once per BACK END, so that one process can run the same caller code against
  "ours"      : gaussmart_b200's drop-in packages (`diff_surfel_rasterization`, `simple_knn._C`, gaussmart_b200.install())
  "reference" : the unmodified reference CUDA build (oracle/_ref/*/_C.so)
and compare.  Test / baseline infrastructure only.

The callers bind their back end at import time (`from diff_surfel_rasterization import ...`,
`from simple_knn._C import distCUDA2`), so each load imports a fresh copy of the caller modules with sys.modules pointing
at the chosen back end, then takes the copies out of sys.modules again.  Third-party modules the callers import but the
hot path never touches and this image lacks (matplotlib, plyfile) are stubbed.
"""
import importlib
import os
import sys
import types

import ref_loader

CALLERS_DIR = os.path.join(ref_loader.REF, "gaussmart")
_CALLER_TOPS = ("gaussian_renderer", "scene", "utils", "arguments")
_BACKEND_MODS = ("diff_surfel_rasterization", "simple_knn", "simple_knn._C")
_cache = {}


def available():
    return ref_loader.available() and os.path.exists(os.path.join(CALLERS_DIR, "gaussian_renderer", "__init__.py"))


def _stub(name, **attrs):
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    return m


def _third_party_stubs():
    stubs = {}
    for name in ("matplotlib", "matplotlib.pyplot", "plyfile"):
        try:
            importlib.import_module(name)
        except Exception:
            stubs[name] = _stub(name, PlyData=object, PlyElement=object)
    if "matplotlib" in stubs and "matplotlib.pyplot" in stubs:
        stubs["matplotlib"].pyplot = stubs["matplotlib.pyplot"]
    return stubs


def _purge(prefixes):
    saved = {}
    for k in list(sys.modules):
        if any(k == p or k.startswith(p + ".") for p in prefixes):
            saved[k] = sys.modules.pop(k)
    return saved


class Callers:
    """The caller modules bound to one back end."""
    def __init__(self, backend, mods):
        self.backend = backend
        self.gaussian_renderer = mods["gaussian_renderer"]
        self.render = mods["gaussian_renderer"].render
        self.GaussianModel = mods["scene.gaussian_model"].GaussianModel
        self.Camera = mods["scene.cameras"].Camera
        self.MiniCam = mods["scene.cameras"].MiniCam
        self.BasicPointCloud = mods["utils.graphics_utils"].BasicPointCloud
        self.loss_utils = mods["utils.loss_utils"]
        self.point_utils = mods["utils.point_utils"]
        self.arguments = mods["arguments"]
        self.dsr = mods["diff_surfel_rasterization"]


def load(backend):
    """backend: "ours" | "reference"."""
    if backend in _cache:
        return _cache[backend]
    import torch  # noqa: F401
    if backend == "ours":
        import gaussmart_b200
        gaussmart_b200.install()
        saved_backend = _purge(_BACKEND_MODS)
        dsr = importlib.import_module("diff_surfel_rasterization")
        knn_c = importlib.import_module("simple_knn._C")
        assert "gaussmart_b200" in os.path.realpath(dsr.__file__), dsr.__file__
        backend_mods = {k: sys.modules[k] for k in _BACKEND_MODS}
    elif backend == "reference":
        saved_backend = _purge(_BACKEND_MODS)
        dsr = ref_loader.load_dsr()
        knn_c = ref_loader.load_knn()
        backend_mods = {"diff_surfel_rasterization": dsr, "simple_knn": sys.modules["ref_knn"], "simple_knn._C": knn_c}
    else:
        raise ValueError(backend)
    saved_callers = _purge(_CALLER_TOPS)
    stubs = _third_party_stubs()
    sys.modules.update(stubs)
    sys.modules.update(backend_mods)
    sys.path.insert(0, CALLERS_DIR)
    try:
        importlib.import_module("gaussian_renderer")
        importlib.import_module("scene.gaussian_model")
        importlib.import_module("scene.cameras")
        importlib.import_module("utils.loss_utils")
        importlib.import_module("arguments")
        mods = {k: v for k, v in sys.modules.items() if any(k == p or k.startswith(p + ".") for p in _CALLER_TOPS)}
        mods["diff_surfel_rasterization"] = dsr
    finally:
        sys.path.remove(CALLERS_DIR)
        _purge(_CALLER_TOPS)
        _purge(_BACKEND_MODS)
        for k in stubs:
            sys.modules.pop(k, None)
        sys.modules.update(saved_callers)
        sys.modules.update(saved_backend)
    for k, m in mods.items():  # every copy really is the install copy of the reference, bound to the chosen back end
        if k.split(".")[0] in _CALLER_TOPS and getattr(m, "__file__", None):
            assert os.path.realpath(m.__file__).startswith(os.path.realpath(CALLERS_DIR)), (k, m.__file__)
    gm = mods["scene.gaussian_model"]
    assert gm.distCUDA2 is knn_c.distCUDA2
    assert mods["gaussian_renderer"].GaussianRasterizer is dsr.GaussianRasterizer
    _cache[backend] = Callers(backend, mods)
    return _cache[backend]
