#!/usr/bin/env python
"""Build the UNMODIFIED reference CUDA extensions for sm_100a into oracle/_ref/.

TEST / BASELINE INFRASTRUCTURE ONLY -- nothing under gaussmart_b200/ may import this.

The reference (alevalve/gaussmart submodules diff-surfel-rasterization and simple-knn) is
CUDA-only, so it cannot execute in the build container (no GPU); it is compiled here from the
sources where they lie under /root/reference (never copied into the repo history) and the
resulting shared objects travel to the B200 box with the gpurun snapshot.  There they serve as
 (1) the bit-exact oracle (radii, sorted keys, tile ranges) and generator of tests/golden/*,
 (2) the `bench.py --impl reference` arm (the ">= 2x" denominator of BASELINE.json).

We do not run the reference's setup.py; this is a direct nvcc/g++ recipe over its source
files.  `-include cstdint` is needed because rasterizer_impl.h:24,40 use std::uintptr_t /
uint32_t without the header under gcc 13.

Outputs (git-ignored, NOT gpurun-ignored):
  oracle/_ref/diff_surfel_rasterization/_C.so  + the reference's __init__.py (install copy)
  oracle/_ref/simple_knn/_C.so
  oracle/_ref/gaussmart/...                    install copy of the reference's Python CALLERS of the hot path
                                               (render(), GaussianModel, Camera, utils/*, arguments): what
                                               `pip install --target` of the reference would put there.  They are the
                                               "unchanged callers" tests/test_reference_callers_gpu.py runs against
                                               both back ends, and the loss / render() code of the bench's reference arm.
"""
import os, shutil, subprocess, sys, sysconfig
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("GAUSSMART_REFERENCE", "/root/reference")
OUT = os.path.join(HERE, "_ref")
OBJ = os.path.join(OUT, "_obj")


def _flags(name):
    import torch
    from torch.utils import cpp_extension as ce
    inc = [f"-I{p}" for p in ce.include_paths("cuda")] + [f"-I{sysconfig.get_paths()['include']}"]
    common = inc + [f"-DTORCH_EXTENSION_NAME=_C", "-DTORCH_API_INCLUDE_EXTENSION_H",
                    "-D_GLIBCXX_USE_CXX11_ABI=" + str(int(torch._C._GLIBCXX_USE_CXX11_ABI)),
                    "-std=c++17", "-O3"]
    libdir = os.path.join(os.path.dirname(torch.__file__), "lib")
    link = [f"-L{libdir}", "-lc10", "-lc10_cuda", "-ltorch_cpu", "-ltorch_cuda", "-ltorch",
            "-ltorch_python", "-L/usr/local/cuda/lib64", "-lcudart", f"-Wl,-rpath,{libdir}"]
    return common, link


def build_one(name, srcdir, sources, extra_nvcc):
    common, link = _flags(name)
    os.makedirs(os.path.join(OBJ, name), exist_ok=True)
    os.makedirs(os.path.join(OUT, name), exist_ok=True)
    objs, cmds = [], []
    for s in sources:
        o = os.path.join(OBJ, name, s.replace("/", "_") + ".o")
        objs.append(o)
        src = os.path.join(srcdir, s)
        if s.endswith(".cu"):
            cmds.append(["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-Xcompiler", "-fPIC",
                         "-include", "cstdint", "--expt-relaxed-constexpr", "-c", src, "-o", o]
                        + common + extra_nvcc)
        else:
            cmds.append(["g++", "-fPIC", "-c", src, "-o", o] + common)
    return objs, cmds, link


# the reference's callers of the hot path -- copied, never edited
CALLERS = ["gaussian_renderer/__init__.py", "scene/gaussian_model.py", "scene/cameras.py", "utils/point_utils.py",
           "utils/sh_utils.py", "utils/general_utils.py", "utils/graphics_utils.py", "utils/loss_utils.py",
           "utils/system_utils.py", "arguments/__init__.py"]


def install_callers():
    """`scene/` and `utils/` are installed WITHOUT their __init__.py (namespace packages): scene/__init__.py pulls in the
    dataset readers (PIL / plyfile / COLMAP), which are not callers of the hot path."""
    dst_root = os.path.join(OUT, "gaussmart")
    shutil.rmtree(dst_root, ignore_errors=True)
    for rel in CALLERS:
        dst = os.path.join(dst_root, rel)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        shutil.copy(os.path.join(REF, rel), dst)
    print("[build_ref] installed", len(CALLERS), "caller modules into", dst_root)


def main():
    if not os.path.isdir(REF):
        print(f"[build_ref] {REF} not present; keeping whatever is prebuilt in {OUT}")
        return 0
    dsr = os.path.join(REF, "submodules/diff-surfel-rasterization")
    knn = os.path.join(REF, "submodules/simple-knn")
    jobs = [
        ("diff_surfel_rasterization", dsr,
         ["cuda_rasterizer/rasterizer_impl.cu", "cuda_rasterizer/forward.cu",
          "cuda_rasterizer/backward.cu", "rasterize_points.cu", "ext.cpp"],
         ["-I" + os.path.join(dsr, "third_party/glm")]),
        ("simple_knn", knn, ["spatial.cu", "simple_knn.cu", "ext.cpp"], []),
    ]
    allcmds, links = [], []
    for name, srcdir, sources, extra in jobs:
        target = os.path.join(OUT, name, "_C.so")
        if os.path.exists(target) and "--force" not in sys.argv:
            print(f"[build_ref] {target} exists, skipping")
            continue
        objs, cmds, link = build_one(name, srcdir, sources, extra)
        allcmds += cmds
        links.append((target, objs, link))
    with ThreadPoolExecutor(8) as ex:
        for r in ex.map(lambda c: subprocess.run(c, capture_output=True, text=True), allcmds):
            if r.returncode != 0:
                print(r.stderr[-4000:])
                return 1
    for target, objs, link in links:
        subprocess.check_call(["g++", "-shared", "-o", target] + objs + link)
        print("[build_ref] built", target)
    # install copy of the reference's python front end (as `pip install --target` would do)
    shutil.copy(os.path.join(dsr, "diff_surfel_rasterization/__init__.py"),
                os.path.join(OUT, "diff_surfel_rasterization/__init__.py"))
    with open(os.path.join(OUT, "simple_knn/__init__.py"), "w") as f:
        f.write("")
    install_callers()
    shutil.rmtree(OBJ, ignore_errors=True)
    return 0


if __name__ == "__main__":
    sys.exit(main())
