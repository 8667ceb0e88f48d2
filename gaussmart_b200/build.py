"""Builds gaussmart_b200/lib/libgaussmart_b200.so (hand-written sm_100a CUDA behind a C ABI).

Plain nvcc, in-tree, no torch headers: the library has no torch types in its interface
(include/gaussmart_b200.h).  `python -m gaussmart_b200.build [--force]`.
"""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIBDIR = os.path.join(HERE, "lib")
LIB = os.path.join(LIBDIR, "libgaussmart_b200.so")
SOURCES = ["api.cu", "preprocess.cu", "binning.cu", "sort.cu", "render_fwd.cu", "render_bwd.cu",
           "preprocess_bwd.cu", "knn.cu", "loss.cu", "adam.cu", "regularizer.cu", "densify.cu", "exchange.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
              "-Xcompiler", "-fPIC"]


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build_native(force=False, verbose=False, defines=(), tag=None):
    """tag/defines: build a tuning variant lib/libgaussmart_b200_<tag>.so with extra -D flags (selected at run
    time with GAUSSMART_B200_LIB=<path>); the default build takes neither."""
    os.makedirs(LIBDIR, exist_ok=True)
    objdir = os.path.join(LIBDIR, "obj" if tag is None else "obj_" + tag)
    lib = LIB if tag is None else os.path.join(LIBDIR, f"libgaussmart_b200_{tag}.so")
    flags = NVCC_FLAGS + ["-D" + d for d in defines]
    os.makedirs(objdir, exist_ok=True)
    headers = [os.path.join(CSRC, "common.cuh"),
               os.path.join(HERE, "..", "include", "gaussmart_b200.h")]
    cmds, objs = [], []
    for s in SOURCES:
        src = os.path.join(CSRC, s)
        obj = os.path.join(objdir, s + ".o")
        objs.append(obj)
        if force or _stale(obj, [src] + headers):
            cmds.append(["nvcc"] + flags + ["-c", src, "-o", obj])
    if cmds:
        with ThreadPoolExecutor(8) as ex:
            for cmd, r in zip(cmds, ex.map(lambda c: subprocess.run(c, capture_output=True, text=True), cmds)):
                if verbose and r.stderr:
                    print(r.stderr)
                if r.returncode != 0:
                    raise RuntimeError("nvcc failed: %s\n%s" % (" ".join(cmd), r.stderr[-6000:]))
    if cmds or not os.path.exists(lib):
        subprocess.check_call(["nvcc", "-shared", "-o", lib] + objs +
                              ["-gencode", "arch=compute_100a,code=sm_100a"])
    return lib


if __name__ == "__main__":
    defs = [a[2:] for a in sys.argv[1:] if a.startswith("-D")]
    tags = [a.split("=", 1)[1] for a in sys.argv[1:] if a.startswith("--tag=")]
    print(build_native(force="--force" in sys.argv, verbose=True, defines=defs, tag=tags[0] if tags else None))
