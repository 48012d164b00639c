#!/usr/bin/env python
"""TEST INFRASTRUCTURE — builds the UNMODIFIED reference CUDA extensions for sm_100.

Compiles, from the sources where they lie under /root/reference (never copied
into this repo), the two native modules the reference's hot path uses:

  * diff-gaussian-rasterization-w-pose  (DGR/ext.cpp, DGR/rasterize_points.cu,
    CR/rasterizer_impl.cu, CR/forward.cu, CR/backward.cu)  -> oracle/_ref/dgr_ref_C*.so
  * simple-knn (KNN/ext.cpp, KNN/spatial.cu, KNN/simple_knn.cu)  -> oracle/_ref/knn_ref_C*.so

The reference's own build system (setup.py / CMake) is NOT run; this is a direct
nvcc invocation.  Two forced includes are needed because the sources rely on
transitive includes that newer toolchains no longer provide:
`-include cstdint` (rasterizer_impl.h) and `-include cfloat` (simple_knn.cu).

Outputs go only to oracle/_ref/ (git-ignored, but shipped to the GPU box).
The products are used by tests/ (GPU parity: new kernels vs the reference's)
and by `bench.py --impl reference`.  Nothing in the product path imports them.
"""
import os
import subprocess
import sys
import sysconfig
import time
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref")
REF = os.environ.get("B2GS_REFERENCE_ROOT", "/root/reference")
DGR = os.path.join(REF, "submodules", "diff-gaussian-rasterization")
KNN = os.path.join(REF, "submodules", "simple-knn")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
ARCH = ["-gencode", "arch=compute_100,code=sm_100"]


def _torch_flags():
    import torch
    from torch.utils import cpp_extension as ce

    inc = []
    for p in ce.include_paths():
        inc += ["-isystem", p]
    inc += ["-isystem", sysconfig.get_paths()["include"]]
    libdir = os.path.join(os.path.dirname(torch.__file__), "lib")
    cxx11 = int(torch._C._GLIBCXX_USE_CXX11_ABI)
    defs = [f"-D_GLIBCXX_USE_CXX11_ABI={cxx11}", "-DTORCH_API_INCLUDE_EXTENSION_H"]
    link = [f"-L{libdir}", "-lc10", "-ltorch", "-ltorch_cpu", "-ltorch_python",
            "-lc10_cuda", "-ltorch_cuda", f"-Xlinker", f"-rpath={libdir}"]
    return inc, defs, link


def _run(cmd):
    print(" ".join(cmd), flush=True)
    subprocess.check_call(cmd)


def _obj(src, name, extra, inc, defs, modname):
    o = os.path.join(OUT, "obj", name + ".o")
    if os.path.exists(o) and os.path.getmtime(o) > os.path.getmtime(src):
        return o
    os.makedirs(os.path.dirname(o), exist_ok=True)
    cmd = [NVCC, "-c", src, "-o", o, "-O3", "-std=c++17", "--expt-relaxed-constexpr",
           "-Xcompiler", "-fPIC", "-w", f"-DTORCH_EXTENSION_NAME={modname}"] + ARCH + extra + inc + defs
    _run(cmd)
    return o


def build(force=False):
    if not os.path.isdir(DGR):
        print(f"[build_ref] {DGR} not present; keeping prebuilt files in {OUT}")
        return False
    os.makedirs(OUT, exist_ok=True)
    inc, defs, link = _torch_flags()
    suffix = sysconfig.get_config_var("EXT_SUFFIX")
    t0 = time.time()

    # --- rasterizer ---
    target = os.path.join(OUT, "dgr_ref_C" + suffix)
    if force or not os.path.exists(target):
        glm = ["-I", os.path.join(DGR, "third_party", "glm"), "-include", "cstdint"]
        srcs = [
            (os.path.join(DGR, "cuda_rasterizer", "rasterizer_impl.cu"), "dgr_rasterizer_impl"),
            (os.path.join(DGR, "cuda_rasterizer", "forward.cu"), "dgr_forward"),
            (os.path.join(DGR, "cuda_rasterizer", "backward.cu"), "dgr_backward"),
            (os.path.join(DGR, "rasterize_points.cu"), "dgr_rasterize_points"),
            (os.path.join(DGR, "ext.cpp"), "dgr_ext"),
        ]
        def one(sn):
            s, n = sn
            extra = list(glm) + (["-x", "cu"] if s.endswith(".cpp") else [])
            return _obj(s, n, extra, inc, defs, "dgr_ref_C")
        with ThreadPoolExecutor(max_workers=len(srcs)) as ex:  # torch headers make each TU minutes long
            objs = list(ex.map(one, srcs))
        _run([NVCC, "-shared", "-o", target] + objs + ARCH + link)

    # --- simple-knn ---
    target = os.path.join(OUT, "knn_ref_C" + suffix)
    if force or not os.path.exists(target):
        extra0 = ["-include", "cfloat", "-include", "cstdint"]
        srcs = [
            (os.path.join(KNN, "simple_knn.cu"), "knn_simple_knn"),
            (os.path.join(KNN, "spatial.cu"), "knn_spatial"),
            (os.path.join(KNN, "ext.cpp"), "knn_ext"),
        ]
        def one_knn(sn):
            s, n = sn
            extra = list(extra0) + (["-x", "cu"] if s.endswith(".cpp") else [])
            return _obj(s, n, extra, inc, defs, "knn_ref_C")
        with ThreadPoolExecutor(max_workers=len(srcs)) as ex:
            objs = list(ex.map(one_knn, srcs))
        _run([NVCC, "-shared", "-o", target] + objs + ARCH + link)
    print(f"[build_ref] done in {time.time() - t0:.1f}s -> {OUT}")
    return True


if __name__ == "__main__":
    build(force="--force" in sys.argv)
