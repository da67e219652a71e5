"""In-tree build of the two native artefacts (explicit nvcc / g++ commands, no JIT cache):

* ``lib/libtsp.so``  -- torch-free C-ABI + sm_100a kernels (``csrc/tsp_abi.cu``)
* ``torch_sputnik.<abi>.so`` -- the ``torch_sputnik`` Python module (``csrc/torch_shim.cpp``),
  linked against libtsp.so with an ``$ORIGIN/lib`` rpath.

Both are git-ignored and travel to the GPU box with the snapshot.
"""
from __future__ import annotations

import os
import subprocess
import sys
import sysconfig

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
LIB_DIR = os.path.join(HERE, "lib")
LIBTSP = os.path.join(LIB_DIR, "libtsp.so")
EXT_SUFFIX = sysconfig.get_config_var("EXT_SUFFIX") or ".so"
SHIM = os.path.join(HERE, "torch_sputnik" + EXT_SUFFIX)

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
    "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr",
    "-diag-suppress", "177,550",
]


def _newer(target: str, sources) -> bool:
    if not os.path.exists(target):
        return False
    t = os.path.getmtime(target)
    return all(os.path.getmtime(s) <= t for s in sources)


def _sources(*dirs_or_files):
    out = []
    for d in dirs_or_files:
        if os.path.isdir(d):
            for base, _, files in os.walk(d):
                out += [os.path.join(base, f) for f in files
                        if f.endswith((".cu", ".cuh", ".cpp", ".h"))]
        else:
            out.append(d)
    return out


def build_libtsp(force: bool = False, verbose: bool = False) -> str:
    srcs = _sources(os.path.join(CSRC, "tsp_abi.cu"), os.path.join(CSRC, "kernels"),
                    os.path.join(ROOT, "include", "tsp.h"))
    if not force and _newer(LIBTSP, srcs):
        return LIBTSP
    os.makedirs(LIB_DIR, exist_ok=True)
    nvcc = os.environ.get("NVCC", "nvcc")
    cmd = [nvcc, *NVCC_FLAGS, "-shared", "-cudart", "shared", "-Xlinker",
           "-rpath,/usr/local/cuda/lib64", "-o", LIBTSP, os.path.join(CSRC, "tsp_abi.cu")]
    if verbose:
        cmd += ["-Xptxas", "-v"]
    subprocess.check_call(cmd, cwd=ROOT)
    return LIBTSP


def build_shim(force: bool = False) -> str:
    src = os.path.join(CSRC, "torch_shim.cpp")
    if not force and _newer(SHIM, [src, os.path.join(CSRC, "many_mask.inc"), os.path.join(ROOT, "include", "tsp.h")]):
        return SHIM
    import torch
    from torch.utils import cpp_extension as ce

    inc = []
    for p in ce.include_paths("cuda") if "device_type" in ce.include_paths.__code__.co_varnames \
            else ce.include_paths(cuda=True):
        inc += ["-isystem", p]
    inc += ["-isystem", sysconfig.get_paths()["include"]]
    libdirs = []
    for p in ce.library_paths("cuda") if "device_type" in ce.library_paths.__code__.co_varnames \
            else ce.library_paths(cuda=True):
        libdirs += ["-L" + p, "-Wl,-rpath," + p]
    abi = int(torch._C._GLIBCXX_USE_CXX11_ABI)
    cmd = ["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-o", SHIM, src,
           "-DTORCH_EXTENSION_NAME=torch_sputnik", "-DTORCH_API_INCLUDE_EXTENSION_H",
           f"-D_GLIBCXX_USE_CXX11_ABI={abi}", "-Wno-deprecated-declarations",
           *inc, *libdirs, "-L" + LIB_DIR, "-Wl,-rpath,$ORIGIN/lib",
           "-ltsp", "-lc10", "-lc10_cuda", "-ltorch_cpu", "-ltorch_cuda", "-ltorch", "-ltorch_python",
           "-lcudart"]
    subprocess.check_call(cmd, cwd=ROOT)
    return SHIM


def build_all(force: bool = False, verbose: bool = False):
    return build_libtsp(force, verbose), build_shim(force)


if __name__ == "__main__":
    paths = build_all(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print("\n".join(paths))
