"""Build recipe for libk8s_place_b200.so (explicit nvcc, in-tree, sm_100a only).

The library is built next to this file so that it travels with the source tree; a JIT
cache under ~/.cache would not.  Flags: -fmad=false keeps nvcc from contracting a*b+c
(the kernels also use explicit round-to-nearest intrinsics), -lineinfo makes ncu's source
page map to our code.
"""
from __future__ import annotations

import os
import shutil
import subprocess
from pathlib import Path

PKG_DIR = Path(__file__).resolve().parent
CSRC = PKG_DIR / "csrc"
LIB_NAME = "libk8s_place_b200.so"
LIB_PATH = PKG_DIR / LIB_NAME
SOURCES = ["kp_api.cu", "kp_snapshot.cu", "kp_greedy.cu", "kp_greedy_fast.cu", "kp_sort.cu", "kp_ranksort.cu", "kp_ingest.cu", "kp_cachekey.cu", "kp_diag.cu", "kp_exchange.cu"]
NVCC_FLAGS = [
    "-shared", "-Xcompiler", "-fPIC", "-O3", "-std=c++17",
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-fmad=false",
]


def find_nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and Path(cand).exists():
            return cand
    raise RuntimeError("nvcc not found (set NVCC=/path/to/nvcc)")


def _stale() -> bool:
    if not LIB_PATH.exists():
        return True
    built = LIB_PATH.stat().st_mtime
    deps = list(CSRC.glob("*.cu")) + list(CSRC.glob("*.cuh")) + \
        [PKG_DIR.parent / "include" / "k8s_place_b200.h"]
    return any(d.stat().st_mtime > built for d in deps)


def build_library(force: bool = False, verbose: bool = False) -> Path:
    """Compile the CUDA library if it is missing or older than its sources."""
    if not force and not _stale():
        return LIB_PATH
    extra = os.environ.get("KP_NVCC_EXTRA", "").split()          # experiments only (e.g. -DKP_UPDATE_MODE=1)
    out = Path(os.environ.get("KP_LIB_OUT", str(LIB_PATH)))
    cmd = [find_nvcc(), *NVCC_FLAGS, *extra, "-o", str(out), *[str(CSRC / s) for s in SOURCES]]
    if verbose:
        cmd.insert(1, "-Xptxas")
        cmd.insert(2, "-v")
        print(" ".join(cmd))
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError(f"nvcc failed:\n{res.stdout}\n{res.stderr}")
    if verbose:
        print(res.stderr)
    return LIB_PATH


def pyobj_path() -> Path:
    import sysconfig
    return PKG_DIR / ("_kp_pyobj" + (sysconfig.get_config_var("EXT_SUFFIX") or ".so"))


def build_pyobj(force: bool = False) -> Path:
    """gcc build of the small CPython extension that walks reference-style record lists in C
    (csrc/kp_pyobj.c); optional: tables.py keeps the same conversions in Python."""
    import sysconfig
    out, src = pyobj_path(), CSRC / "kp_pyobj.c"
    if not force and out.exists() and out.stat().st_mtime >= src.stat().st_mtime:
        return out
    cc = os.environ.get("CC") or shutil.which("gcc") or shutil.which("cc")
    if not cc:
        raise RuntimeError("no C compiler for the record-walker extension")
    cmd = [cc, "-O2", "-shared", "-fPIC", "-I", sysconfig.get_paths()["include"], str(src), "-o", str(out), "-lm"]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError(f"{cc} failed:\n{res.stdout}\n{res.stderr}")
    return out


TORCH_EXT_NAME = "libk8s_place_torch.so"
TORCH_EXT_PATH = PKG_DIR / TORCH_EXT_NAME


def build_torch_ext(force: bool = False) -> Path:
    """g++ build of the PyTorch-extension form of the ops (csrc/kp_torch_ext.cpp: TORCH_LIBRARY(k8s_place, ...)),
    in-tree and linked against libk8s_place_b200.so next to it (rpath $ORIGIN).  No nvcc: the file holds no kernels."""
    import torch
    from torch.utils import cpp_extension as ce
    src = CSRC / "kp_torch_ext.cpp"
    deps = [src, PKG_DIR.parent / "include" / "k8s_place_b200.h"]
    if not force and TORCH_EXT_PATH.exists() and all(TORCH_EXT_PATH.stat().st_mtime >= d.stat().st_mtime for d in deps):
        return TORCH_EXT_PATH
    build_library()
    cxx = os.environ.get("CXX") or shutil.which("g++") or shutil.which("c++")
    if not cxx:
        raise RuntimeError("no C++ compiler for the PyTorch extension")
    cuda_home = Path(find_nvcc()).resolve().parents[1]
    inc = [f"-I{p}" for p in ce.include_paths()] + [f"-I{cuda_home / 'include'}"]
    libdir = ce.library_paths()[0]
    cmd = [cxx, "-O2", "-std=c++17", "-shared", "-fPIC", f"-D_GLIBCXX_USE_CXX11_ABI={int(torch._C._GLIBCXX_USE_CXX11_ABI)}",
           *inc, str(src), "-o", str(TORCH_EXT_PATH), f"-L{PKG_DIR}", f"-l:{LIB_NAME}", f"-L{libdir}",
           "-ltorch", "-ltorch_cpu", "-lc10", "-ltorch_cuda", "-lc10_cuda", "-Wl,-rpath,$ORIGIN", f"-Wl,-rpath,{libdir}"]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError(f"{cxx} failed:\n{res.stdout}\n{res.stderr}")
    return TORCH_EXT_PATH


if __name__ == "__main__":
    print(build_library(force=True, verbose=True))
    print(build_pyobj(force=True))
    print(build_torch_ext(force=True))
