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
SOURCES = ["kp_api.cu", "kp_snapshot.cu", "kp_greedy.cu", "kp_greedy_fast.cu", "kp_sort.cu", "kp_ranksort.cu", "kp_ingest.cu", "kp_cachekey.cu", "kp_diag.cu", "kp_exchange.cu", "kp_decide.cu"]
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


def _digest(paths, extra: str = "") -> str:
    """Content hash of a target's inputs.  A built file is current iff the hash stored next to it matches: file
    times say nothing after the tree has been copied to another machine (the GPU box), where a time-based check
    made every rank of a multi-process run rebuild -- and overwrite -- the library at the same moment."""
    import hashlib
    h = hashlib.sha256(extra.encode())
    for p in sorted(paths):
        h.update(p.name.encode())
        h.update(p.read_bytes())
    return h.hexdigest()


def _current(target: Path, digest: str) -> bool:
    stamp = target.with_name(target.name + ".srchash")
    return target.exists() and stamp.exists() and stamp.read_text().strip() == digest


def _stamp(target: Path, digest: str) -> None:
    stamp = target.with_name(target.name + ".srchash")
    tmp = stamp.with_name(stamp.name + f".{os.getpid()}")
    tmp.write_text(digest + "\n")
    os.replace(tmp, stamp)


class _BuildLock:
    """One builder at a time per tree (several ranks may start together); the others wait and then find the
    target current."""

    def __enter__(self):
        import fcntl
        self.f = open(PKG_DIR / ".build.lock", "w")
        fcntl.flock(self.f, fcntl.LOCK_EX)
        return self

    def __exit__(self, *exc):
        import fcntl
        fcntl.flock(self.f, fcntl.LOCK_UN)
        self.f.close()


def _library_digest() -> str:
    deps = list(CSRC.glob("*.cu")) + list(CSRC.glob("*.cuh")) + [PKG_DIR.parent / "include" / "k8s_place_b200.h"]
    return _digest(deps, " ".join(NVCC_FLAGS + SOURCES))


def build_library(force: bool = False, verbose: bool = False) -> Path:
    """Compile the CUDA library if it is missing or was built from other sources (content hash, see _digest)."""
    extra = os.environ.get("KP_NVCC_EXTRA", "").split()          # experiments only (e.g. -DKP_UPDATE_MODE=1)
    out = Path(os.environ.get("KP_LIB_OUT", str(LIB_PATH)))
    variant = bool(extra) or out != LIB_PATH
    digest = _library_digest()
    if not force and not variant and _current(LIB_PATH, digest):
        return LIB_PATH
    with _BuildLock():
        if not force and not variant and _current(LIB_PATH, digest):     # another process built it while we waited
            return LIB_PATH
        tmp = out.with_name(out.name + f".{os.getpid()}.tmp")
        cmd = [find_nvcc(), *NVCC_FLAGS, *extra, "-o", str(tmp), *[str(CSRC / s) for s in SOURCES]]
        if verbose:
            cmd.insert(1, "-Xptxas")
            cmd.insert(2, "-v")
            print(" ".join(cmd))
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            tmp.unlink(missing_ok=True)
            raise RuntimeError(f"nvcc failed:\n{res.stdout}\n{res.stderr}")
        os.replace(tmp, out)                                          # readers never see a half-written library
        if not variant:
            _stamp(LIB_PATH, digest)
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
    digest = _digest([src], sysconfig.get_config_var("EXT_SUFFIX") or "")
    if not force and _current(out, digest):
        return out
    cc = os.environ.get("CC") or shutil.which("gcc") or shutil.which("cc")
    if not cc:
        raise RuntimeError("no C compiler for the record-walker extension")
    with _BuildLock():
        if not force and _current(out, digest):
            return out
        tmp = out.with_name(out.name + f".{os.getpid()}.tmp")
        cmd = [cc, "-O2", "-shared", "-fPIC", "-I", sysconfig.get_paths()["include"], str(src), "-o", str(tmp), "-lm"]
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            tmp.unlink(missing_ok=True)
            raise RuntimeError(f"{cc} failed:\n{res.stdout}\n{res.stderr}")
        os.replace(tmp, out)
        _stamp(out, digest)
    return out


TORCH_EXT_NAME = "libk8s_place_torch.so"
TORCH_EXT_PATH = PKG_DIR / TORCH_EXT_NAME


def build_torch_ext(force: bool = False) -> Path:
    """g++ build of the PyTorch-extension form of the ops (csrc/kp_torch_ext.cpp: TORCH_LIBRARY(k8s_place, ...)),
    in-tree and linked against libk8s_place_b200.so next to it (rpath $ORIGIN).  No nvcc: the file holds no kernels."""
    import torch
    from torch.utils import cpp_extension as ce
    src = CSRC / "kp_torch_ext.cpp"
    digest = _digest([src, PKG_DIR.parent / "include" / "k8s_place_b200.h"], torch.__version__)
    if not force and _current(TORCH_EXT_PATH, digest):
        return TORCH_EXT_PATH
    build_library()
    cxx = os.environ.get("CXX") or shutil.which("g++") or shutil.which("c++")
    if not cxx:
        raise RuntimeError("no C++ compiler for the PyTorch extension")
    cuda_home = Path(find_nvcc()).resolve().parents[1]
    inc = [f"-I{p}" for p in ce.include_paths()] + [f"-I{cuda_home / 'include'}"]
    libdir = ce.library_paths()[0]
    with _BuildLock():
        if not force and _current(TORCH_EXT_PATH, digest):
            return TORCH_EXT_PATH
        tmp = TORCH_EXT_PATH.with_name(TORCH_EXT_NAME + f".{os.getpid()}.tmp")
        cmd = [cxx, "-O2", "-std=c++17", "-shared", "-fPIC", f"-D_GLIBCXX_USE_CXX11_ABI={int(torch._C._GLIBCXX_USE_CXX11_ABI)}",
               *inc, str(src), "-o", str(tmp), f"-L{PKG_DIR}", f"-l:{LIB_NAME}", f"-L{libdir}",
               "-ltorch", "-ltorch_cpu", "-lc10", "-ltorch_cuda", "-lc10_cuda", "-Wl,-rpath,$ORIGIN", f"-Wl,-rpath,{libdir}"]
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            tmp.unlink(missing_ok=True)
            raise RuntimeError(f"{cxx} failed:\n{res.stdout}\n{res.stderr}")
        os.replace(tmp, TORCH_EXT_PATH)
        _stamp(TORCH_EXT_PATH, digest)
    return TORCH_EXT_PATH


if __name__ == "__main__":
    print(build_library(force=True, verbose=True))
    print(build_pyobj(force=True))
    print(build_torch_ext(force=True))
