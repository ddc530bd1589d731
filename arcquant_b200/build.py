"""In-tree build of libarcq_b200.so (hand-written sm_100a CUDA + the C ABI of include/arcq.h).

``python -m arcquant_b200.build`` or ``arcquant_b200.build.build()``.  nvcc cross-compiles without a
GPU; the .so is git-ignored but travels to the GPU box with the gpurun snapshot.
"""
from __future__ import annotations

import hashlib
import os
import subprocess
import sys

_HERE = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_HERE)
CSRC = os.path.join(_HERE, "csrc")
INCLUDE = os.path.join(_ROOT, "include")
LIB = os.path.join(_HERE, "libarcq_b200.so")
EXT = os.path.join(_HERE, "_agemm_ext.so")          # compiled pybind11/torch module over the C ABI (csrc/agemm_ext.cpp)
EXT_NAME = "_agemm_ext"
# (tile rows R, rows per item RT, threads, CTAs per SM) of the TMA-staged quantise kernel: one object file each
QUANT_CFGS = [(4, 4, 256, 2), (4, 4, 512, 1), (2, 2, 256, 2), (2, 2, 512, 1), (1, 1, 512, 1), (8, 8, 512, 1), (8, 4, 512, 1)]
# source file, or (source file, extra nvcc flags, object name)
SOURCES = ["runtime.cu", "plan.cu", "quant.cu", "quant_tma.cu", "gemm.cu", "gemm_pair.cu", "gemm_splitk.cu", "noalloc.cu", "peer.cu"] + [
    ("quant_tma_cfg.cu", ["-DARCQ_CFG_R=%d" % r, "-DARCQ_CFG_RT=%d" % rt, "-DARCQ_CFG_TH=%d" % th, "-DARCQ_CFG_MB=%d" % mb],
     "quant_tma_cfg_%d_%d_%d_%d.o" % (r, rt, th, mb)) for r, rt, th, mb in QUANT_CFGS]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC", "-I" + INCLUDE, "-I" + CSRC,
]


def _nvcc() -> str:
    for c in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if c and (os.path.isabs(c) and os.path.exists(c) or not os.path.isabs(c)):
            return c
    return "nvcc"


def _fingerprint() -> str:
    h = hashlib.sha256()
    for root in (CSRC, INCLUDE):
        for name in sorted(os.listdir(root)):
            if name.endswith((".cu", ".cuh", ".h", ".cpp")):
                with open(os.path.join(root, name), "rb") as f:
                    h.update(name.encode() + b"\0" + f.read())
    h.update(" ".join(NVCC_FLAGS).encode() + repr(SOURCES).encode())
    return h.hexdigest()


def is_current() -> bool:
    stamp = LIB + ".stamp"
    if not (os.path.exists(LIB) and os.path.exists(EXT) and os.path.exists(stamp)):
        return False
    with open(stamp) as f:
        return f.read().strip() == _fingerprint()


def can_build() -> bool:
    """nvcc present and the sources writable (the GPU box runs the prebuilt .so files of the snapshot)."""
    import shutil
    return (os.path.exists("/usr/local/cuda/bin/nvcc") or shutil.which("nvcc") is not None) and os.access(_HERE, os.W_OK)


def build_ext(verbose: bool = False) -> str:
    """g++ the pybind11/torch module `_agemm_ext` (csrc/agemm_ext.cpp) against libarcq_b200.so ($ORIGIN rpath)."""
    import sysconfig
    import torch
    from torch.utils import cpp_extension as ce
    tlib = os.path.join(os.path.dirname(torch.__file__), "lib")
    inc = ["-I" + p for p in ce.include_paths()] + ["-I" + sysconfig.get_paths()["include"], "-I" + INCLUDE,
                                                     "-I/usr/local/cuda/include"]
    abi = int(getattr(torch._C, "_GLIBCXX_USE_CXX11_ABI", True))
    tmp = EXT + ".tmp%d" % os.getpid()
    cmd = ["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-fvisibility=hidden", "-DTORCH_EXTENSION_NAME=" + EXT_NAME,
           "-DTORCH_API_INCLUDE_EXTENSION_H", "-D_GLIBCXX_USE_CXX11_ABI=%d" % abi] + inc + [
           os.path.join(CSRC, "agemm_ext.cpp"), "-o", tmp, "-L" + _HERE, "-l:libarcq_b200.so", "-L" + tlib, "-ltorch_python",
           "-ltorch", "-ltorch_cpu", "-ltorch_cuda", "-lc10", "-lc10_cuda", "-Wl,-rpath,$ORIGIN", "-Wl,-rpath," + tlib]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if verbose or r.returncode:
        sys.stderr.write(r.stdout)
    if r.returncode:
        raise RuntimeError("g++ failed: " + " ".join(cmd))
    os.replace(tmp, EXT)
    return EXT


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile every .cu under csrc/ for sm_100a into one shared library (+ the pybind11 module); returns the library path.
    Serialised across processes with a file lock (concurrent ranks / xdist workers share arcquant_b200/build/*.o)."""
    import fcntl
    os.makedirs(os.path.join(_HERE, "build"), exist_ok=True)
    with open(os.path.join(_HERE, "build", ".lock"), "w") as lock:
        fcntl.flock(lock, fcntl.LOCK_EX)
        try:
            return _build_locked(force, verbose)
        finally:
            fcntl.flock(lock, fcntl.LOCK_UN)


def _build_locked(force: bool, verbose: bool) -> str:
    if not force and is_current():
        return LIB
    objs = []
    objdir = os.path.join(_HERE, "build")
    os.makedirs(objdir, exist_ok=True)
    procs = []
    for entry in SOURCES:
        src, extra, oname = entry if isinstance(entry, tuple) else (entry, [], entry.replace(".cu", ".o"))
        path = os.path.join(CSRC, src)
        if not os.path.exists(path):
            raise FileNotFoundError(path)
        obj = os.path.join(objdir, oname)
        cmd = [_nvcc()] + NVCC_FLAGS + extra + (["-Xptxas", "-v"] if verbose else []) + ["-c", path, "-o", obj]
        procs.append((cmd, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
        objs.append(obj)
    for cmd, p in procs:
        out, _ = p.communicate()
        if verbose or p.returncode:
            sys.stderr.write(out)
        if p.returncode:
            raise RuntimeError("nvcc failed: " + " ".join(cmd))
    tmp = LIB + ".tmp%d" % os.getpid()
    subprocess.check_call([_nvcc(), "-shared", "-Wno-deprecated-gpu-targets", "-o", tmp] + objs)
    os.replace(tmp, LIB)
    build_ext(verbose)
    with open(LIB + ".stamp", "w") as f:
        f.write(_fingerprint())
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
