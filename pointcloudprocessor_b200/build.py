"""Builds libpcp_b200.so (hand-written CUDA for sm_100a + the C ABI) in-tree with nvcc.

The library travels to the GPU box with the repo snapshot; nothing is JIT-compiled at run time.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

_HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(_HERE, "csrc")
LIB_PATH = os.path.join(_HERE, "libpcp_b200.so")
SOURCES = ["pcp_kernels.cu", "pcp_capi.cu", "pcp_frames.cu", "pcp_nid.cu", "pcp_nccl.cu", "pcp_prebin.cu", "pcp_image.cu", "pcp_jpeg.cu", "pcp_process.cu", "pcp_calib.cu", "pcp_splat.cu", "pcp_noalloc.cu"]
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]


def _nvcc() -> str:
    for cand in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: the CUDA library cannot be built (there is no CPU fallback)")


def _host_cxx() -> str:
    return shutil.which("g++") or "g++"


def _newer(target: str, deps) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build_library(force: bool = False, verbose: bool = False, defines=(), out: str | None = None) -> str:
    nvcc = _nvcc()
    srcs = [s for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".hpp", ".h"))]
    headers.append(os.path.join(_HERE, "..", "include", "pcp", "pcp.h"))
    lib_path = out or LIB_PATH
    tag = "" if not defines else "_" + "_".join(d.replace("=", "") for d in defines)
    objdir = os.path.join(_HERE, "build" + tag)
    os.makedirs(objdir, exist_ok=True)
    common = ["-O3", "-std=c++17", "-lineinfo", "-fmad=false", "-ccbin", _host_cxx(), "-Xcompiler", "-fPIC,-ffp-contract=off"] + ARCH
    common += [f"-D{d}" for d in defines]
    if verbose:
        common += ["-Xptxas", "-v"]
    objs = []
    rebuilt = False
    for s in srcs:
        src = os.path.join(CSRC, s)
        obj = os.path.join(objdir, s.replace(".cu", ".o"))
        if force or _newer(obj, [src] + headers):
            cmd = [nvcc, "-c", src, "-o", obj] + common
            r = subprocess.run(cmd, capture_output=True, text=True)
            if verbose or r.returncode != 0:
                sys.stderr.write(r.stdout + r.stderr)
            if r.returncode != 0:
                raise RuntimeError(f"nvcc failed on {s}")
            rebuilt = True
        objs.append(obj)
    if rebuilt or not os.path.exists(lib_path):
        cmd = [nvcc, "-shared", "-o", lib_path] + objs + ARCH + ["-ccbin", _host_cxx(), "-ldl", "-lz", "-Xlinker", "--no-undefined"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            sys.stderr.write(r.stdout + r.stderr)
            raise RuntimeError("link failed")
    if out is None:
        build_cli(lib_path)
    return lib_path


def build_cli(lib_path: str = LIB_PATH) -> str:
    """pcp_colorize: the reference's executable (src/main.cpp) over the C++ facade."""
    exe = os.path.join(_HERE, "pcp_colorize")
    src = os.path.join(CSRC, "pcp_main.cpp")
    deps = [src, os.path.join(_HERE, "..", "include", "pcp", "pcp.hpp"), os.path.join(_HERE, "..", "include", "pcp", "pcp.h"), lib_path]
    if _newer(exe, deps):
        cmd = [_host_cxx(), "-O2", "-std=c++17", "-Wall", "-I" + os.path.join(_HERE, "..", "include"), src, "-o", exe, "-L" + _HERE, "-lpcp_b200",
               "-Wl,-rpath,$ORIGIN"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            sys.stderr.write(r.stdout + r.stderr)
            raise RuntimeError("pcp_colorize: link failed")
    return exe


if __name__ == "__main__":
    _defs = [a[2:] for a in sys.argv[1:] if a.startswith("-D")]
    _out = next((a[6:] for a in sys.argv[1:] if a.startswith("--out=")), None)
    print(build_library(force="--force" in sys.argv, verbose="-v" in sys.argv, defines=_defs, out=_out))
