"""Builds libwkv6_b200.so (hand-written sm_100a CUDA + the C ABI) and libwkv6_b200_torch.so
(the TORCH_LIBRARY(wkv6) shim) in-tree with nvcc / g++.  No JIT cache: the .so files travel with the repo."""
from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libwkv6_b200.so")
TRACE_LIB = os.path.join(HERE, "libwkv6_b200_trace.so")   # -DWKV6_TRACE_PHASES=1; loaded when WKV6_B200_LIB points at it
TORCH_LIB = os.path.join(HERE, "libwkv6_b200_torch.so")
CUDA_SOURCES = ["wkv6_scan.cu", "wkv6_chunk_lean.cu", "wkv6_chunk_bwd.cu", "wkv6_tmix.cu", "convmod.cu", "wkv6_capi.cu"]
HEADERS = ["wkv6_common.cuh", "wkv6_scan.cuh", "wkv6_umma.cuh", "wkv6_fwd_tiles.cuh", "../../include/wkv6_b200.h"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", "-Xptxas", "-v", "--expt-relaxed-constexpr"]


def _newer(target: str, deps) -> bool:
    if not os.path.exists(target):
        return False
    t = os.path.getmtime(target)
    return all(os.path.getmtime(d) <= t for d in deps)


def _nvcc() -> str:
    for c in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if c and (os.path.isabs(c) and os.path.exists(c) or not os.path.isabs(c)):
            return c
    return "nvcc"


def build_cuda(force: bool = False, verbose: bool = False, trace: bool = False) -> str:
    """trace=True builds libwkv6_b200_trace.so: the same sources with the in-loop clock64() phase stamps of the chunk kernels
    compiled in (scripts/trace_phases.py loads it through WKV6_B200_LIB); the shipped library is built without them."""
    deps = [os.path.join(CSRC, s) for s in CUDA_SOURCES + HEADERS]
    lib = TRACE_LIB if trace else LIB          # the trace build is a file of its own: it never replaces the shipped library
    if not force and _newer(lib, deps):
        return lib
    objs = []
    procs = []
    for s in CUDA_SOURCES:
        o = os.path.join(CSRC, s.replace(".cu", "_trace.o" if trace else ".o"))
        objs.append(o)
        cmd = [_nvcc()] + NVCC_FLAGS + (["-DWKV6_TRACE_PHASES=1"] if trace else []) + ["-c", os.path.join(CSRC, s), "-o", o]
        procs.append((s, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    log = []
    for s, pr in procs:
        out, _ = pr.communicate()
        log.append(f"==== {s}\n{out}")
        if pr.returncode != 0:
            raise RuntimeError(f"nvcc failed on {s}:\n{out}")
    if not trace:
        with open(os.path.join(CSRC, "ptxas_v.log"), "w") as f:
            f.write("\n".join(log))
    if verbose:
        print("\n".join(log))
    cmd = [_nvcc(), "-shared", "-o", lib] + objs + ["-gencode", "arch=compute_100a,code=sm_100a", "-lcudart"]
    subprocess.run(cmd, check=True)
    return lib


def build_torch_shim(force: bool = False) -> str:
    src = os.path.join(CSRC, "wkv6_torch.cpp")
    deps = [src, os.path.join(HERE, "..", "include", "wkv6_b200.h")]
    if not force and _newer(TORCH_LIB, deps):
        return TORCH_LIB
    import torch
    from torch.utils import cpp_extension as ce
    inc = []
    for p in ce.include_paths():
        inc += ["-isystem", p]
    tl = os.path.join(os.path.dirname(torch.__file__), "lib")
    abi = int(getattr(torch._C, "_GLIBCXX_USE_CXX11_ABI", True))
    cmd = ["g++", "-O2", "-std=c++17", "-fPIC", "-shared", f"-D_GLIBCXX_USE_CXX11_ABI={abi}",
           "-I/usr/local/cuda/include", *inc, src, "-o", TORCH_LIB,
           f"-L{tl}", "-lc10", "-ltorch_cpu", "-ltorch", "-lc10_cuda", "-ltorch_cuda",
           f"-L{HERE}", "-l:libwkv6_b200.so", "-Wl,-rpath,$ORIGIN", f"-Wl,-rpath,{tl}"]
    subprocess.run(cmd, check=True)
    return TORCH_LIB


def build_all(force: bool = False, verbose: bool = False, trace: bool = False):
    if trace:
        return build_cuda(force, verbose, True), None
    return build_cuda(force, verbose), build_torch_shim(force)


if __name__ == "__main__":
    build_all(force="--force" in sys.argv, verbose=True, trace="--trace" in sys.argv)
    print("built", LIB, TORCH_LIB)
