"""Build libgcta_b200.so (C ABI, hand-written sm_100a CUDA) in-tree with nvcc.

    python -m gcta_b200.build            # or: from gcta_b200.build import build; build()
nvcc cross-compiles without a GPU.  The .so is git-ignored but travels with gpurun snapshots.
"""
from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libgcta_b200.so")
SOURCES = ["decode.cu", "decode2.cu", "nonmiss.cu", "syrk.cu", "syrk2.cu", "paircount.cu", "finalize.cu", "plan.cu", "exchange.cu", "multi.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", "--fmad=true"]


def _stale(target: str, deps) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build(verbose: bool = False, force: bool = False) -> str:
    nvcc = os.environ.get("NVCC", "nvcc")
    objdir = os.path.join(HERE, "build")
    os.makedirs(objdir, exist_ok=True)
    hdrs = [os.path.join(CSRC, "common.cuh"), os.path.join(HERE, "..", "include", "gcta_b200.h")]
    objs = []
    procs = []
    for s in SOURCES:
        src = os.path.join(CSRC, s)
        obj = os.path.join(objdir, s.replace(".cu", ".o"))
        objs.append(obj)
        if force or _stale(obj, [src] + hdrs):
            cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-c", src, "-o", obj]
            procs.append((s, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    for s, pr in procs:
        out, _ = pr.communicate()
        if verbose or pr.returncode:
            print(out)
        if pr.returncode:
            raise RuntimeError(f"nvcc failed on {s}")
    if force or procs or _stale(LIB, objs):
        cmd = [nvcc, "-shared", "-o", LIB] + objs + ["-gencode", "arch=compute_100a,code=sm_100a", "-cudart", "static"]
        subprocess.run(cmd, check=True)
    build_host(force=force)
    return LIB


HOST_BIN = os.path.join(HERE, "gcta64_b200")
HOST_SOURCES = ["main.cpp", "gcta_host.cpp", "gcta_host_grm.cpp", "gcta_host_grmio.cpp", "grm_text_gz.cpp"]


def build_host(force: bool = False) -> str:
    """C++ front end (same CLI as the reference for the --make-grm path), linked against the C ABI."""
    hdir = os.path.join(HERE, "host")
    srcs = [os.path.join(hdir, s) for s in HOST_SOURCES]
    deps = srcs + [os.path.join(hdir, "gcta_host.h"), os.path.join(hdir, "grm_text_gz.h"), os.path.join(hdir, "grm_bin_writer.h"), os.path.join(hdir, "gz_member_writer.h"), os.path.join(hdir, "big_alloc.h"), os.path.join(HERE, "..", "include", "gcta_b200.h"), LIB]
    if force or _stale(HOST_BIN, deps):
        cxx = os.environ.get("CXX", "g++")
        cmd = [cxx, "-O2", "-std=c++17", "-pthread", "-o", HOST_BIN] + srcs + \
              ["-L" + HERE, "-lgcta_b200", "-lz", "-Wl,-rpath,$ORIGIN"]
        subprocess.run(cmd, check=True)
    return HOST_BIN


if __name__ == "__main__":
    print(build(verbose="-v" in sys.argv, force="-f" in sys.argv))
