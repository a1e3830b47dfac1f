"""Builds libspecviterbi_b200.so (hand-written sm_100a CUDA + C ABI) in-tree with nvcc.

Usage: python -m spec_viterbi_b200.build [--force] [--verbose]
The library is git-ignored but travels to the GPU box with the repo snapshot.
"""
from __future__ import annotations

import os
import subprocess
import sys
from pathlib import Path

PKG = Path(__file__).resolve().parent
ROOT = PKG.parent
LIB = PKG / "libspecviterbi_b200.so"

PROBE_LIB = PKG / "libsvprobe_b200.so"  # micro-benchmarks: a separate library, never loaded by the product

SOURCES = [
    "csrc/sv_api.cu",
    "csrc/sv_multi.cu",
    "csrc/sv_plain_generic.cu",
    "csrc/sv_noalloc.cu",
    "csrc/sv_plain_es.cu",
    "csrc/sv_plain_ls.cu",
    "csrc/sv_plain_chain.cu",
    "csrc/sv_spec.cu",
    "csrc/sv_spec_chain.cu",
    "csrc/sv_es_plan.cpp",
    "csrc/sv_ls_plan.cpp",
    "csrc/sv_chain_plan.cpp",
    "csrc/sv_plan_cache.cpp",
    "csrc/sv_io.cpp",
    "host/data_reader.cpp",
]
PROBE_SOURCES = ["csrc/sv_probe.cu"]
HEADERS = ["csrc/sv_internal.h", "csrc/sv_ls_device.cuh", "host/data_reader.h", "host/HMM.h",
           "../include/spec_viterbi_b200.h", "../include/sv_probe_b200.h"]

NVCC_FLAGS = [
    "-O3", "-std=c++17",
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo",
    "-Xcompiler", "-fPIC,-O2,-Wall",
    "--threads", "0",
    # no fast-math anywhere: the add order and IEEE adds are the parity contract
]


def nvcc() -> str:
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.isabs(cand) and os.path.exists(cand) or not os.path.isabs(cand)):
            return cand
    raise RuntimeError("nvcc not found")


def needs_build() -> bool:
    if not LIB.exists() or not PROBE_LIB.exists():
        return True
    t = min(LIB.stat().st_mtime, PROBE_LIB.stat().st_mtime)
    return any((PKG / s).stat().st_mtime > t for s in SOURCES + PROBE_SOURCES + HEADERS if (PKG / s).exists())


def build(force: bool = False, verbose: bool = False) -> Path:
    if not force and not needs_build():
        return LIB
    objs = []
    objdir = PKG / "build"
    objdir.mkdir(exist_ok=True)
    procs = []
    for s in SOURCES + PROBE_SOURCES:
        obj = objdir / (Path(s).stem + ".o")
        cmd = [nvcc(), *NVCC_FLAGS, "-I", str(ROOT / "include"), "-I", str(PKG / "csrc"),
               "-c", str(PKG / s), "-o", str(obj)]
        if verbose:
            cmd += ["-Xptxas", "-v"]
        cmd += os.environ.get("SV_NVCC_EXTRA", "").split()  # e.g. -DSV_LS_DEBUG (instrumented build)
        procs.append((s, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
        if s in SOURCES:
            objs.append(str(obj))
    failed = False
    for s, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0 or verbose:
            sys.stderr.write(f"--- {s}\n{out}\n")
        failed |= p.returncode != 0
    if failed:
        raise RuntimeError("nvcc failed")
    cmd = [nvcc(), "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", str(LIB), *objs, "-lcudart",
           "-Xcompiler", "-pthread"]
    subprocess.run(cmd, check=True)
    cmd = [nvcc(), "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", str(PROBE_LIB),
           *[str(objdir / (Path(s).stem + ".o")) for s in PROBE_SOURCES], "-lcudart"]
    subprocess.run(cmd, check=True)
    return LIB


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose="--verbose" in sys.argv)
    print(LIB)
