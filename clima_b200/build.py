"""Builds clima_b200/libclima_rt.so (the C-ABI library of include/clima_rt.h) with nvcc for sm_100a.

No torch, no JIT cache: one in-tree shared object that travels with the repo snapshot."""
from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libclima_rt.so")
SOURCES = ["capi.cu", "prep.cu", "optics_rorr.cu", "optics_rorr_fast.cu", "optics_rorr_exact.cu", "optics_rorr_cut.cu", "jacobian_delta.cu", "adiabat_profile.cu", "twostream.cu", "fp64_peak.cu"]
HEADERS = ["rt_device.cuh", "rt_kernels.h", "rorr_impl.cuh", "dop853_tables.h", "adiabat_group.cuh", os.path.join("..", "..", "include", "clima_rt.h")]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
         "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr", "--extended-lambda"]


def _stale() -> bool:
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, s) for s in SOURCES + HEADERS] + [os.path.abspath(__file__)]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and not _stale():
        return LIB
    objs = []
    procs = []
    for s in SOURCES:
        o = os.path.join(CSRC, s.replace(".cu", ".o"))
        cmd = [NVCC] + FLAGS + os.environ.get("CLIMA_RT_NVCC_EXTRA", "").split() + (["-Xptxas", "-v"] if verbose else []) + ["-c", os.path.join(CSRC, s), "-o", o]
        procs.append((s, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
        objs.append(o)
    for s, p in procs:
        out, _ = p.communicate()
        if verbose or p.returncode != 0:
            sys.stderr.write(f"--- {s}\n{out}\n")
        if p.returncode != 0:
            raise RuntimeError(f"nvcc failed on {s}")
    cmd = [NVCC, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", LIB] + objs + ["-lcudart"]
    subprocess.run(cmd, check=True)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
