"""Builds pilottai_b200/libpilott_b200.so in-tree with nvcc for sm_100a.

No torch involvement: the library is plain CUDA runtime + driver entry points behind the C ABI
in include/pilott_b200.h.  `python -m pilottai_b200.build [--force]`.
"""
from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "libpilott_b200.so")
# host-side helper (CPython extension, plain C): task dicts -> routing kernel columns
TASKENC_SRC = os.path.join(CSRC, "task_encode.c")
TASKENC_OUT = os.path.join(HERE, "_taskenc" + (__import__("sysconfig").get_config_var("EXT_SUFFIX") or ".so"))
SOURCES = ["abi.cu", "store_kernels.cu", "scan_generic.cu", "scan_tc.cu", "finish.cu", "route.cu", "route_tc.cu", "xchg.cu"]
HEADERS = ["common.cuh", "internal.h", "tc_ptx.cuh", os.path.join("..", "..", "include", "pilott_b200.h")]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
    "-Xcompiler", "-fPIC",
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.sep not in cand or os.path.exists(cand)):
            return cand
    raise RuntimeError("nvcc not found")


def needs_build() -> bool:
    if not os.path.exists(OUT):
        return True
    t = os.path.getmtime(OUT)
    deps = [os.path.join(CSRC, f) for f in SOURCES + HEADERS] + [os.path.abspath(__file__)]
    return any(os.path.getmtime(d) > t for d in deps)


def build_taskenc(force: bool = False) -> str:
    import sysconfig
    if not force and os.path.exists(TASKENC_OUT) and os.path.getmtime(TASKENC_OUT) >= os.path.getmtime(TASKENC_SRC):
        return TASKENC_OUT
    cmd = [os.environ.get("CC", "gcc"), "-O2", "-fPIC", "-shared", "-Wall", "-I", sysconfig.get_paths()["include"],
           TASKENC_SRC, "-o", TASKENC_OUT]
    subprocess.check_call(cmd)
    return TASKENC_OUT


def build(force: bool = False, verbose: bool = False) -> str:
    build_taskenc(force)
    if not force and not needs_build():
        return OUT
    objs = []
    flags = list(NVCC_FLAGS)
    if os.environ.get("PILOTT_DEBUG") == "1":          # developer build: scan-kernel experiment switches and counters
        flags.append("-DPILOTT_DEBUG")
    build_dir = os.path.join(HERE, "build")
    os.makedirs(build_dir, exist_ok=True)
    procs = []
    for src in SOURCES:
        obj = os.path.join(build_dir, src.replace(".cu", ".o"))
        cmd = [_nvcc(), *flags, "-c", os.path.join(CSRC, src), "-o", obj]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
        objs.append(obj)
    for src, p in procs:
        out, _ = p.communicate()
        if verbose or p.returncode:
            sys.stderr.write(out)
        if p.returncode:
            raise RuntimeError(f"nvcc failed on {src}")
    link = [_nvcc(), "-shared", "-o", OUT, *objs, "-gencode", "arch=compute_100a,code=sm_100a",
            "-Xcompiler", "-fPIC", "-cudart", "static"]
    subprocess.check_call(link)
    return OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
