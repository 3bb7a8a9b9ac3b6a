"""Build librf_b200.so in-tree with nvcc for sm_100a.

One translation unit per (dtype, N) tile-kernel instantiation plus the API unit,
compiled in parallel; objects are cached by source mtime under csrc/build/.
Run as ``python -m randomfield_b200.build`` or through ``__graft_entry__.build()``.
"""
from __future__ import annotations

import concurrent.futures
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
BUILD = os.path.join(CSRC, "build")
LIB = os.path.join(HERE, "librf_b200.so")

SIZES = (64, 128, 256, 512, 1024, 2048, 4096)
PIPE_SIZES = (64, 128, 256, 512, 1024, 2048, 4096)
DTYPES = (("f32", "float"), ("f64", "double"))

NVCC_FLAGS = [
    "-std=c++17", "-O3", "-lineinfo",
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-Xcompiler", "-fPIC",
]


def _nvcc():
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found")


def _headers_mtime():
    m = 0.0
    for name in os.listdir(CSRC):
        if name.endswith((".cuh", ".h")):
            m = max(m, os.path.getmtime(os.path.join(CSRC, name)))
    inc = os.path.join(os.path.dirname(HERE), "include", "rf_b200.h")
    if os.path.exists(inc):
        m = max(m, os.path.getmtime(inc))
    return m


def _compile(job):
    src, obj, defs, verbose = job
    cmd = [_nvcc(), *NVCC_FLAGS, *defs, "-c", src, "-o", obj]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed: %s\n%s\n%s" % (" ".join(cmd), r.stdout, r.stderr))
    return obj, r.stderr


def build(force: bool = False, verbose: bool = False, jobs: int | None = None) -> str:
    os.makedirs(BUILD, exist_ok=True)
    hm = _headers_mtime()
    todo, objs = [], []

    def want(src, obj):
        if force or not os.path.exists(obj):
            return True
        return os.path.getmtime(obj) < max(hm, os.path.getmtime(src))

    api_src = os.path.join(CSRC, "rf_api.cu")
    api_obj = os.path.join(BUILD, "rf_api.o")
    objs.append(api_obj)
    if want(api_src, api_obj):
        todo.append((api_src, api_obj, [], verbose))
    inst_src = os.path.join(CSRC, "rf_inst.cu")
    for tag, ctype in DTYPES:
        for n in SIZES:
            obj = os.path.join(BUILD, "inst_%s_%d.o" % (tag, n))
            objs.append(obj)
            if want(inst_src, obj):
                todo.append((inst_src, obj, ["-DRF_INST_T=%s" % ctype, "-DRF_INST_N=%d" % n], verbose))
    pipe_src = os.path.join(CSRC, "rf_pipe_inst.cu")
    for n in PIPE_SIZES:
        obj = os.path.join(BUILD, "pipe_%d.o" % n)
        objs.append(obj)
        if want(pipe_src, obj):
            todo.append((pipe_src, obj, ["-DRF_PIPE_N=%d" % n], verbose))
    fused_src = os.path.join(CSRC, "rf_fused_inst.cu")
    for n in PIPE_SIZES:
        if n > 1024:
            continue
        obj = os.path.join(BUILD, "fused_%d.o" % n)
        objs.append(obj)
        if want(fused_src, obj):
            todo.append((fused_src, obj, ["-DRF_PIPE_N=%d" % n], verbose))
    if todo:
        nj = jobs or min(len(todo), os.cpu_count() or 4)
        with concurrent.futures.ThreadPoolExecutor(nj) as ex:
            for obj, log in ex.map(_compile, todo):
                if verbose and log:
                    sys.stderr.write("== %s\n%s\n" % (os.path.basename(obj), log))
    if todo or not os.path.exists(LIB):
        cmd = [_nvcc(), "-shared", "-o", LIB, *objs, "-gencode", "arch=compute_100a,code=sm_100a",
               "-Xcompiler", "-fPIC", "-lcudart"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("link failed:\n%s\n%s" % (r.stdout, r.stderr))
    return LIB


if __name__ == "__main__":
    path = build(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print(path)
