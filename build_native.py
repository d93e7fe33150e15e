#!/usr/bin/env python
"""Build libtadnn.so (every csrc/*.cu) for sm_100a with plain nvcc, in-tree.

The shared object lands in  torch-automatic-distributed-neural-network_b200/lib/libtadnn.so
so it travels with the repo snapshot to the GPU box (built artefacts are git-ignored,
not gpurun-ignored).  Incremental: a source is recompiled only if it or a header is
newer than its object file.

    python build_native.py            # build
    python build_native.py --force    # rebuild everything
    python build_native.py --sass     # also dump SASS listings under profiles/sass/
"""
from __future__ import annotations

import argparse
import concurrent.futures as cf
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "torch-automatic-distributed-neural-network_b200")
CSRC = os.path.join(ROOT, "csrc")
OBJ = os.path.join(ROOT, "build", "obj")
LIBDIR = os.path.join(PKG, "lib")
LIB = os.path.join(LIBDIR, "libtadnn.so")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17",
    "-Xcompiler", "-fPIC,-Wall,-Wno-unused-function",
    "--expt-relaxed-constexpr",
    "-Xptxas", "-v",
    "-I", os.path.join(ROOT, "include"),
]


def nvcc() -> str:
    exe = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(exe):
        raise RuntimeError("nvcc not found")
    return exe


def _newest_header() -> float:
    t = 0.0
    for d in (CSRC, os.path.join(ROOT, "include")):
        for f in os.listdir(d):
            if f.endswith((".cuh", ".h", ".hpp")):
                t = max(t, os.path.getmtime(os.path.join(d, f)))
    return t


def _compile(src: str, obj: str) -> tuple[str, int, str]:
    cmd = [nvcc(), *NVCC_FLAGS, "-c", src, "-o", obj]
    p = subprocess.run(cmd, capture_output=True, text=True)
    return src, p.returncode, p.stdout + p.stderr


def build(force: bool = False, verbose: bool = False) -> str:
    os.makedirs(OBJ, exist_ok=True)
    os.makedirs(LIBDIR, exist_ok=True)
    srcs = sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(".cu"))
    hdr_t = max(_newest_header(), os.path.getmtime(os.path.abspath(__file__)))
    jobs, objs = [], []
    for s in srcs:
        o = os.path.join(OBJ, os.path.basename(s)[:-3] + ".o")
        objs.append(o)
        if force or not os.path.exists(o) or os.path.getmtime(o) < max(os.path.getmtime(s), hdr_t):
            jobs.append((s, o))
    log_dir = os.path.join(ROOT, "build", "logs")
    os.makedirs(log_dir, exist_ok=True)
    if jobs:
        with cf.ThreadPoolExecutor(max_workers=min(8, len(jobs))) as ex:
            for src, rc, out in ex.map(lambda so: _compile(*so), jobs):
                with open(os.path.join(log_dir, os.path.basename(src) + ".log"), "w") as f:
                    f.write(out)
                if verbose or rc != 0:
                    sys.stderr.write(out)
                if rc != 0:
                    raise RuntimeError(f"nvcc failed on {src}")
    if jobs or not os.path.exists(LIB):
        cmd = [nvcc(), "-shared", "-o", LIB, *objs, "-lcudart"]
        p = subprocess.run(cmd, capture_output=True, text=True)
        if p.returncode != 0:
            sys.stderr.write(p.stdout + p.stderr)
            raise RuntimeError("link failed")
    return LIB


def dump_sass() -> None:
    """cuobjdump -sass of every object under profiles/sass/.  The GEMM object holds 24 template
    instantiations (24 MB of listing); only two representative ones are kept in the tree."""
    out_dir = os.path.join(ROOT, "profiles", "sass")
    os.makedirs(out_dir, exist_ok=True)
    # forward (TN; also the reduce-scatter / all-gather push epilogues, which are run-time branches of it), dgrad with
    # fused tanh' (the K13 producer), wgrad (MN-major operands; the K13 consumer with the flag-watcher warp)
    keep_gemm = ("ILb0ELb0ELi256ELb0E", "ILb0ELb1ELi256ELb1E", "ILb1ELb1ELi256ELb0E")
    for f in sorted(os.listdir(OBJ)):
        if not f.endswith(".o"):
            continue
        p = subprocess.run(["cuobjdump", "-sass", os.path.join(OBJ, f)], capture_output=True, text=True)
        if p.returncode != 0 or not p.stdout.strip():
            continue
        text = p.stdout
        if f.startswith("gemm_tcgen05"):
            parts = text.split("\t\tFunction : ")
            head, funcs = parts[0], parts[1:]
            kept = [fn for fn in funcs if any(k in fn.split("\n", 1)[0] for k in keep_gemm)]
            text = head + "".join("\t\tFunction : " + fn for fn in kept)
            text += f"\n// ({len(funcs) - len(kept)} further instantiations of gemm_bf16_tcgen05_kernel omitted)\n"
        with open(os.path.join(out_dir, f[:-2] + ".sass"), "w") as fh:
            fh.write(text)


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--force", action="store_true")
    ap.add_argument("--verbose", action="store_true")
    ap.add_argument("--sass", action="store_true")
    a = ap.parse_args()
    print(build(a.force, a.verbose))
    if a.sass:
        dump_sass()
