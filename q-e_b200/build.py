"""Builds q-e_b200/libqeb200.so (sm_100a) in-tree with nvcc. Used by
__graft_entry__.build(); the .so is git-ignored but travels with gpurun."""
from __future__ import annotations

import concurrent.futures as cf
import hashlib
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "libqeb200.so")
BUILD = os.path.join(HERE, "build")
SOURCES = ["context.cu", "operators.cu", "pipeline.cu", "graphs.cu", "multigpu.cu", "dav_blocks.cu", "vloc.cu", "gemm.cu", "fullfft.cu", "distfft.cu", "nccl_glue.cu", "descriptor.cu", "uspp.cu", "davidson.cu", "solver.cu", "static_scratch.cu", "density_nc.cu"] + [
    f"vloc_fast_{n}.cu" for n in (15, 25, 27, 64, 125, 135, 144, 200, "200b", "200c", "200d")] + [
    f"vloc_plane_{a}_{b}.cu" for a, b in ((125, 64), (25, 25), (27, 27), (15, 15), (64, 64), ("125b", 64), ("125b", "64b"))] + [
    f"vloc_plane2_{a}_{b}.cu" for a, b in ((125, 64), ("125b", 64), (25, 25), (27, 27), (15, 15), (64, 64))]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC", "-Xcompiler", "-O2", "--use_fast_math=false",
]
NVCC_FLAGS = [f for f in NVCC_FLAGS if f != "--use_fast_math=false"]


def _nvcc() -> str:
    for cand in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found")


def _digest(paths) -> str:
    h = hashlib.sha256()
    for p in sorted(paths):
        h.update(p.encode())
        h.update(open(p, "rb").read())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def _compile(nvcc, src, obj):
    cmd = [nvcc, *NVCC_FLAGS, "-I", CSRC, "-c", os.path.join(CSRC, src), "-o", obj]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
    return obj


def build_variant(tag: str, defines: list[str], verbose: bool = True) -> str:
    """Tuning aid: builds q-e_b200/libqeb200_<tag>.so with extra -D defines (select it with QEB_LIB=<path>)."""
    nvcc = _nvcc()
    bdir = os.path.join(BUILD, "variant_" + tag)
    os.makedirs(bdir, exist_ok=True)
    srcs = [s for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]
    flags = [*NVCC_FLAGS, *[f"-D{d}" for d in defines]]

    def one(src):
        obj = os.path.join(bdir, src.replace(".cu", ".o"))
        r = subprocess.run([nvcc, *flags, "-I", CSRC, "-c", os.path.join(CSRC, src), "-o", obj], capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
        return obj

    with cf.ThreadPoolExecutor(max_workers=8) as ex:
        objs = list(ex.map(one, srcs))
    out = os.path.join(HERE, f"libqeb200_{tag}.so")
    r = subprocess.run([nvcc, "-shared", "-o", out, *objs, "-gencode", "arch=compute_100a,code=sm_100a"], capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    if verbose:
        print("built", out)
    return out


def _deps(path, seen=None):
    """the file plus every `#include "..."` it reaches (resolved against its own directory and csrc/)"""
    import re

    seen = set() if seen is None else seen
    path = os.path.normpath(path)
    if path in seen or not os.path.exists(path):
        return seen
    seen.add(path)
    for inc in re.findall(r'^\s*#\s*include\s+"([^"]+)"', open(path).read(), flags=re.M):
        for base in (os.path.dirname(path), CSRC):
            cand = os.path.normpath(os.path.join(base, inc))
            if os.path.exists(cand):
                _deps(cand, seen)
                break
    return seen


def build(force: bool = False, verbose: bool = False) -> str:
    """Incremental: an object is recompiled when its source or a header it includes (transitively) changed."""
    srcs = [s for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]
    nvcc = _nvcc()
    os.makedirs(BUILD, exist_ok=True)
    todo, objs, digs = [], [], {}
    for s in srcs:
        obj = os.path.join(BUILD, s.replace(".cu", ".o"))
        dig = _digest(_deps(os.path.join(CSRC, s)))
        stamp = obj + ".stamp"
        objs.append(obj)
        digs[obj] = dig
        if force or not os.path.exists(obj) or not os.path.exists(stamp) or open(stamp).read() != dig:
            todo.append((s, obj))
    if not todo and os.path.exists(OUT) and os.path.getmtime(OUT) >= max(os.path.getmtime(o) for o in objs):
        return OUT
    with cf.ThreadPoolExecutor(max_workers=min(8, max(1, len(todo)))) as ex:
        futs = [ex.submit(_compile, nvcc, s, obj) for s, obj in todo]
        for f in futs:
            f.result()
    for _, obj in todo:
        open(obj + ".stamp", "w").write(digs[obj])
    cmd = [nvcc, "-shared", "-o", OUT, *objs, "-gencode", "arch=compute_100a,code=sm_100a"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    if verbose:
        print("built", OUT, f"({len(todo)} of {len(srcs)} objects recompiled)")
    return OUT


if __name__ == "__main__":
    if "--variant" in sys.argv:
        i = sys.argv.index("--variant")
        build_variant(sys.argv[i + 1], sys.argv[i + 2:])
    else:
        build(force="--force" in sys.argv, verbose=True)
