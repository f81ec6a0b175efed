"""Builds rayca_b200/librayca_cuda.so in-tree with nvcc for sm_100a.

  --fmad=false            device code must not contract a*b+c (bit parity with the Rust reference,
                          which never fuses; the explicit fmaf() calls still emit FFMA)
  -Xcompiler -ffp-contract=off   same rule for the host flattening / BVH code
  -lineinfo               so ncu's source page maps to these files
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "librayca_cuda.so")
SOURCES = ["rayca_cuda.cu", "host_scene.cpp", "bvh_build.cpp", "fast_bvh.cpp", "gpu_fast_build.cu", "png_encode.cpp", "wavefront.cu"]
HEADERS = ["rmath.h", "gpu_types.h", "host_scene.h", "gpu_fast_build.h", "trace.cuh", "shade.cuh", "render.cuh", "wavefront.h",
           os.path.join("..", "..", "include", "rayca_cuda.h")]


def nvcc_path() -> str:
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", shutil.which("nvcc")):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found")


def is_stale() -> bool:
    if not os.path.exists(OUT):
        return True
    t = os.path.getmtime(OUT)
    return any(os.path.getmtime(os.path.join(CSRC, f)) > t for f in SOURCES + HEADERS) or \
        os.path.getmtime(os.path.abspath(__file__)) > t


def build(force: bool = False, verbose: bool = False, out: str = OUT, defines=()) -> str:
    """`out` / `defines` build a tuning variant next to the product library (see tools/variants.py)."""
    if out == OUT and not force and not is_stale():
        return OUT
    host_cxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
    common = [nvcc_path(), "-std=c++17", "-O3", "-lineinfo",
              "-gencode", "arch=compute_100a,code=sm_100a", "--fmad=false",
              "-ccbin", host_cxx, "-x", "cu",
              "-Xcompiler", "-fPIC,-ffp-contract=off,-fno-fast-math,-fopenmp,-O2",
              "-Xptxas", "-v" if verbose else "-O3"]
    common += [f"-D{d}" for d in defines]
    # one nvcc per source, side by side (the two kernel translation units take over a minute each), then one link
    import tempfile
    from concurrent.futures import ThreadPoolExecutor

    # objects are kept under .objcache/ keyed by the compile line, the source and every header: a change to one
    # translation unit recompiles that unit only (rayca_cuda.cu alone takes over two minutes)
    import hashlib
    import types
    cache = os.path.join(os.path.dirname(os.path.abspath(__file__)), ".objcache")
    os.makedirs(cache, exist_ok=True)
    header_digest = hashlib.sha1()
    for h in sorted(HEADERS):
        with open(os.path.join(CSRC, h), "rb") as fh:
            header_digest.update(h.encode() + b"\0" + fh.read())

    with tempfile.TemporaryDirectory(prefix="rayca_build_") as tmp:
        def compile_one(src):
            with open(os.path.join(CSRC, src), "rb") as fh:
                key = hashlib.sha1(" ".join(common).encode() + b"\0" + header_digest.digest() + fh.read()).hexdigest()
            obj = os.path.join(tmp if defines else cache, f"{os.path.splitext(src)[0]}-{key}.o")  # variants are not cached
            if os.path.exists(obj) and not verbose:
                return src, obj, types.SimpleNamespace(returncode=0, stdout="", stderr="")
            part = os.path.join(tmp, os.path.splitext(src)[0] + ".o")
            res = subprocess.run(common + ["-c", os.path.join(CSRC, src), "-o", part], capture_output=True, text=True)
            if res.returncode == 0:
                shutil.move(part, obj)
                stem = os.path.splitext(src)[0] + "-"
                for old in os.listdir(cache):  # one object per source and flag set would pile up 15 MB at a time
                    if not defines and old.startswith(stem) and os.path.join(cache, old) != obj:
                        os.remove(os.path.join(cache, old))
            return src, obj, res

        with ThreadPoolExecutor(max_workers=len(SOURCES)) as ex:
            results = list(ex.map(compile_one, SOURCES))
        log = "".join(r.stdout + r.stderr for _, _, r in results)
        if any(r.returncode != 0 for _, _, r in results):
            sys.stderr.write(log)
            raise RuntimeError("nvcc failed")
        link = [nvcc_path(), "-shared", "-o", out, "-ccbin", host_cxx, "-gencode", "arch=compute_100a,code=sm_100a"]
        link += [obj for _, obj, _ in results] + ["-lgomp", "-lz"]  # zlib: deflate + crc32 of the PNG encoder (png_encode.cpp)
        res = subprocess.run(link, capture_output=True, text=True)
        log += res.stdout + res.stderr
        if res.returncode != 0:
            sys.stderr.write(log)
            raise RuntimeError("nvcc link failed")
    if verbose:
        sys.stderr.write(log)
    return out


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
