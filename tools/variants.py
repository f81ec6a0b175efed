#!/usr/bin/env python
"""Builds tuning variants of librayca_cuda.so (extra -D defines) into rayca_b200/variants/ for A/B runs
on the GPU box: RAYCA_CUDA_LIB=rayca_b200/variants/<name>.so python bench.py ...

    python tools/variants.py name1:DEF1,DEF2=3 name2:DEF ...
"""
import os
import sys
from concurrent.futures import ThreadPoolExecutor

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from rayca_b200 import build as B  # noqa: E402


def one(spec):
    name, _, defs = spec.partition(":")
    out = os.path.join(ROOT, "rayca_b200", "variants", name + ".so")
    os.makedirs(os.path.dirname(out), exist_ok=True)
    B.build(force=True, out=out, defines=[d for d in defs.split(",") if d] + ["RAYCA_FAST_BUILD"])
    return out


if __name__ == "__main__":
    with ThreadPoolExecutor(max_workers=6) as ex:
        for o in ex.map(one, sys.argv[1:]):
            print(o)
