"""ctypes binding of oracle/_build/librayca_oracle.so (built by oracle/Makefile).

TEST INFRASTRUCTURE ONLY — see oracle/__init__.py.  The POD structs are shared with the product's
host mirror (rayca_b200.model) because both sides consume include/rayca_cuda.h.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

from rayca_b200 import model as M

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "librayca_oracle.so")

FAITHFUL = 1
COUNT = 2
OPTIMISED = 4  # ordered, culled walk over a binned-SAH tree of pre-transformed triangles ("optimised CPU" arm)


class rayca_oracle_stats(C.Structure):
    _fields_ = [("build_ms", C.c_double), ("trace_ms", C.c_double), ("rays_primary", C.c_uint64),
                ("rays_shadow", C.c_uint64), ("rays_secondary", C.c_uint64), ("ord_nodes", C.c_uint64),
                ("ord_tris", C.c_uint64), ("exh_nodes", C.c_uint64), ("exh_tris", C.c_uint64),
                ("threads", C.c_uint32), ("reserved", C.c_uint32), ("opt_build_ms", C.c_double)]


def build(force: bool = False) -> str:
    srcs = [os.path.join(_HERE, "rayca_oracle.cpp"), os.path.join(_HERE, "oracle_math.h"),
            os.path.join(_HERE, "..", "include", "rayca_cuda.h")]
    stale = force or not os.path.exists(_SO) or any(
        os.path.exists(s) and os.path.getmtime(s) > os.path.getmtime(_SO) for s in srcs)
    if stale:
        subprocess.check_call(["make", "-C", _HERE], stdout=subprocess.DEVNULL)
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(build())
        _lib.rayca_oracle_scene_create.argtypes = [C.c_void_p, C.c_uint32, C.POINTER(C.c_void_p)]
        _lib.rayca_oracle_scene_destroy.argtypes = [C.c_void_p]
        _lib.rayca_oracle_scene_destroy.restype = None
        _lib.rayca_oracle_scene_bvh_info.argtypes = [C.c_void_p, C.POINTER(M.rayca_bvh_info)]
        _lib.rayca_oracle_scene_bvh_read.argtypes = [C.c_void_p] + [C.c_void_p] * 4
        _lib.rayca_oracle_render.argtypes = [C.c_void_p, C.POINTER(M.rayca_config), C.POINTER(M.rayca_options),
                                             C.c_uint32, C.c_uint32, C.c_void_p, C.POINTER(M.rayca_aux_out),
                                             C.c_uint32, C.c_int, C.POINTER(C.c_uint32),
                                             C.POINTER(rayca_oracle_stats)]
        _lib.rayca_oracle_kat_quantize.restype = C.c_uint8
        _lib.rayca_oracle_kat_quantize.argtypes = [C.c_float]
        _lib.rayca_oracle_kat_aabb.restype = C.c_float
        _lib.rayca_oracle_kat_rng.restype = C.c_float
        _lib.rayca_oracle_kat_rng.argtypes = [C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32]
    return _lib


class AuxPlanes:
    """Host planes matching rayca_aux_out."""

    def __init__(self, width, height):
        n = width * height
        self.hit_blas = np.full(n, 0xFFFFFFFF, dtype=np.uint32)
        self.hit_prim = np.full(n, 0xFFFFFFFF, dtype=np.uint32)
        self.hit_src = np.full(n, 0xFFFFFFFF, dtype=np.uint32)
        self.depth = np.full(n, np.inf, dtype=np.float32)
        self.uv = np.zeros(n * 2, dtype=np.float32)
        self.rgba_f32 = np.zeros(n * 4, dtype=np.float32)
        self.c = M.rayca_aux_out(self.hit_blas.ctypes.data, self.hit_prim.ctypes.data, self.hit_src.ctypes.data,
                                 self.depth.ctypes.data, self.uv.ctypes.data, self.rgba_f32.ctypes.data)


class OracleScene:
    def __init__(self, scene_desc: M.SceneDesc, bvh: bool = True):
        self._h = C.c_void_p()
        rc = lib().rayca_oracle_scene_create(C.cast(scene_desc.ptr(), C.c_void_p), 1 if bvh else 0, C.byref(self._h))
        if rc != 0:
            raise RuntimeError(f"rayca_oracle_scene_create failed: {rc}")

    def close(self):
        if self._h:
            lib().rayca_oracle_scene_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        self.close()

    def bvh(self):
        info = M.rayca_bvh_info()
        lib().rayca_oracle_scene_bvh_info(self._h, C.byref(info))
        nodes = np.zeros((info.node_count, 8), dtype=np.float32)
        no = np.zeros(info.blas_count + 1, dtype=np.uint32)
        po = np.zeros(info.blas_count + 1, dtype=np.uint32)
        src = np.zeros(info.triangle_count, dtype=np.uint32)
        lib().rayca_oracle_scene_bvh_read(self._h, nodes.ctypes.data, no.ctypes.data, po.ctypes.data, src.ctypes.data)
        return info, nodes, no, po, src

    def render(self, config: M.rayca_config, width, height, options=None, flags=0, threads=0, rows=None,
               want_aux=True):
        rgba8 = np.zeros((height, width, 4), dtype=np.uint8)
        aux = AuxPlanes(width, height) if want_aux else None
        stats = rayca_oracle_stats()
        rows_c = (C.c_uint32 * 2)(*rows) if rows is not None else None
        rc = lib().rayca_oracle_render(self._h, C.byref(config), C.byref(options) if options is not None else None,
                                       width, height, rgba8.ctypes.data, C.byref(aux.c) if aux else None, flags,
                                       threads, rows_c, C.byref(stats))
        if rc != 0:
            raise RuntimeError(f"rayca_oracle_render failed: {rc}")
        return rgba8, aux, stats
