"""Host-side mirror of rayca-soft's renderer interface over the C ABI of librayca_cuda.

Reference interface mirrored here (same names, argument meaning and failure behaviour):
    trait Draw { fn draw(&mut self, scene: &Scene, image: &mut Image) }   rayca-soft/src/draw.rs:7-9
    SoftRenderer::default() / new_with_config(Config)                      rayca-soft/src/scene.rs:11-14,57-61
    Config (bon builder defaults)                                          rayca-soft/src/config.rs:10-49
    IntegratorStrategy / SamplerStrategy                                   integrator/mod.rs:32-41, sampler/mod.rs:41-50

There is no CPU fallback: constructing a renderer without the built CUDA library, or without a
CUDA device, raises.  Nothing in this module imports or calls the oracle.
"""
from __future__ import annotations

import ctypes as C
import enum
import os
from dataclasses import dataclass
from typing import Optional

import numpy as np

from . import model as M

_HERE = os.path.dirname(os.path.abspath(__file__))
# RAYCA_CUDA_LIB selects a tuning variant of the same library (kernel experiments); never a fallback
LIB_PATH = os.environ.get("RAYCA_CUDA_LIB") or os.path.join(_HERE, "librayca_cuda.so")


class IntegratorStrategy(enum.IntEnum):
    Scratcher = 0
    Raytracer = 1
    Flat = 2
    AnalyticDirect = 3
    Direct = 4
    Pathtracer = 5


class SamplerStrategy(enum.IntEnum):
    NONE = 0
    Nee = 1
    Hemisphere = 2
    Cosine = 3
    Brdf = 4
    Mis = 5


TRAVERSAL_ORDERED, TRAVERSAL_EXHAUSTIVE, TRAVERSAL_REFERENCE_ORDERED = 0, 1, 2
FLAG_COUNT_WALK = 1
FLAG_PERSISTENT = 2
FLAG_DIRECT_ROWS = 4
FLAG_NO_WAVEFRONT = 16  # keep the fused kernel for bounce frames (A/B, tests)
FLAG_SHARD_GROUPS = 8  # with DIRECT_ROWS: shard by 16x8 pixel groups in a diagonal checkerboard instead of row tiles
STATIC_FRAME_BYTES = 8 << 20  # RAYCA_STATIC_FRAME_BYTES: largest width * height * 4 rayca_cuda_render_static accepts

_ERRORS = {-1: "RAYCA_E_INVALID_ARG", -2: "RAYCA_E_UNSUPPORTED", -3: "RAYCA_E_OOM", -4: "RAYCA_E_CUDA",
           -5: "RAYCA_E_NO_CAMERA", -6: "RAYCA_E_TIMEOUT"}


class RaycaError(RuntimeError):
    def __init__(self, code: int, detail: str):
        super().__init__(f"{_ERRORS.get(code, code)}: {detail}")
        self.code = code


@dataclass
class Config:
    """rayca-soft/src/config.rs:10-49 with the builder defaults."""
    bvh: bool = True
    light_samples: int = 1
    light_stratify: bool = False
    samples_per_pixel: int = 1
    russian_roulette: bool = False
    direct_sampler: SamplerStrategy = SamplerStrategy.Nee
    indirect_sampler: SamplerStrategy = SamplerStrategy.Cosine
    integrator: IntegratorStrategy = IntegratorStrategy.Pathtracer
    max_depth: int = 5
    gamma: float = 1.0

    def apply(self, sdtf: M.SdtfConfig) -> None:
        """Config::apply (config.rs:58-71)."""
        self.max_depth = 16 if sdtf.max_depth == -1 else sdtf.max_depth
        self.light_samples = sdtf.light_samples
        self.light_stratify = sdtf.light_stratify
        self.samples_per_pixel = sdtf.samples_per_pixel
        self.direct_sampler = {"none": SamplerStrategy.NONE, "nee": SamplerStrategy.Nee, "mis": SamplerStrategy.Mis,
                               "hemisphere": SamplerStrategy.Hemisphere, "cosine": SamplerStrategy.Cosine,
                               "brdf": SamplerStrategy.Brdf}[sdtf.direct_sampler]
        self.indirect_sampler = {"hemisphere": SamplerStrategy.Hemisphere, "cosine": SamplerStrategy.Cosine,
                                 "brdf": SamplerStrategy.Brdf}[sdtf.indirect_sampler]
        self.integrator = {"raytracer": IntegratorStrategy.Raytracer,
                           "analyticdirect": IntegratorStrategy.AnalyticDirect, "direct": IntegratorStrategy.Direct,
                           "pathtracer": IntegratorStrategy.Pathtracer}[sdtf.integrator]
        self.gamma = sdtf.gamma

    def to_c(self) -> M.rayca_config:
        return M.rayca_config(int(self.bvh), self.light_samples, int(self.light_stratify), self.samples_per_pixel,
                              int(self.russian_roulette), int(self.direct_sampler), int(self.indirect_sampler),
                              int(self.integrator), self.max_depth, self.gamma)


class PinnedBuffer:
    """Page-locked host memory from rayca_cuda_host_alloc, viewed as a numpy array."""

    def __init__(self, shape, dtype=np.uint8):
        self._p = C.c_void_p()
        n = int(np.prod(shape)) * np.dtype(dtype).itemsize
        _check(load_library().rayca_cuda_host_alloc(n, C.byref(self._p)))
        self.array = np.ctypeslib.as_array(C.cast(self._p, C.POINTER(C.c_uint8)), shape=(n,)).view(dtype).reshape(shape)

    def close(self):
        if self._p:
            self.array = None
            load_library().rayca_cuda_host_free(self._p)
            self._p = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


@dataclass
class Image:
    """rayca-model/src/image.rs:26-61 restricted to ColorType::RGBA8: row-major, top-left origin.
    `pinned=True` takes the pixel buffer from rayca_cuda_host_alloc (overlapped read-back)."""
    width: int
    height: int
    buffer: Optional[np.ndarray] = None
    pinned: bool = False

    def __post_init__(self):
        if self.buffer is None:
            if self.pinned:
                self._pin = PinnedBuffer((self.height, self.width, 4))
                self.buffer = self._pin.array
                self.buffer[...] = 0
            else:
                self.buffer = np.zeros((self.height, self.width, 4), dtype=np.uint8)

    def dump_png(self, path) -> None:
        """Image::dump_png (image.rs:160-169), encoded by the library (rayca_cuda_write_png)."""
        buf = np.ascontiguousarray(self.buffer)
        _check(load_library().rayca_cuda_write_png(os.fspath(path).encode(), buf.ctypes.data, self.width, self.height))

    def encode_png(self) -> bytes:
        """The frame as PNG bytes (rayca_cuda_encode_png)."""
        lib = load_library()
        buf = np.ascontiguousarray(self.buffer)
        n = C.c_size_t()
        lib.rayca_cuda_encode_png(buf.ctypes.data, self.width, self.height, None, 0, C.byref(n))
        out = C.create_string_buffer(n.value)
        _check(lib.rayca_cuda_encode_png(buf.ctypes.data, self.width, self.height, out, n.value, C.byref(n)))
        return out.raw[:n.value]


_lib = None


def load_library():
    """Loads rayca_b200/librayca_cuda.so (built in-tree by rayca_b200/build.py).  Fails loudly."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(f"{LIB_PATH} is missing: run `python -m rayca_b200.build` (there is no CPU fallback)")
    lib = C.CDLL(LIB_PATH)
    vp, u32 = C.c_void_p, C.c_uint32
    lib.rayca_cuda_host_alloc.argtypes = [C.c_size_t, C.POINTER(vp)]
    lib.rayca_cuda_host_free.argtypes = [vp]
    lib.rayca_cuda_host_free.restype = None
    lib.rayca_cuda_host_register.argtypes = [vp, C.c_size_t, C.POINTER(vp)]
    lib.rayca_cuda_host_unregister.argtypes = [vp]
    lib.rayca_cuda_probe_read_bandwidth.argtypes = [vp, C.c_size_t, u32, C.POINTER(C.c_double)]
    lib.rayca_cuda_frame_create.argtypes = [vp, u32, u32, C.POINTER(vp)]
    lib.rayca_cuda_frame_export.argtypes = [vp, C.c_char_p]
    lib.rayca_cuda_frame_import.argtypes = [vp, C.c_char_p, u32, u32, C.POINTER(vp)]
    lib.rayca_cuda_frame_pixels.argtypes = [vp]
    lib.rayca_cuda_frame_pixels.restype = vp
    lib.rayca_cuda_frame_destroy.argtypes = [vp]
    lib.rayca_cuda_frame_destroy.restype = None
    lib.rayca_cuda_render_to_frame.argtypes = [vp, vp, C.POINTER(M.rayca_config), C.POINTER(M.rayca_options), vp, C.c_uint64, vp]
    lib.rayca_cuda_frame_wait.argtypes = [vp, vp, u32, C.c_uint64, vp]
    lib.rayca_cuda_frame_release.argtypes = [vp, vp, C.c_uint64, vp]
    lib.rayca_cuda_frame_acquire.argtypes = [vp, vp, C.c_uint64, vp]
    lib.rayca_cuda_frame_timed_out.argtypes = [vp]
    lib.rayca_cuda_frame_set_timeout.argtypes = [vp, C.c_double]
    lib.rayca_cuda_frame_sync.argtypes = [vp, vp, vp]
    lib.rayca_cuda_frame_clear_timeout.argtypes = [vp]
    lib.rayca_cuda_frame_clear_timeout.restype = None
    lib.rayca_cuda_init.argtypes = [C.c_int, C.POINTER(vp)]
    lib.rayca_cuda_init_devices.argtypes = [C.POINTER(C.c_int), C.c_int, C.POINTER(vp)]
    lib.rayca_cuda_device_count.argtypes = [vp]
    lib.rayca_cuda_shutdown.argtypes = [vp]
    lib.rayca_cuda_shutdown.restype = None
    lib.rayca_cuda_scene_create.argtypes = [vp, vp, u32, C.POINTER(vp), C.POINTER(M.rayca_stats)]
    lib.rayca_cuda_scene_create_host.argtypes = [vp, u32, C.POINTER(vp), C.POINTER(M.rayca_stats)]
    lib.rayca_cuda_scene_destroy.argtypes = [vp]
    lib.rayca_cuda_scene_destroy.restype = None
    lib.rayca_cuda_render.argtypes = [vp, vp, C.POINTER(M.rayca_config), C.POINTER(M.rayca_options), u32, u32, vp,
                                      C.POINTER(M.rayca_aux_out), C.POINTER(M.rayca_stats)]
    lib.rayca_cuda_render_static.argtypes = [vp, vp, C.POINTER(M.rayca_config), C.POINTER(M.rayca_options), u32, u32, vp,
                                             C.POINTER(M.rayca_stats)]
    lib.rayca_cuda_render_device.argtypes = [vp, vp, C.POINTER(M.rayca_config), C.POINTER(M.rayca_options), u32, u32,
                                             vp, C.POINTER(M.rayca_aux_out), vp, C.POINTER(M.rayca_stats)]
    lib.rayca_cuda_last_counters.argtypes = [vp, C.POINTER(M.rayca_stats)]
    lib.rayca_cuda_draw.argtypes = [vp, C.POINTER(M.rayca_config), u32, u32, vp]
    lib.rayca_cuda_scene_bvh_info.argtypes = [vp, C.POINTER(M.rayca_bvh_info)]
    lib.rayca_cuda_scene_bvh_read.argtypes = [vp, vp, vp, vp, vp]
    lib.rayca_cuda_accum_create.argtypes = [vp, u32, u32, C.POINTER(vp)]
    lib.rayca_cuda_accum_destroy.argtypes = [vp]
    lib.rayca_cuda_accum_destroy.restype = None
    lib.rayca_cuda_accum_reset.argtypes = [vp]
    lib.rayca_cuda_accum_reset.restype = None
    lib.rayca_cuda_accum_samples.argtypes = [vp]
    lib.rayca_cuda_accum_samples.restype = u32
    lib.rayca_cuda_render_accumulate.argtypes = [vp, vp, C.POINTER(M.rayca_config), C.POINTER(M.rayca_options), vp, u32, vp,
                                                 C.POINTER(M.rayca_stats)]
    lib.rayca_cuda_write_png.argtypes = [C.c_char_p, vp, u32, u32]
    lib.rayca_cuda_encode_png.argtypes = [vp, u32, u32, vp, C.c_size_t, C.POINTER(C.c_size_t)]
    lib.rayca_cuda_strerror.restype = C.c_char_p
    lib.rayca_cuda_strerror.argtypes = [C.c_int]
    lib.rayca_cuda_last_error.restype = C.c_char_p
    lib.rayca_cuda_abi_version.restype = u32
    _lib = lib
    return lib


def _check(rc: int) -> None:
    if rc != 0:
        raise RaycaError(rc, load_library().rayca_cuda_last_error().decode())


class AuxPlanes:
    """Host planes for rayca_aux_out (parity checks)."""

    def __init__(self, width: int, height: int):
        n = width * height
        self.hit_blas = np.full(n, 0xFFFFFFFF, dtype=np.uint32)
        self.hit_prim = np.full(n, 0xFFFFFFFF, dtype=np.uint32)
        self.hit_src = np.full(n, 0xFFFFFFFF, dtype=np.uint32)
        self.depth = np.full(n, np.inf, dtype=np.float32)
        self.uv = np.zeros(n * 2, dtype=np.float32)
        self.rgba_f32 = np.zeros(n * 4, dtype=np.float32)
        self.c = M.rayca_aux_out(self.hit_blas.ctypes.data, self.hit_prim.ctypes.data, self.hit_src.ctypes.data,
                                 self.depth.ctypes.data, self.uv.ctypes.data, self.rgba_f32.ctypes.data)


class Context:
    """rayca_ctx.  `Context(device)`: one GPU (one process per GPU).  `Context(devices=[...])`: one context over several
    GPUs driven by this thread (rayca_cuda_init_devices; `devices=[]` = every visible device): DeviceScene replicates the
    scene and DeviceScene.render shards the frame over them."""

    def __init__(self, device: int = 0, devices=None):
        self._h = C.c_void_p()
        if devices is None:
            _check(load_library().rayca_cuda_init(device, C.byref(self._h)))
            self.device = device
        else:
            arr = (C.c_int * max(len(devices), 1))(*devices)
            _check(load_library().rayca_cuda_init_devices(arr if devices else None, len(devices), C.byref(self._h)))
            self.device = devices[0] if devices else 0
        self.device_count = load_library().rayca_cuda_device_count(self._h)

    def probe_read_bandwidth(self, nbytes: int, reps: int) -> float:
        """GB/s of a streaming read over an nbytes buffer (L2-resident when small): bench.py's L2 roofline."""
        out = C.c_double()
        _check(load_library().rayca_cuda_probe_read_bandwidth(self._h, nbytes, reps, C.byref(out)))
        return out.value

    def close(self):
        if self._h:
            load_library().rayca_cuda_shutdown(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class HostFrame:
    """An RGBA8 frame in POSIX shared memory that every one-process-per-GPU rank of a node maps and page-locks
    (rayca_cuda_host_register): with FLAG_DIRECT_ROWS each rank's kernel stores its tiles straight into the frame
    over its own PCIe link, so the frame is complete in HOST memory once every rank's launch is — no gather, no
    device-to-host copy.  Behind the pixels sit 64 int64 slots for host-side flags (per-rank "frame k stored",
    slot 63 = "consumed up to k").  `HostFrame(w, h)` creates (owner), `HostFrame(w, h, name=...)` attaches."""

    SLOTS = 64

    def __init__(self, width: int, height: int, name: Optional[str] = None):
        from multiprocessing import shared_memory

        self.width, self.height, self.owner = width, height, name is None
        nbytes = width * height * 4 + 8 * self.SLOTS
        self._shm = shared_memory.SharedMemory(create=True, size=nbytes) if name is None else shared_memory.SharedMemory(name=name)
        self.name = self._shm.name
        if name is not None:  # the owner unlinks; keep this process's resource tracker from doing it again at exit
            try:
                from multiprocessing import resource_tracker

                resource_tracker.unregister(self._shm._name, "shared_memory")
            except Exception:
                pass
        raw = np.frombuffer(self._shm.buf, dtype=np.uint8, count=nbytes)
        self.pixels = raw[: width * height * 4].reshape(height, width, 4)
        self.flags = raw[width * height * 4:].view(np.int64)
        if self.owner:
            raw[...] = 0
        self._addr = raw.ctypes.data
        dv = C.c_void_p()
        _check(load_library().rayca_cuda_host_register(self._addr, nbytes, C.byref(dv)))
        self.device_ptr = dv.value
        self._registered = True

    def close(self):
        if getattr(self, "_registered", False):
            load_library().rayca_cuda_host_unregister(self._addr)
            self._registered = False
            self.pixels = self.flags = None
            try:
                self._shm.close()
                if self.owner:
                    self._shm.unlink()
            except Exception:
                pass

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class SharedFrame:
    """rayca_frame: an RGBA8 frame in the owner GPU's HBM that every rank's kernel writes into over NVLink
    (CUDA IPC mapping), with one completion flag per rank.  `SharedFrame(ctx, w, h)` creates (owner),
    `SharedFrame(ctx, w, h, handle=...)` imports the owner's `export()`ed handle."""

    def __init__(self, ctx: Context, width: int, height: int, handle: Optional[bytes] = None):
        self.ctx, self.width, self.height, self.owner = ctx, width, height, handle is None
        self._h = C.c_void_p()
        lib = load_library()
        if handle is None:
            _check(lib.rayca_cuda_frame_create(ctx._h, width, height, C.byref(self._h)))
        else:
            _check(lib.rayca_cuda_frame_import(ctx._h, handle, width, height, C.byref(self._h)))

    def export(self) -> bytes:
        buf = C.create_string_buffer(64)
        _check(load_library().rayca_cuda_frame_export(self._h, buf))
        return buf.raw

    @property
    def pixels_ptr(self) -> int:
        return load_library().rayca_cuda_frame_pixels(self._h)

    @property
    def __cuda_array_interface__(self):
        return {"shape": (self.height, self.width, 4), "typestr": "|u1", "data": (self.pixels_ptr, False), "version": 3}

    def wait(self, rank_count: int, frame_id: int, stream: int = 0) -> None:
        _check(load_library().rayca_cuda_frame_wait(self.ctx._h, self._h, rank_count, frame_id, stream))

    def release(self, frame_id: int, stream: int = 0) -> None:
        """Owner: frames up to frame_id have been consumed (stream order)."""
        _check(load_library().rayca_cuda_frame_release(self.ctx._h, self._h, frame_id, stream))

    def acquire(self, frame_id: int, stream: int = 0) -> None:
        """Any rank: wait (stream order) until the owner released frame_id before overwriting the frame."""
        _check(load_library().rayca_cuda_frame_acquire(self.ctx._h, self._h, frame_id, stream))

    def timed_out(self) -> bool:
        return bool(load_library().rayca_cuda_frame_timed_out(self._h))

    def set_timeout(self, milliseconds: float) -> None:
        """Wall-clock limit of the flag waits of this frame object (default 2000 ms)."""
        _check(load_library().rayca_cuda_frame_set_timeout(self._h, milliseconds))

    def sync(self, stream: int = 0) -> None:
        """Waits for `stream`; raises RaycaError(RAYCA_E_TIMEOUT) if a flag wait gave up: the frame is then undefined."""
        _check(load_library().rayca_cuda_frame_sync(self.ctx._h, self._h, stream))

    def clear_timeout(self) -> None:
        load_library().rayca_cuda_frame_clear_timeout(self._h)

    def close(self):
        if self._h:
            load_library().rayca_cuda_frame_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Accumulator:
    """rayca_accum: running per-pixel sums for progressive rendering (DeviceScene.render_accumulate)."""

    def __init__(self, ctx: Context, width: int, height: int):
        self.ctx, self.width, self.height = ctx, width, height
        self._h = C.c_void_p()
        _check(load_library().rayca_cuda_accum_create(ctx._h, width, height, C.byref(self._h)))

    @property
    def samples(self) -> int:
        return load_library().rayca_cuda_accum_samples(self._h)

    def reset(self) -> None:
        load_library().rayca_cuda_accum_reset(self._h)

    def close(self):
        if self._h:
            load_library().rayca_cuda_accum_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class DeviceScene:
    """rayca_scene: the flattened scene + BVH resident in HBM (or host-only when ctx is None)."""

    def __init__(self, ctx: Optional[Context], desc: M.SceneDesc, bvh: bool = True):
        self.ctx = ctx
        self._h = C.c_void_p()
        self.stats = M.rayca_stats()
        lib = load_library()
        d = C.cast(desc.ptr(), C.c_void_p)
        if ctx is None:
            _check(lib.rayca_cuda_scene_create_host(d, int(bvh), C.byref(self._h), C.byref(self.stats)))
        else:
            _check(lib.rayca_cuda_scene_create(ctx._h, d, int(bvh), C.byref(self._h), C.byref(self.stats)))

    def close(self):
        if self._h:
            load_library().rayca_cuda_scene_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def bvh(self):
        lib = load_library()
        info = M.rayca_bvh_info()
        _check(lib.rayca_cuda_scene_bvh_info(self._h, C.byref(info)))
        nodes = np.zeros((info.node_count, 8), dtype=np.float32)
        no = np.zeros(info.blas_count + 1, dtype=np.uint32)
        po = np.zeros(info.blas_count + 1, dtype=np.uint32)
        src = np.zeros(info.triangle_count, dtype=np.uint32)
        _check(lib.rayca_cuda_scene_bvh_read(self._h, nodes.ctypes.data, no.ctypes.data, po.ctypes.data,
                                             src.ctypes.data))
        return info, nodes, no, po, src

    def render(self, config: Config, width: int, height: int, options: Optional[M.rayca_options] = None,
               want_aux: bool = False, out: Optional[np.ndarray] = None):
        """rayca_cuda_render with HOST buffers; returns (rgba8 HxWx4, aux or None, stats).
        `out`: caller-owned HxWx4 uint8 buffer (e.g. Image.buffer), written in place."""
        if out is not None:
            assert out.shape == (height, width, 4) and out.dtype == np.uint8 and out.flags["C_CONTIGUOUS"]
        rgba8 = out if out is not None else np.zeros((height, width, 4), dtype=np.uint8)
        aux = AuxPlanes(width, height) if want_aux else None
        stats = M.rayca_stats()
        cfg = config.to_c()
        _check(load_library().rayca_cuda_render(self.ctx._h, self._h, C.byref(cfg),
                                                C.byref(options) if options is not None else None, width, height,
                                                rgba8.ctypes.data, C.byref(aux.c) if aux else None, C.byref(stats)))
        return rgba8, aux, stats

    def render_static(self, config: Config, width: int, height: int, options: Optional[M.rayca_options] = None,
                      out: Optional[np.ndarray] = None):
        """rayca_cuda_render_static: the allocation-free render (scratch in a static device array of the library, frames
        up to STATIC_FRAME_BYTES).  Returns (rgba8 HxWx4, stats).  Calls on one device must not overlap."""
        if out is not None:
            assert out.shape == (height, width, 4) and out.dtype == np.uint8 and out.flags["C_CONTIGUOUS"]
        rgba8 = out if out is not None else np.zeros((height, width, 4), dtype=np.uint8)
        stats = M.rayca_stats()
        cfg = config.to_c()
        _check(load_library().rayca_cuda_render_static(self.ctx._h, self._h, C.byref(cfg),
                                                       C.byref(options) if options is not None else None, width, height,
                                                       rgba8.ctypes.data, C.byref(stats)))
        return rgba8, stats

    def render_accumulate(self, config: Config, accum: Accumulator, sample_count: int,
                          options: Optional[M.rayca_options] = None, out: Optional[np.ndarray] = None):
        """rayca_cuda_render_accumulate: the next `sample_count` of config.samples_per_pixel samples; returns
        (rgba8 of the mean so far, stats)."""
        rgba8 = out if out is not None else np.zeros((accum.height, accum.width, 4), dtype=np.uint8)
        stats = M.rayca_stats()
        cfg = config.to_c()
        _check(load_library().rayca_cuda_render_accumulate(self.ctx._h, self._h, C.byref(cfg),
                                                           C.byref(options) if options is not None else None, accum._h,
                                                           sample_count, rgba8.ctypes.data, C.byref(stats)))
        return rgba8, stats

    def render_device(self, config: Config, width: int, height: int, d_rgba8: int, stream: int = 0,
                      options: Optional[M.rayca_options] = None, d_aux: Optional[M.rayca_aux_out] = None):
        """rayca_cuda_render_device: outputs are device pointers; enqueues on `stream` without syncing."""
        cfg = config.to_c()
        _check(load_library().rayca_cuda_render_device(self.ctx._h, self._h, C.byref(cfg),
                                                       C.byref(options) if options is not None else None, width,
                                                       height, d_rgba8, C.byref(d_aux) if d_aux is not None else None,
                                                       stream, None))

    def render_to_frame(self, config: Config, frame: SharedFrame, frame_id: int, stream: int = 0,
                        options: Optional[M.rayca_options] = None):
        """rayca_cuda_render_to_frame: this rank's tiles go straight into the (possibly remote) shared frame."""
        cfg = config.to_c()
        _check(load_library().rayca_cuda_render_to_frame(self.ctx._h, self._h, C.byref(cfg),
                                                         C.byref(options) if options is not None else None, frame._h,
                                                         frame_id, stream))

    def last_counters(self) -> M.rayca_stats:
        stats = M.rayca_stats()
        _check(load_library().rayca_cuda_last_counters(self.ctx._h, C.byref(stats)))
        return stats


class CudaRenderer:
    """Drop-in for `SoftRenderer` (rayca-soft/src/scene.rs:11-14,57-61,88-154)."""

    def __init__(self, config: Optional[Config] = None, device: int = 0, devices=None):
        """`devices`: list of CUDA devices one `draw` call fans out over ([] = all visible); default: `device` alone."""
        self.config = config if config is not None else Config()
        self._ctx = Context(device, devices=devices)

    @staticmethod
    def new_with_config(config: Config) -> "CudaRenderer":
        return CudaRenderer(config)

    @staticmethod
    def create_default_model() -> M.Model:
        return M.create_default_model()

    def draw(self, scene: M.Scene, image: Image) -> None:
        """`Draw::draw`: rebuilds the acceleration structure on every call, like the reference."""
        desc = scene.to_desc()
        dev = DeviceScene(self._ctx, desc, self.config.bvh)
        try:
            dev.render(self.config, image.width, image.height, out=image.buffer)
        finally:
            dev.close()
