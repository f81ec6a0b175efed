#!/usr/bin/env python
"""bench.py — Mrays/s (primary + shadow) of the raytracing hot path on N B200s, beside the CPU path.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload atrium|soup]

One step = one frame: every pixel's primary ray, its shadow rays and shading, through the CUDA path.
Workload (config.workload): BASELINE.json quotes its metric on 1080p Sponza; `sponza.bin` is absent
from the reference checkout and nothing of the reference exists on the GPU box, so the default is the
procedural Sponza-class atrium of rayca_b200/procedural.py (~262k triangles, one BLAS, textured PBR,
the reference test's root rotation / camera / default lights), 1920x1080, 1 spp,
Pathtracer max_depth=1 (= the deterministic primary + shadow slice of SoftRenderer::default()).

Keys: `value` = device-timed whole-job throughput with the scene resident in HBM; `e2e` = the same
metric through rayca_cuda_render with HOST output buffers (device->host copy of the frame inside the
timed region; the per-step host->device input is the camera/config parameter block — the scene is
uploaded once by scene_create, exactly as the reference's own timer excludes its BVH build,
rayca-soft/src/scene.rs:101,152); `draw_e2e` = the full Draw::draw equivalent (flatten + BVH build +
upload + render + read-back).  `cpu_baseline` = the oracle (a port of rayca-soft) on the host cores.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

def metric_name(args):
    """BASELINE.json's metric on the workload actually run (the default is the 1080p Sponza-class frame)."""
    rays = "primary+shadow" if args.max_depth <= 1 else "primary+shadow+bounce"
    scene = {"atrium": "Sponza-class scene", "flighthelmet": "FlightHelmet textured PBR model"}.get(args.workload, f"{args.workload} scene")
    res = "1080p" if (args.width, args.height) == (1920, 1080) else f"{args.width}x{args.height}"
    return f"Mrays/s ({rays}, {res} {scene}" + (f", {args.spp} spp" if args.spp > 1 else "") + ")"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="atrium", choices=["atrium", "soup", "flighthelmet"])
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--height", type=int, default=1080)
    ap.add_argument("--spp", type=int, default=1)
    ap.add_argument("--triangles", type=int, default=0, help="override the workload's triangle count")
    ap.add_argument("--traversal", type=int, default=0)
    ap.add_argument("--flags", type=int, default=0, help="rayca_options.flags for the timed launches")
    ap.add_argument("--max-depth", type=int, default=1, help="Pathtracer max_depth (2 = one cosine bounce, BASELINE M5)")
    ap.add_argument("--gather", choices=["peer", "nccl", "host"], default="host",
                    help="N>1: host = kernels store tiles straight into a shared page-locked host frame (zero copy, default); "
                         "peer = kernels store tiles into rank 0's device frame over NVLink; nccl = all_gather_into_tensor")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-rows", type=int, default=0, help="rows of the frame the CPU baseline renders")
    return ap.parse_args()


class _FixtureScene:
    """A committed POD scene descriptor (tests/golden/*.scene.npz) standing in for a Scene object."""

    def __init__(self, name):
        from rayca_b200 import model as M

        self.desc = M.SceneDesc.load(os.path.join(ROOT, "tests", "golden", f"{name}.scene.npz"))

    def to_desc(self):
        return self.desc


def make_scene(args):
    from rayca_b200 import procedural as P

    if args.workload == "soup":
        n = args.triangles or 10_000_000
        return P.soup(n), f"procedural triangle soup, {n} triangles (BASELINE.md M4 generator)"
    if args.workload == "flighthelmet":
        # BASELINE config 2 asks for DamagedHelmet, which the reference checkout does not hold; its in-repo textured PBR
        # model is FlightHelmet with the camera of the reference's `flight` test (rayca-soft/tests/gltf.rs:296-317)
        return _FixtureScene("flighthelmet"), ("FlightHelmet.gltf (94,722 triangles, 6 PBR materials with albedo + normal + "
                                               "metallic/roughness maps, textures point-subsampled to 256x256; tests/golden/"
                                               "flighthelmet.scene.npz), camera of the reference's `flight` test — stand-in for "
                                               "BASELINE config 2 (DamagedHelmet is not in the reference checkout)")
    n = args.triangles or 262_267
    return P.atrium(n), ("procedural Sponza-class atrium (sponza.bin absent from the reference checkout), "
                         "one BLAS, textured PBR, reference sponza-test camera/lights")


def render_config(args):
    from rayca_b200.renderer import Config

    return Config(max_depth=args.max_depth, samples_per_pixel=args.spp)


class ClockSampler(threading.Thread):
    """Samples SM clocks and throttle reasons with nvidia-smi while the timed region runs."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, device_index):
        super().__init__(daemon=True)
        self.dev = device_index
        self.samples = []
        self.stop = threading.Event()
        self.ready = threading.Event()  # set once the first sample is in: the caller starts timing after that

    def run_nvml(self):
        """In-process NVML polling (every 0.5 ms): a 20-step timed region lasts tens of milliseconds, which one
        nvidia-smi process per sample cannot resolve."""
        import pynvml as N

        N.nvmlInit()
        h = N.nvmlDeviceGetHandleByIndex(self.dev)
        mx = N.nvmlDeviceGetMaxClockInfo(h, N.NVML_CLOCK_SM)
        get_reasons = getattr(N, "nvmlDeviceGetCurrentClocksEventReasons", None) or N.nvmlDeviceGetCurrentClocksThrottleReasons
        bits = ((0x8, 5), (0x40, 6), (0x20, 7), (0x4, 8))  # hw_slowdown, hw_thermal, sw_thermal, sw_power_cap
        last = False
        while not last:  # one more sample after stop: a timed region of a few milliseconds still gets its closing sample
            last = self.stop.is_set()
            r = int(get_reasons(h))
            f = [str(self.dev), str(N.nvmlDeviceGetClockInfo(h, N.NVML_CLOCK_SM)), str(mx), "", hex(r), "", "", "", ""]
            for bit, slot in bits:
                f[slot] = "Active" if r & bit else "Not Active"
            self.samples.append(f)
            self.ready.set()
            self.stop.wait(0.0005)

    def run(self):
        try:
            self.run_nvml()
            return
        except Exception:
            pass
        while not self.stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                      "-i", str(self.dev)], capture_output=True, text=True, timeout=5).stdout
                f = [x.strip() for x in out.strip().split(",")]
                if len(f) >= 9:
                    self.samples.append(f)
                    self.ready.set()
            except Exception:
                pass
            self.stop.wait(0.1)

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unsampled"]}
        sm = sorted(float(s[1]) for s in self.samples)
        reasons = set()
        for s in self.samples:
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), s[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(self.samples[0][2]), "reasons": sorted(reasons),
                "samples": len(self.samples)}


def sample_rows(args, rows_total):
    """The fixed, evenly spaced image rows every CPU leg renders (the same rows in every step and every arm)."""
    return [int(y) for y in ((np.arange(rows_total) + 0.5) / rows_total * args.height).astype(int) % args.height]


RNG_SEED = 0x5EED  # key of the Philox counter RNG, the same on both sides


def oracle_options():
    from rayca_b200 import model as M

    return M.rayca_options(RNG_SEED, 0, 0, 0, 0, 0, 0)


def cpu_leg(osc, cfg, args, ys, flags, threads=0, keep=False):
    """Times the oracle's pixel loop on image rows `ys` (one call per row; the oracle hands out pixels, not rows, to
    its threads — rayca-soft/src/scene.rs:117-122 — so one row of 1920 pixels keeps every core busy).
    Returns (rays, ms, threads that rendered pixels, {row: (rgba8 row, aux planes)} when keep)."""
    from oracle import binding as O  # noqa: F401

    rays, ms, th, kept = 0, 0.0, 0, {}
    for y in ys:
        rgba8, aux, st = osc.render(cfg.to_c(), args.width, args.height, options=oracle_options(), flags=flags, threads=threads,
                                    rows=(int(y), int(y) + 1), want_aux=keep)
        rays += st.rays_primary + st.rays_shadow + st.rays_secondary
        ms += st.trace_ms
        th = max(th, st.threads)
        if keep:
            W = args.width
            sl = slice(y * W, (y + 1) * W)
            kept[y] = {"rgba8": rgba8[y].copy(), "hit_src": aux.hit_src[sl].copy(), "hit_prim": aux.hit_prim[sl].copy(),
                       "hit_blas": aux.hit_blas[sl].copy(), "depth": aux.depth[sl].copy(),
                       "uv": aux.uv[2 * y * W:2 * (y + 1) * W].copy(), "rgba_f32": aux.rgba_f32[4 * y * W:4 * (y + 1) * W].copy()}
    return rays, ms, th, kept


def host_threads():
    """Host cores this process may use (torchrun exports OMP_NUM_THREADS=1: the oracle gets an explicit count)."""
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def cpu_legs(desc, cfg, args, rows, gpu=None):
    os.environ.setdefault("OMP_NUM_THREADS", str(host_threads()))
    """Every CPU figure of the bench line, all on the same fixed rows of the same frame:
      faithful   = the reference's algorithm (exhaustive unordered walk of the reference BVH, vertices re-transformed
                   per test) on all host threads — "rayca-soft on the host cores";
      1 thread   = the same on one thread, one row;
      optimised  = ordered, distance-culled walk over a binned-SAH tree of pre-transformed triangles (oracle OptTree),
                   whole frame, all host threads — what a tuned CPU renderer would do, so the speed-up is not inflated.
    With `gpu` = (rgba8 frame, aux planes) of the CUDA path the oracle rows are compared with it (parity_rows)."""
    from oracle import binding as O

    W, H = args.width, args.height
    nthr = host_threads()
    osc = O.OracleScene(desc, bvh=True)
    ys = sample_rows(args, rows)
    out = {}
    r, m, th, kept = cpu_leg(osc, cfg, args, ys, O.FAITHFUL, threads=nthr, keep=gpu is not None)
    faithful = r / (m * 1e-3) / 1e6
    out["cpu_baseline"] = {"value": faithful, "unit": "Mrays/s", "cores": th, "kind": "port",
                           "sample": f"image rows {ys} of the {W}x{H} frame (pixels of a row spread over the threads), pixel loop only "
                                     "(BVH build excluded), reference-faithful mode: exhaustive walk of the reference BVH, vertices "
                                     "re-transformed per test", "host_threads_available": nthr}
    r1, m1, th1, _ = cpu_leg(osc, cfg, args, ys[len(ys) // 2:len(ys) // 2 + 1], O.FAITHFUL, threads=1)
    out["cpu_1thread"] = {"value": r1 / (m1 * 1e-3) / 1e6, "unit": "Mrays/s", "cores": th1, "kind": "port",
                          "sample": f"image row {ys[len(ys) // 2]}, reference-faithful mode, one thread"}
    out["cpu_baseline"]["scaling_vs_1thread"] = faithful / max(out["cpu_1thread"]["value"], 1e-30)
    opt_frame = None
    try:
        rgba8, aux, st = osc.render(cfg.to_c(), W, H, options=oracle_options(), flags=O.OPTIMISED, threads=nthr, want_aux=gpu is not None)
        rays = st.rays_primary + st.rays_shadow + st.rays_secondary
        out["cpu_optimised"] = {"value": rays / (st.trace_ms * 1e-3) / 1e6, "unit": "Mrays/s", "cores": st.threads, "kind": "port",
                                "tree_build_ms": st.opt_build_ms,
                                "sample": f"the whole {W}x{H} frame, ordered + distance-culled walk over a binned-SAH tree of "
                                          "pre-transformed triangles (oracle/rayca_oracle.cpp OptTree); same pixels as the faithful mode"}
        opt_frame = (rgba8, aux)
    except RuntimeError as e:  # sphere scenes: the optimised arm covers triangle scenes
        out["cpu_optimised"] = {"unavailable": str(e)}
    if gpu is not None:
        g8, gaux = gpu
        stochastic = args.max_depth > 1  # bounce directions go through cosf / sinf / acosf: ulp differences, rare different hits

        def rgb_err(a, b):  # per pixel, absolute (relative above 1.0 for the unclamped HDR values of stochastic frames)
            e = np.abs(a.reshape(-1, 4) - b.reshape(-1, 4))
            if stochastic:
                e = e / np.maximum(np.abs(b.reshape(-1, 4)), 1.0)
            return np.nan_to_num(e, nan=0.0).max(axis=1) if len(e) else np.zeros(0)

        idm, derr, u8, errs = 0, 0, 0, []
        for y, k in kept.items():
            sl = slice(y * W, (y + 1) * W)
            idm += int((gaux.hit_src[sl] != k["hit_src"]).sum() + (gaux.hit_prim[sl] != k["hit_prim"]).sum() +
                       (gaux.hit_blas[sl] != k["hit_blas"]).sum())
            derr += int((gaux.depth[sl].view(np.uint32) != k["depth"].view(np.uint32)).sum() +
                        (gaux.uv[2 * y * W:2 * (y + 1) * W].view(np.uint32) != k["uv"].view(np.uint32)).sum())
            errs.append(rgb_err(gaux.rgba_f32[4 * y * W:4 * (y + 1) * W], k["rgba_f32"]))
            u8 = max(u8, int(np.abs(g8[y].astype(np.int16) - k["rgba8"].astype(np.int16)).max()))
        errs = np.concatenate(errs)
        out["parity_rows"] = {"rows": ys, "against": "oracle, reference-faithful exhaustive walk", "pixels": len(ys) * W,
                              "id_mismatch": idm, "depth_uv_bit_mismatch": derr, "max_rgb_err": float(errs.max()),
                              "pixels_beyond_1e-3": int((errs > 1e-3).sum()), "max_rgba8_diff": u8}
        if opt_frame is not None:
            o8, oaux = opt_frame
            ferr = rgb_err(gaux.rgba_f32, oaux.rgba_f32)
            out["parity_frame"] = {
                "against": "oracle, optimised arm (checked against the exhaustive walk on the sampled rows), whole frame",
                "pixels": W * H,
                "id_mismatch": int((gaux.hit_src != oaux.hit_src).sum() + (gaux.hit_prim != oaux.hit_prim).sum()),
                "depth_uv_bit_mismatch": int((gaux.depth.view(np.uint32) != oaux.depth.view(np.uint32)).sum() +
                                             (gaux.uv.view(np.uint32) != oaux.uv.view(np.uint32)).sum()),
                "max_rgb_err": float(ferr.max()), "pixels_beyond_1e-3": int((ferr > 1e-3).sum()),
                "max_rgba8_diff": int(np.abs(g8.astype(np.int16) - o8.astype(np.int16)).max()),
                "rgba8_bytes_differing": int((g8 != o8).sum()),
                "optimised_equals_exhaustive_on_rows": all(
                    np.array_equal(oaux.hit_src[y * W:(y + 1) * W], k["hit_src"]) and
                    np.array_equal(oaux.depth[y * W:(y + 1) * W].view(np.uint32), k["depth"].view(np.uint32)) and
                    np.array_equal(oaux.rgba_f32[4 * y * W:4 * (y + 1) * W].view(np.uint32), k["rgba_f32"].view(np.uint32))
                    for y, k in kept.items())}
        # verdict: ids / depth / uv bit-exact everywhere; RGB within 1e-3 (stochastic frames: on >= 99 % of the pixels)
        bad = idm + derr
        allowed = 0.01 if stochastic else 0.0
        bad += int(out["parity_rows"]["pixels_beyond_1e-3"] > allowed * len(errs))
        if "parity_frame" in out:
            pf = out["parity_frame"]
            bad += pf["id_mismatch"] + pf["depth_uv_bit_mismatch"] + int(pf["pixels_beyond_1e-3"] > allowed * W * H)
            bad += int(not pf["optimised_equals_exhaustive_on_rows"])
        out["parity_ok"] = bad == 0
    osc.close()
    return out


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path on the host cores.  The Rust
    crate cannot be built here (nightly toolchain absent), so this is the oracle port in
    reference-faithful mode on all host threads; every step renders the SAME fixed rows of the frame."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    # torchrun exports OMP_NUM_THREADS=1; the oracle's scene build (and libgomp's pool) reads it when the library loads
    os.environ["OMP_NUM_THREADS"] = str(host_threads())
    from oracle import binding as O

    scene, what = make_scene(args)
    desc = scene.to_desc()
    cfg = render_config(args)
    rows = args.cpu_rows or 4
    ys = sample_rows(args, rows)
    nthr = host_threads()
    osc = O.OracleScene(desc, bvh=True)
    for _ in range(min(args.warmup, 1)):
        cpu_leg(osc, cfg, args, ys[:1], O.FAITHFUL, threads=nthr)
    rays, ms, threads = 0, 0.0, 1
    for k in range(args.steps):
        r, m, threads, _ = cpu_leg(osc, cfg, args, ys, O.FAITHFUL, threads=nthr)
        rays += r
        ms += m
    osc.close()
    value = rays / (ms * 1e-3) / 1e6
    line = {"impl": "reference", "metric": metric_name(args), "value": value, "unit": "Mrays/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / max(args.steps, 1),
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": bench_config(args, what, desc),
            "cpu_baseline": {"value": value, "unit": "Mrays/s", "cores": threads, "kind": "port",
                             "sample": f"image rows {ys} of the {args.width}x{args.height} frame in every step (pixels of a row spread "
                                       "over the threads), pixel loop only (BVH build excluded), reference-faithful mode",
                             "host_threads_available": nthr},
            "e2e": {"value": value, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


def bench_config(args, what, desc):
    """`config` of the JSON line: derived from the arguments only, so both arms print the same object (the reference arm
    renders a bounded sample of this workload; the sample is described under cpu_baseline.sample)."""
    G, tile = args.gpus, 8
    gather = {"host": "tiles stored by the trace kernel straight into a shared page-locked HOST frame (POSIX shm registered by every rank), no gather",
              "peer": "tiles stored into rank 0's device frame by the trace kernel over NVLink (CUDA IPC), per-rank completion flags",
              "nccl": "NCCL all_gather of RGBA8 tiles"}[args.gather]
    return {"workload": what, "width": args.width, "height": args.height, "spp": args.spp, "triangles": desc.triangle_count,
            "integrator": f"Pathtracer max_depth={args.max_depth} (NEE point lights" +
                          (", cosine bounce, Philox keyed RNG)" if args.max_depth > 1 else ")"),
            "parallelism": (f"image row-tiles of {tile} rows interleaved over {G} GPUs, scene replicated, " + gather) if G > 1
                           else "one GPU, whole frame in one launch",
            "l2": "GPU arm: flushed between timed iterations (256 MiB fill)", "traversal": args.traversal}


def ncu_profile_name(args, G):
    """profiles/ file that holds the `ncu --set full` summary of this exact bench command (tools/ncu_summary.py)."""
    if G != 1:
        return "r02_ncu_none.json"
    if args.workload == "soup":
        return "r02_ncu_soup.json"
    if args.workload == "flighthelmet":
        return "r02_ncu_flighthelmet.json"
    return "r02_ncu_bounce_split_shade.json" if args.max_depth > 1 else "r02_ncu_atrium.json"


def run_ours(args):
    import torch

    from rayca_b200 import build, model as M, renderer as R

    build.build()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device: there is no CPU fallback for the product path")
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    G = world
    W, H = args.width, args.height
    scene, what = make_scene(args)
    desc = scene.to_desc()
    cfg = render_config(args)
    ctx = R.Context(local_rank)
    t0 = time.time()
    dev = R.DeviceScene(ctx, desc, bvh=True)
    build_s = time.time() - t0

    tile = 8
    tiles = (H + tile - 1) // tile
    per = (tiles + G - 1) // G
    local_rows = per * tile if G > 1 else H
    local = torch.zeros((local_rows, W, 4), dtype=torch.uint8, device="cuda")
    gathered = torch.zeros((G, local_rows, W, 4), dtype=torch.uint8, device="cuda") if G > 1 else None
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")  # > 126 MB L2
    opts = M.rayca_options(0x5EED, args.traversal, rank, G, tile, args.flags, 0)
    stream = torch.cuda.current_stream().cuda_stream

    # N > 1, --gather peer: rank 0 owns the frame; every rank's trace kernel stores its tiles
    # straight into it over NVLink (CUDA IPC mapping) and publishes a per-rank completion flag, rank 0
    # waits on the flags on its stream.  --gather nccl: local tiles + all_gather_into_tensor.
    frame, frame_id = None, [0]
    peer = G > 1 and args.gather == "peer"
    frames = []
    if peer:
        for _ in range(2):  # two frames: the e2e loop traces frame k+1 while rank 0 still reads frame k back
            handle = [None]
            if rank == 0:
                frames.append(R.SharedFrame(ctx, W, H))
                handle[0] = frames[-1].export()
            dist.broadcast_object_list(handle, src=0)
            if rank != 0:
                frames.append(R.SharedFrame(ctx, W, H, handle=handle[0]))
        frame = frames[0]

    # N > 1, --gather host: the frame lives in POSIX shared memory that every rank page-locks; each rank's kernel stores
    # its tiles straight into it over its own PCIe link (zero copy): no gather and no device-to-host copy at all.
    host_mode = G > 1 and args.gather == "host"
    hframes = []
    if host_mode:
        for _ in range(2):
            nm = [None]
            if rank == 0:
                hframes.append(R.HostFrame(W, H))
                nm[0] = hframes[-1].name
            dist.broadcast_object_list(nm, src=0)
            if rank != 0:
                hframes.append(R.HostFrame(W, H, name=nm[0]))
    hopts = M.rayca_options(0x5EED, args.traversal, rank, G, tile, args.flags | R.FLAG_DIRECT_ROWS, 0)

    def trace():
        if host_mode:
            dev.render_device(cfg, W, H, hframes[0].device_ptr, stream=stream, options=hopts)
        elif peer:
            frame_id[0] += 1
            dev.render_to_frame(cfg, frame, frame_id[0], stream=stream, options=opts)
        else:
            dev.render_device(cfg, W, H, local.data_ptr(), stream=stream, options=opts)

    def gather():
        if host_mode:
            pass  # nothing to collect: the pixels are already at their place in host memory when the launches are done
        elif peer:
            if rank == 0:
                frame.wait(G, frame_id[0], stream=stream)
        elif G > 1:
            dist.all_gather_into_tensor(gathered, local)

    def step():
        trace()
        gather()

    def barrier():
        if G > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        step()
    barrier()
    if host_mode:
        # the host frame must equal the NCCL-gathered tiles byte for byte once every rank's launch is complete
        dev.render_device(cfg, W, H, local.data_ptr(), stream=stream, options=opts)
        dist.all_gather_into_tensor(gathered, local)
        expect = gathered.view(G, per, tile, W, 4).permute(1, 0, 2, 3, 4).reshape(per * G * tile, W, 4)[:H].cpu().numpy()
        for _ in range(4):
            if rank == 0:
                hframes[0].pixels[...] = 0
            barrier()
            step()
            barrier()
            if rank == 0 and not np.array_equal(hframes[0].pixels, expect):
                raise RuntimeError("host frame differs from the NCCL-gathered frame")
            barrier()
    if peer:
        # The shared frame must equal the NCCL-gathered tiles, byte for byte, at the moment the flag wait
        # returns: rank 0 clears the frame before every round, so a flag that overtook its pixels shows up
        # as black tiles.
        dev.render_device(cfg, W, H, local.data_ptr(), stream=stream, options=opts)
        dist.all_gather_into_tensor(gathered, local)
        expect = gathered.view(G, per, tile, W, 4).permute(1, 0, 2, 3, 4).reshape(per * G * tile, W, 4)[:H].contiguous()
        shared = torch.as_tensor(frame, device="cuda") if rank == 0 else None
        for _ in range(16):
            if rank == 0:
                shared.zero_()
            barrier()
            step()
            if rank == 0 and not torch.equal(shared, expect):  # same stream as the wait: ordered after it
                raise RuntimeError("shared frame differs from the NCCL-gathered frame")
            barrier()
        if rank == 0 and frame.timed_out():
            raise RuntimeError("shared frame: a rank's completion flag never arrived")
    # ray counts of one frame (deterministic): this rank's share, summed over ranks
    st = dev.last_counters()
    rays_local = torch.tensor([st.rays_primary, st.rays_shadow, st.rays_secondary], dtype=torch.int64, device="cuda")
    rays_local_n = st.rays_primary + st.rays_shadow + st.rays_secondary
    if G > 1:
        dist.all_reduce(rays_local)
    rays_frame = int(rays_local.sum().item())

    # algorithmic bytes of one launch (untimed counting pass)
    copts = M.rayca_options(0x5EED, args.traversal, rank, G, tile, R.FLAG_COUNT_WALK, 0)
    dev.render_device(cfg, W, H, local.data_ptr(), stream=stream, options=copts)
    torch.cuda.synchronize()
    cst = dev.last_counters()
    px_local = st.rays_primary // max(args.spp, 1)
    algo_bytes = cst.node_bytes + cst.triangle_bytes + 4 * px_local  # + RGBA8 store per pixel

    sampler = ClockSampler(local_rank)
    sampler.start()
    sampler.ready.wait(5.0)
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    kev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    barrier()
    del sampler.samples[:]  # keep the samples taken from here on: under load
    for k in range(args.steps):
        flush.fill_(k & 0xFF)  # L2 flush between timed iterations (outside the timed events)
        ev[k][0].record()
        kev[k][0].record()
        trace()
        kev[k][1].record()
        gather()
        ev[k][1].record()
    barrier()
    sampler.stop.set()
    sampler.join()
    total_ms = sum(a.elapsed_time(b) for a, b in ev)
    launches_per_frame = max(int(dev.last_counters().kernel_launches), 1)  # of the last timed frame
    kernel_ms = sum(a.elapsed_time(b) for a, b in kev) / args.steps
    t = torch.tensor([total_ms], dtype=torch.float64, device="cuda")
    if G > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms = float(t.item())
    value = rays_frame * args.steps / (total_ms * 1e-3) / 1e6

    # e2e: the frame lands in (pinned) HOST memory inside the timed region, every step.
    #   N == 1: rayca_cuda_render, the reference-facing entry point (row bands traced on one stream and
    #           read back on another);
    #   N  > 1: each rank traces its tiles, NCCL all_gather, rank 0 copies the gathered frame to the host.
    # Wall clock between barriers, max over ranks.  Per-step host->device input = the kernel parameter
    # block (camera, config, pointers); the scene was uploaded once by scene_create, which the
    # reference's own timer excludes too (scene.rs:101,152).
    param_bytes = 1024
    frame_crc = None
    n_e2e = max(3, args.steps)
    draw_e2e = None
    if G == 1:
        img = R.Image(W, H, pinned=True)
        for _ in range(3):
            dev.render(cfg, W, H, out=img.buffer, options=opts)
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        for _ in range(n_e2e):
            dev.render(cfg, W, H, out=img.buffer, options=opts)
        e2e_s = (time.perf_counter() - t1) / n_e2e
        d2h = W * H * 4
        import zlib

        frame_crc = zlib.crc32(np.ascontiguousarray(img.buffer).tobytes())  # A/B runs: same pixels, byte for byte
        ab_ref = os.environ.get("RAYCA_AB_FRAME")  # tools/ab2.sh: the first variant saves its frame, the others diff against it
        if ab_ref:
            if os.path.exists(ab_ref):
                dq = np.abs(np.load(ab_ref).astype(np.int16) - img.buffer.astype(np.int16))
                frame_crc = f"{frame_crc} maxdiff_u8={int(dq.max())} differing_bytes={int((dq > 0).sum())}"
            else:
                np.save(ab_ref, np.ascontiguousarray(img.buffer))
        e2e_what = "rayca_cuda_render: RGBA8 frame stored by the kernel straight into the caller's pinned host buffer (zero copy over PCIe), scene resident (scene_create excluded, like the reference timer)"
    else:
        srcs = [torch.as_tensor(f, device="cuda") for f in frames] if (peer and rank == 0) else [gathered]
        host = torch.empty(tuple(srcs[0].shape), dtype=torch.uint8).pin_memory() if rank == 0 else None
        ids = [frame_id[0], 0]
        if peer and rank == 0:
            frames[0].release(ids[0], stream=stream)  # the timed loop above consumed nothing: catch up
        e2e_k = [0]

        uses = [0, 0]  # host mode: how often each of the two host frames has been filled (same on every rank)

        def e2e_step():
            if host_mode:
                # every rank: wait until rank 0 consumed the frame's previous content, store this rank's tiles through
                # rayca_cuda_render (synchronous: returns when the pixels are in host memory), publish a host flag;
                # rank 0: spin on the G flags — then the whole frame is in host memory — and mark it consumed
                b = e2e_k[0] & 1
                e2e_k[0] += 1
                f = hframes[b]
                while f.flags[R.HostFrame.SLOTS - 1] < uses[b]:
                    pass
                dev.render(cfg, W, H, out=f.pixels, options=hopts)
                uses[b] += 1
                f.flags[rank] = uses[b]
                if rank == 0:
                    while (f.flags[:G] < uses[b]).any():
                        pass
                    f.flags[R.HostFrame.SLOTS - 1] = uses[b]
                return
            if peer:
                # frames used alternately; a rank overwrites a frame only after rank 0 released its last use
                b = e2e_k[0] & 1
                e2e_k[0] += 1
                f = frames[b]
                f.acquire(ids[b], stream=stream)
                ids[b] += 1
                dev.render_to_frame(cfg, f, ids[b], stream=stream, options=opts)
                if rank == 0:
                    f.wait(G, ids[b], stream=stream)
                    host.copy_(srcs[b], non_blocking=True)
                    f.release(ids[b], stream=stream)
            else:
                step()
                if rank == 0:
                    host.copy_(srcs[0], non_blocking=True)
            torch.cuda.synchronize()  # rank 0: the frame is on the host before the next one starts

        for _ in range(4):
            e2e_step()
        torch.cuda.synchronize()
        if host_mode:
            barrier()
            if rank == 0 and not (np.array_equal(hframes[0].pixels, expect) and np.array_equal(hframes[1].pixels, expect)):
                raise RuntimeError("e2e: host frame differs from the NCCL-gathered frame")
        elif rank == 0 and not torch.equal(host, srcs[0].cpu()):
            raise RuntimeError("e2e: host frame differs from the device frame")
        barrier()
        t1 = time.perf_counter()
        for _ in range(n_e2e):
            e2e_step()
        barrier()
        if peer and any(f.timed_out() for f in frames):
            raise RuntimeError("shared frame: a flag wait timed out during the e2e loop")
        e2e_t = torch.tensor([(time.perf_counter() - t1) / n_e2e], dtype=torch.float64, device="cuda")
        dist.all_reduce(e2e_t, op=dist.ReduceOp.MAX)
        e2e_s = float(e2e_t.item())
        d2h = int(host.numel()) if rank == 0 else 0
        if host_mode:
            d2h = W * H * 4 // G  # each rank stores its own tiles into the host frame
        e2e_what = ("rayca_cuda_render on every rank with RAYCA_FLAG_DIRECT_ROWS: each rank's kernel stores its tiles straight into a shared, page-locked host frame over its own PCIe link (zero copy, two frames alternating, host-side completion / consumed flags); no gather, no device-to-host copy"
                    if host_mode else "render_to_frame on every rank (peer stores into rank 0's frame, two frames alternating with acquire / release flags) + flag wait + rank-0 copy to pinned host memory"
                    if peer else "render_device on every rank + NCCL all_gather + rank-0 copy of the frame to pinned host memory")
    e2e = {"value": rays_frame / e2e_s / 1e6, "unit": "Mrays/s", "h2d_bytes_per_step": param_bytes,
           "d2h_bytes_per_step": d2h, "ms_per_step": e2e_s * 1e3, "what": e2e_what}
    if G == 1 and not isinstance(scene, _FixtureScene):
        img = R.Image(W, H)
        rnd = R.CudaRenderer.__new__(R.CudaRenderer)
        rnd.config, rnd._ctx = cfg, ctx
        draws = []
        for _ in range(4):  # the first call pays one-time allocations; report the median of the next three
            t1 = time.perf_counter()
            rnd.draw(scene, img)
            draws.append((time.perf_counter() - t1) * 1e3)
        lib_draws = []  # the library's share: what a Rust caller pays after its own FlatScene::from (rayca_cuda_draw-equivalent)
        for _ in range(4):
            t1 = time.perf_counter()
            d2 = R.DeviceScene(ctx, desc, cfg.bvh)
            d2.render(cfg, W, H, out=img.buffer)
            d2.close()
            lib_draws.append((time.perf_counter() - t1) * 1e3)
        draw_e2e = {"ms": sorted(draws[1:])[1], "first_ms": draws[0], "library_ms": sorted(lib_draws[1:])[1],
                    "what": "Draw::draw, median of 3 after one warm-up call: to_desc (Python harness) + flatten + reference BLAS and "
                            "performance tree (both built on the GPU) + upload + render + read-back; library_ms = the same without "
                            "to_desc: rayca_cuda_scene_create + rayca_cuda_render + rayca_cuda_scene_destroy through the C ABI"}

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
        achieved = algo_bytes / (kernel_ms * 1e-3) / 1e9
        # the hot set is L2 / L1 resident, so the governing memory bound is the L2 -> SM read path.
        # Probe: 128-bit ld.global.cg (L1 bypassed) over a 64 MiB L2-resident buffer from all SMs; the physical
        # ceiling is 148 SMs x 64 B/clk x f_SM, which the answer must not exceed.
        l2_peak = ctx.probe_read_bandwidth(64 << 20, 32)
        clk = sampler.summary()
        l2_ceiling = 148 * 64 * float(clk.get("sm_max_mhz") or 1965.0) * 1e6 / 1e9
        # what physically moved per launch: the committed `ncu --set full` summary of this exact command, if any
        prof_name = ncu_profile_name(args, G)
        phys, traffic, issue_pct, simt = None, None, None, None
        try:
            prof = json.load(open(os.path.join(ROOT, "profiles", prof_name)))
            k0 = prof["launches"][0]

            def num(key):
                v, u = (k0[key].split() + [""])[:2]
                return float(v) * {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1.0)

            dur = num("gpu__time_duration.sum") * {"us": 1e-6, "ms": 1e-3, "ns": 1e-9}.get(k0["gpu__time_duration.sum"].split()[1], 1e-6)
            l2_read = num("lts__t_sectors_srcunit_tex_op_read.sum") * 32.0
            l1_read = num("l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum") * 32.0
            dram = num("dram__bytes_read.sum") + num("dram__bytes_write.sum")
            traffic = int(dram)
            issue_pct = num("smsp__issue_active.avg.pct_of_peak_sustained_active")
            simt = num("smsp__thread_inst_executed_per_inst_executed.ratio")
            phys = {"source": "profiles/" + prof_name, "kernel_us_under_ncu": dur * 1e6,
                    "l1_read_bytes": int(l1_read), "l2_to_l1_read_bytes": int(l2_read),
                    "dram_read_bytes": int(num("dram__bytes_read.sum")), "dram_write_bytes": int(num("dram__bytes_write.sum")),
                    "l2_gbs": l2_read / dur / 1e9, "l2_frac": l2_read / dur / 1e9 / l2_peak,
                    "dram_gbs": dram / dur / 1e9, "dram_frac": dram / dur / 1e9 / hbm_peak,
                    "logical_over_l2_bytes": algo_bytes / max(l2_read, 1.0)}
        except Exception:
            pass
        line = {"metric": metric_name(args), "value": value, "unit": "Mrays/s", "n_gpus": G, "steps": args.steps,
                "warmup": max(args.warmup, 3), "ms_per_step": total_ms / args.steps, "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": bench_config(args, what, desc),
                "rays_per_frame": rays_frame, "kernel_ms": kernel_ms, "build_s": build_s,
                "clocks": clk,
                # this library's kernels per frame as the library counts them (1 for the fused kernel; fill + stages + copy
                # for the wavefront path, cub's sort kernels not counted) x steps; peer: + flag signal (+ flag wait on rank 0)
                "gpu_launches": args.steps * (launches_per_frame + ((2 if rank == 0 else 1) if peer else 0)),
                "roofline": {"bound": "l2", "achieved": achieved, "peak": l2_peak, "unit": "GB/s", "frac": achieved / l2_peak,
                             "traffic": traffic, "peak_source": "measured in this run (rayca_cuda_probe_read_bandwidth: ld.global.cg "
                                                                "over a 64 MiB L2-resident buffer, best of 5)",
                             "peak_ceiling_148sm_x_64B_per_clk": l2_ceiling,
                             "hbm": {"peak": hbm_peak, "frac_logical": achieved / hbm_peak,
                                     "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback",
                                     "note": "logical bytes against HBM: not a roofline fraction (the bytes are served by L1 / L2), kept for continuity with round 1"},
                             "algorithmic_bytes_per_launch": int(algo_bytes),
                             "physical": phys, "issue_active_pct": issue_pct, "simt_lanes_per_instruction": simt,
                             "governing": "issue slots (see issue_active_pct and profiles/): neither L2 nor HBM is near its peak in physical bytes",
                             "note": "achieved = LOGICAL bytes per launch (32 B per node record + 16 B per triangle plane + 48 B per "
                                     "triangle whose vertices are fetched, counted per lane by the kernel in a COUNT launch, + 4 B per "
                                     "pixel) / kernel time; lanes of a warp that read the same record count it 32 times, L1 serves "
                                     "most of them - `physical` holds the bytes ncu saw move (per launch, committed capture)"},
                "e2e": e2e, "draw_e2e": draw_e2e, "frame_crc32": frame_crc,
                "walk": {"node_records_per_ray": cst.node_bytes / 32 / max(rays_local_n, 1),
                         "triangle_bytes_per_ray": cst.triangle_bytes / max(rays_local_n, 1),
                         "algorithmic_bytes_per_ray": algo_bytes / max(rays_local_n, 1)}}
        big = desc.triangle_count > 2_000_000 and not args.cpu_rows
        if big:
            line["cpu_baseline"] = {"unavailable": "the oracle's verbatim reference builder needs many minutes on this triangle count; "
                                                   "pass --cpu-rows N to run the CPU legs anyway (parity of this workload's kernel path: "
                                                   "tests/test_cuda_bench_frames.py::test_soup_16spp_rows_against_the_oracle)"}
        if not args.no_cpu_baseline and G == 1 and not big:
            # the CUDA frame with its parity planes, through the reference-facing entry point, for the oracle comparison
            g8, gaux, _ = dev.render(cfg, W, H, options=opts, want_aux=True)
            line.update(cpu_legs(desc, cfg, args, args.cpu_rows or 8, gpu=(g8, gaux)))
            bad = not line["parity_ok"]
            print(json.dumps(line))
            if bad:
                sys.stderr.write("bench.py: the CUDA frame differs from the oracle (see parity_rows / parity_frame)\n")
                sys.exit(3)
        else:
            print(json.dumps(line))
    for f in hframes:
        f.close()
    if G > 1:
        dist.destroy_process_group()


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
