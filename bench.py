#!/usr/bin/env python
"""bench.py - the headline benchmark of BASELINE.json: Gsamples/s and frames/s of the ray march.

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --gpus N --steps K ...  # the reference's render math on host cores

Workload (BASELINE.json configs[1], "example1_grad-shaped"): synthetic ViBE-Z-shaped 1024^3 fp32
emission volume E, reflection R = E^2, a distinct half-resolution absorption volume A (resident;
as written the reference samples absorption from E), precomputed gradient volumes
Gx,Gy,Gz = MATLAB gradient(E), 2 Henyey-Greenstein lights, 2048x2048 mono image, camera of
examples/example1_grad.m.  One step = one frame; the camera turns 1 degree per step
(example2-style) so no frame is a replay of the previous one.

N > 1: image-space mode - every rank holds all volumes, 64x64 tiles are dealt round-robin to the
ranks, partial images are SUM-reduced to rank 0 over NCCL inside the timed region (strong
scaling of one frame).
"""
import argparse
import ctypes as C
import json
import math
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

SEED = 0x5EED01BE
# DRAM traffic of k_raymarch per frame from this round's ncu --set full captures (dram__bytes_read.sum + dram__bytes_write.sum),
# keyed by (volume edge, image edge, gradient method, dtype, texture unit on/off).  Not measured live: bench.py scales it to
# the frames of its own launches and names the profile it came from.
DRAM_PROFILE = {
    (1024, 2048, "lookup", "f32", 1): {"bytes_per_frame": 1.46e9, "profile": "profiles/r02_final_summary.md"},
    (1024, 2048, "lookup", "f32", 0): {"bytes_per_frame": 2.6e9, "profile": "profiles/r01_v5_summary.md"},
}


def parse():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=0, help="frames (default 16; example2: 360)")
    p.add_argument("--warmup", type=int, default=3)
    p.add_argument("--impl", default="b200", choices=["b200", "reference"])
    p.add_argument("--size", type=int, default=1024, help="volume edge (default 1024 = the metric's config)")
    p.add_argument("--config", default="example1_grad", choices=sorted(CONFIGS) + ["example3"],
                   help="BASELINE.json configuration: example1_grad (default, the metric's), example1, example2 (360-frame 1920x1080 movie), example3 (two channels, stereo, partial updates)")
    p.add_argument("--image", type=int, default=0, help="square image edge (default: the configuration's own resolution)")
    p.add_argument("--gradient", default=None, choices=["lookup", "compute"], help="default: the configuration's")
    p.add_argument("--lights", type=int, default=None, help="default: the configuration's")
    p.add_argument("--dtype", default="f32", choices=["f32", "bf16"])
    p.add_argument("--no-cpu-baseline", action="store_true")
    p.add_argument("--cpu-seconds", type=float, default=15.0, help="target CPU time of the cpu_baseline sample")
    p.add_argument("--variant", type=int, default=0)
    p.add_argument("--skip", type=int, default=2)
    p.add_argument("--tile", type=int, default=64)
    p.add_argument("--e2e-batch", type=int, default=10, help="frames per pipelined batch of the end-to-end leg (the host ring holds two batches)")
    p.add_argument("--e2e-tile", type=int, default=32, help="N > 1, end-to-end leg: width of the tile columns dealt round-robin to the ranks")
    p.add_argument("--batch", type=int, default=32, help="frames per launch at most (one work queue per batch; the steps are cut into equal batches)")
    p.add_argument("--wshape", type=int, default=3, help="warp footprint (1<<s) x (32>>s) pixels")
    p.add_argument("--no-upload-probe", action="store_true", help="skip the host->resident upload measurement")
    p.add_argument("--no-gather-probe", action="store_true", help="skip the gather-ceiling microbenchmark (roofline.gather)")
    p.add_argument("--interleave", type=int, default=1, help="1: Gx,Gy,Gz resident as one interleaved storage (default); 0: three volumes")
    p.add_argument("--texture-unit", type=int, default=2, help="VR_OPT_TEXTURE_UNIT: 2 (default) texture unit where the context can, 1 required, 0 loads + fp32 lerps")
    p.add_argument("--arith", type=int, default=-1, help="VR_OPT_ARITHMETIC: 0 exact, 1 reference, 2 fast; -1 = the library's default")
    p.add_argument("--no-load-leg", action="store_true", help="skip the extra load-path measurement of the default run")
    p.add_argument("--order", type=int, default=1, help="1 = tiles handed out centre-first (default), 0 = raster order")
    p.add_argument("--mode", default="image", choices=["image", "volume"],
                   help="image = tiles over replicated volumes (the metric's config); volume = bricks sharded over the ranks, sort-last composite (config 5)")
    p.add_argument("--opt", action="append", default=[], metavar="K=V", help="--mode volume: extra vr_set_option(K, V) calls (A/B and debugging)")
    p.add_argument("--vs-slices", type=int, default=0, help="--mode volume: VR_OPT_VS_SLICES (0 = the library's automatic choice, 1 = whole rays)")
    p.add_argument("--exchange", default="p2p", choices=["p2p", "nccl"],
                   help="N > 1: p2p = every rank's kernel stores its tiles into rank 0's image over NVLink (default); nccl = SUM-reduce of partial images")
    a = p.parse_args()
    cfg = CONFIGS.get(a.config, {})
    if a.gradient is None:
        a.gradient = cfg.get("gradient", "compute")
    if not a.steps:
        a.steps = cfg.get("steps") or 16
    if a.config in CONFIGS and cfg["volumes"] == "aliased" and a.gradient == "lookup":
        p.error("--config %s renders with on-the-fly gradients" % a.config)
    return a


# ---- the scene (examples/example1_grad.m:36-51,79-82 on synthetic data) ---------------------------
def rotation(alpha, beta, gamma, R=None):
    """VolumeRender.rotate (VolumeRender.m:239-262)"""
    c, s = lambda a: math.cos(math.radians(a)), lambda a: math.sin(math.radians(a))
    Rx = np.array([[1, 0, 0], [0, c(alpha), -s(alpha)], [0, s(alpha), c(alpha)]])
    Ry = np.array([[c(beta), 0, s(beta)], [0, 1, 0], [-s(beta), 0, c(beta)]])
    Rz = np.array([[c(gamma), -s(gamma), 0], [s(gamma), c(gamma), 0], [0, 0, 1]])
    R = np.eye(3) if R is None else R
    return R @ Rx @ Ry @ Rz


# BASELINE.json configs as bench workloads (--config).  "volumes": six = distinct E, R = E^2, A and Gx, Gy, Gz (configs[1]);
# aliased = Em = Ab = E as one object, default 1x1x1 reflection volume (VolumeRender.m:131), on-the-fly gradient.
EX1_LIGHTS = [([500, 1000, 550], [0, 1, 1]), ([550, 1000, 500], [1, 0, 1])]  # example1.m:36
CONFIGS = {
    # configs[1], examples/example1_grad.m:36-51,79-82 - the configuration the metric is quoted on (default)
    "example1_grad": dict(image=(2048, 2048), gradient="lookup", lights=EX1_LIGHTS, factors=[0.1, 0.1, 0.4], thr=0.9, color=[1, 1, 1],
                          focal=3.0, dist=6.0, R0=lambda: rotation(125, 25, 0), volumes="six", steps=None),
    # configs[0], examples/example1.m:28-60: 1 HG light, on-the-fly gradients, 1024 x 1024 mono
    "example1": dict(image=(1024, 1024), gradient="compute", lights=EX1_LIGHTS[:1], factors=[1.0, 0.4, 0.6], thr=0.9, color=[1, 1, 0],
                     focal=3.0, dist=6.0, R0=lambda: rotation(125, 25, 0), volumes="aliased", steps=None),
    # configs[2], examples/example2.m:28,53-66: 360-degree movie, 360 frames at 1920 x 1080, volumes resident
    "example2": dict(image=(1920, 1080), gradient="compute", lights=[([1500, 1500, 0], [1, 1, 1])], factors=[10.0, 0.3, 1.0], thr=0.95, color=[1, 1, 1],
                     focal=3.0, dist=6.0, R0=lambda: rotation(-15, -15, 15, rotation(90, 0, 0)), volumes="aliased", steps=360),
}


class Scene:
    def __init__(self, args):
        cfg = CONFIGS[args.config]
        self.cfg = cfg
        self.W, self.H = (args.image, args.image) if args.image else cfg["image"]  # ImageResolution = [W, H]
        self.factors = np.array(cfg["factors"], dtype=np.float32)  # [fEm, fRe, fAb] (VolumeRender.m:558)
        self.es = np.array([1, 1, 1], dtype=np.float32)
        self.res = np.array([self.H, self.W], dtype=np.uint64)
        self.props = np.array([0.0, cfg["focal"], cfg["dist"]], dtype=np.float32)  # CameraXOffset, FocalLength, DistanceToObject
        self.thr = cfg["thr"]
        self.color = np.array(cfg["color"], dtype=np.float32)
        self.R0 = cfg["R0"]()
        self.lights = cfg["lights"][: args.lights] if args.lights is not None else cfg["lights"]
        self.aliased = cfg["volumes"] == "aliased"

    def rot(self, step):
        R = rotation(0, float(step), 0, self.R0)
        return np.asfortranarray(np.flipud(R).astype(np.float32))  # flip(RotationMatrix), VolumeRender.m:560


def f32p(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


# ---- clocks during the timed region ----------------------------------------------------------------
class ClockSampler:
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.rows = []
        self.proc = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "100", "-i", str(index)],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), line.strip()))

    def stop(self, t0, t1):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, power = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ts, line in self.rows:
            if ts < t0 or ts > t1 + 0.2:
                continue
            f = [x.strip() for x in line.split(",")]
            try:
                sm.append(float(f[0])), mx.append(float(f[1])), power.append(float(f[2]))
            except Exception:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(power) if power else None, "samples": len(sm), "reasons": sorted(reasons)}


# ---- oracle (CPU) legs: cpu_baseline of the b200 arm, and the whole reference arm --------------------
def oracle_lib():
    import ctypes

    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import vrt

    return vrt, ctypes


def host_cores():
    """Cores this process may run on (torchrun exports OMP_NUM_THREADS=1 to its workers: the CPU legs ask for all of them explicitly)."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except Exception:
        return max(1, os.cpu_count() or 1)


def frames_per_launch(steps, batch):
    """Equal batches: 20 steps with --batch 16 are 10 + 10 frames, not 16 + 4 (a short last batch is nearly all tail)."""
    steps, batch = max(1, steps), max(1, batch)
    launches = -(-steps // batch)
    return -(-steps // launches)


def mem_available_gib():
    try:
        for line in open("/proc/meminfo"):
            if line.startswith("MemAvailable"):
                return int(line.split()[1]) / (1024.0 * 1024.0)
    except Exception:
        pass
    return 0.0


class CpuRender:
    """The reference's per-ray render math (oracle/oracle.cpp, OpenMP over 16x16 tiles, all host
    cores) on host copies of the workload's volumes."""

    def __init__(self, scene, args, vols, fast=True, filter_bits=0):
        vrt, _ = oracle_lib()
        self.vrt = vrt
        self.scene = scene
        from volume_renderer_b200 import _abi, api

        self.abi, self.api = _abi, api
        self.lib = vrt.oracle(fast=fast)
        self.h = C.c_void_p(self.lib.orc_create())
        self.lib.orc_set_option(self.h, vrt.ORC_OPT_BORROW_HOST_DATA, 1)
        self.lib.orc_set_option(self.h, vrt.ORC_OPT_THREADS, host_cores())  # all host cores, whatever OMP_NUM_THREADS says
        if filter_bits:
            self.lib.orc_set_option(self.h, vrt.ORC_OPT_FILTER_FRAC_BITS, filter_bits)
        self.keep = []
        self.vols = vols
        vs = {k: (_abi.volume_struct(v, 1, self.keep) if v is not None else None) for k, v in vols.items()}
        ptr = lambda k: C.byref(vs[k]) if vs.get(k) is not None else None
        rc = self.lib.orc_sync_volumes(self.h, 0, ptr("E"), ptr("R"), ptr("A"), ptr("Gx"), ptr("Gy"), ptr("Gz"))
        assert rc == 0, self.lib.orc_last_error(self.h)
        self.lut = vrt.oracle_lut(64)
        self.lutv = _abi.volume_struct(self.lut, 1, self.keep)
        self.lights = (_abi.VrLight * len(scene.lights))()
        for i, (p, c) in enumerate(scene.lights):
            self.lights[i].position[:] = p
            self.lights[i].color[:] = c
        self.out = np.zeros((scene.H, scene.W, 3), dtype=np.float32, order="F")

    def render_window(self, step, half, centre=None):
        s = self.scene
        cx, cy = centre if centre is not None else (s.W // 2, s.H // 2)
        win = (C.c_uint64 * 4)(max(cx - half, 0), min(cx + half, s.W), max(cy - half, 0), min(cy + half, s.H))
        self.window = [int(v) for v in win]
        rot = s.rot(step)
        n = len(s.lights)
        rc = self.lib.orc_render(self.h, self.lights if n else None, n, C.byref(self.lutv) if n else None, f32p(s.factors), f32p(s.es),
                                 s.res.ctypes.data_as(C.POINTER(C.c_uint64)), f32p(rot), f32p(s.props), C.c_float(s.thr), f32p(s.color),
                                 win, self.out.ctypes.data)
        assert rc == 0
        st = self.vrt.OrcStats()
        self.lib.orc_get_stats(self.h, C.byref(st))
        return st.samples_marched, st.seconds, st.threads

    def calibrated_half(self, target_s):
        """Window half-size so that one window costs about target_s seconds."""
        n, sec, _ = self.render_window(0, 16)
        rate = n / max(sec, 1e-6)
        per_px = n / (32.0 * 32.0)
        px = max(rate * target_s / max(per_px, 1.0), 1024.0)
        half = int(min(math.sqrt(px) / 2, self.scene.W // 2))
        return max(16, half)


def host_volumes_cpu(args):
    """Generate the workload's volumes on the host with the oracle's bit-identical twins of the
    device generators (no GPU involved): used by the reference arm."""
    vrt, _ = oracle_lib()
    lib = vrt.oracle(fast=False)
    n = args.size
    dims = (C.c_uint64 * 3)(n, n, n)
    E = np.zeros((n, n, n), dtype=np.float32, order="F")
    lib.orc_synth_volume.argtypes = [C.c_int, C.c_uint32, C.POINTER(C.c_uint64), C.c_void_p]
    lib.orc_volume_gradient.argtypes = [C.c_void_p, C.POINTER(C.c_uint64)] + [C.c_void_p] * 3
    lib.orc_synth_volume(0, SEED, dims, E.ctypes.data)
    if CONFIGS[args.config]["volumes"] == "aliased":
        return {"E": E, "R": np.ones((1, 1, 1), dtype=np.float32, order="F"), "A": E}  # the SAME array: Em = Ab aliased
    A = np.asfortranarray(E[::2, ::2, ::2])  # distinct half-resolution absorption volume, normalised
    A = np.asfortranarray((A - A.min()) / (A.max() - A.min()))
    vols = {"E": E, "R": np.asfortranarray(E * E), "A": A}
    if args.gradient == "lookup":
        g = [np.zeros((n, n, n), dtype=np.float32, order="F") for _ in range(3)]
        lib.orc_volume_gradient(E.ctypes.data, dims, g[0].ctypes.data, g[1].ctypes.data, g[2].ctypes.data)
        vols.update({"Gx": g[0], "Gy": g[1], "Gz": g[2]})
    return vols


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    scene = Scene(args)
    need = (5 if args.gradient == "lookup" else 2) * 4.0 * args.size ** 3 / 2 ** 30
    size_used = args.size
    if mem_available_gib() < need * 1.5 + 4:
        args.size = 512
    vols = host_volumes_cpu(args)
    cpu = CpuRender(scene, args, vols)  # asks the oracle for host_cores() threads (torchrun exports OMP_NUM_THREADS=1)
    half = cpu.calibrated_half(10.0)
    n0, s0, _ = cpu.render_window(0, half)
    if s0 < 6.0 and half < scene.W // 2:
        half = int(min(half * math.sqrt(10.0 / max(s0, 1e-3)), scene.W // 2))
    for w in range(args.warmup):
        cpu.render_window(w, 16)
    tot_n, tot_s, threads = 0, 0.0, 1
    for k in range(args.steps):
        n, sec, threads = cpu.render_window(k, half)
        tot_n += n
        tot_s += sec
    value = tot_n / tot_s / 1e9
    sample = "centred %dx%d pixel window of each %dx%d frame over the %d^3 volume, %d frames; software tex3D by the documented filter formula (the model of the hardware filter is 30 %% slower on a CPU)" % (
        2 * half, 2 * half, scene.W, scene.H, args.size, args.steps)
    line = {
        "impl": "reference", "metric": "Gsamples/s (%d^2 image, %d^3 vol)" % (args.image, size_used), "value": value, "unit": "Gsamples/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * tot_s / args.steps,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, scene, size_used),
        "cpu_baseline": {"value": value, "unit": "Gsamples/s", "cores": int(threads), "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "Gsamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)
    return 0


def workload_config(args, scene, size=None):
    size = size or args.size
    return {
        "workload": "%s-shaped: %d^3 %s synthetic ViBE-Z-like E (%s, %s), %d HG light(s), %dx%d mono, F=%g D=%g thr=%g, +1deg/step" % (
            args.config, size, args.dtype, "Em = Ab aliased, constant reflection" if scene.aliased else "+R=E^2, A",
            "Gx,Gy,Gz gradient lookup" if args.gradient == "lookup" else "on-the-fly gradient", len(scene.lights), scene.W, scene.H,
            float(scene.props[1]), float(scene.props[2]), scene.thr),
        "baseline_config": args.config,
        "volume": [size] * 3, "image": [scene.W, scene.H], "gradient": args.gradient, "lights": len(scene.lights),
        "multi_gpu": "image-space tiles %dx%d round-robin over replicated volumes; %s" % (
            args.tile, args.tile, "kernels store their tiles straight into rank 0's image over NVLink (CUDA IPC), one barrier per frame" if args.exchange == "p2p" else "NCCL sum-reduce of partial images to rank 0"),
        "frames_per_launch": frames_per_launch(args.steps, args.batch),
        "fetch": "texture unit (tex3D on fp32 cudaArrays, hardware linear filter as the reference configures it) where the context can, else loads" if getattr(args, "texture_unit", 2) else "loads + fp32 lerps",
        "l2": "inputs (resident volumes, GiBs) are larger than the 126 MB L2; no flush between frames",
    }


# ---- the CUDA arm ---------------------------------------------------------------------------------
def run_b200(args):
    import torch
    import torch.distributed as dist

    from volume_renderer_b200 import _abi, multigpu

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product has no CPU path (use --impl reference for the CPU baseline)")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    lib = _abi.lib()
    ctx = C.c_void_p()
    _abi.check(lib.vr_create_on_device(C.byref(ctx), local))
    stream = torch.cuda.current_stream()
    _abi.check(lib.vr_set_stream(ctx, C.c_void_p(stream.cuda_stream)), ctx)
    _abi.check(lib.vr_set_option(ctx, _abi.OPT_KERNEL_VARIANT, args.variant), ctx)
    _abi.check(lib.vr_set_option(ctx, _abi.OPT_EMPTY_SPACE_SKIP, args.skip), ctx)
    multigpu.configure_partition(ctx, rank, world, args.tile, args.tile)
    _abi.check(lib.vr_set_option(ctx, _abi.OPT_TILE_ORDER, args.order), ctx)
    _abi.check(lib.vr_set_option(ctx, _abi.OPT_WARP_SHAPE, args.wshape), ctx)
    lib.vr_set_option(ctx, _abi.OPT_GRADIENT_INTERLEAVE, args.interleave)  # unchecked: older A/B builds lack the option
    lib.vr_set_option(ctx, _abi.OPT_TEXTURE_UNIT, args.texture_unit)  # unchecked: older A/B builds lack the option
    if args.arith >= 0:
        _abi.check(lib.vr_set_option(ctx, _abi.OPT_ARITHMETIC, args.arith), ctx)
    scene = Scene(args)
    n = args.size
    dt = _abi.DTYPE_F32 if args.dtype == "f32" else _abi.DTYPE_BF16
    tdt = torch.float32 if dt == _abi.DTYPE_F32 else torch.bfloat16
    dims = (C.c_uint64 * 3)(n, n, n)
    dev = torch.device("cuda", local)

    # synthetic volumes generated straight into HBM (torch only owns the memory)
    E = torch.empty(n * n * n, dtype=tdt, device=dev)
    _abi.check(lib.vr_synth_volume(ctx, 0, SEED, dims, dt, C.c_void_p(E.data_ptr())), ctx)
    m = n // 2
    one_host = np.ones(1, dtype=np.float32)
    if scene.aliased:
        Rv = A = None
        keep = [E]
    else:
        Rv = E * E
        A = E.view(n, n, n)[::2, ::2, ::2].contiguous().view(-1)  # distinct half-resolution absorption volume
        A = (A - A.min()) / (A.max() - A.min())
        keep = [E, Rv, A]

    def vol(t, d, stamp):
        v = _abi.VrVolume()
        v.data = t.data_ptr()
        v.dims[0], v.dims[1], v.dims[2] = d
        v.last_update = stamp
        v.dtype = dt
        v.memspace = _abi.MEM_DEVICE
        return v

    if scene.aliased:  # Em = Ab = the same Volume object (example1.m:53-54), default reflection volume = Volume(1)
        vE = vol(E, (n, n, n), 1)
        vA = vE
        vR = _abi.volume_struct((one_host.ctypes.data, (1, 1, 1), _abi.DTYPE_F32, _abi.MEM_HOST), 2)
    else:
        vE, vR, vA = vol(E, (n, n, n), 1), vol(Rv, (n, n, n), 2), vol(A, (m, m, m), 3)
    g = [None, None, None]
    if args.gradient == "lookup":
        G = [torch.empty(n * n * n, dtype=tdt, device=dev) for _ in range(3)]
        _abi.check(lib.vr_volume_gradient(ctx, C.byref(vE), _abi.MEM_DEVICE, *[C.c_void_p(t.data_ptr()) for t in G]), ctx)
        keep += G
        g = [vol(G[i], (n, n, n), 4 + i) for i in range(3)]
    t_up = time.perf_counter()
    _abi.check(lib.vr_sync_volumes(ctx, 0, C.byref(vE), C.byref(vR), C.byref(vA), *[C.byref(x) if x is not None else None for x in g]), ctx)
    torch.cuda.synchronize()
    t_up = time.perf_counter() - t_up

    # LUT: HenyeyGreenstein(64) generated on the GPU, then handed over as the MATLAB array it would be
    lut = np.zeros(64 * 64 * 64, dtype=np.float32)
    _abi.check(lib.vr_henyey_greenstein(64, 0.8, lut.ctypes.data))
    lutv = _abi.VrVolume()
    lutv.data = lut.ctypes.data
    lutv.dims[0] = lutv.dims[1] = lutv.dims[2] = 64
    lutv.last_update = 9
    lutv.dtype = _abi.DTYPE_F32
    lutv.memspace = _abi.MEM_HOST
    nl = len(scene.lights)
    lights = (_abi.VrLight * max(nl, 1))()
    for i, (p, c) in enumerate(scene.lights):
        lights[i].position[:] = p
        lights[i].color[:] = c
    B = frames_per_launch(args.steps, args.batch)
    npx = scene.W * scene.H * 3
    resp = scene.res.ctypes.data_as(C.POINTER(C.c_uint64))
    p2p = world > 1 and args.exchange == "p2p"
    shared = token = d_img = None
    if p2p:
        # the batch's images live in rank 0's HBM, mapped into every rank (CUDA IPC): kernels store their tiles into them
        shared = multigpu.SharedImage(ctx, scene.H, scene.W, rank, world, root=0, count=B)
        out_ptrs = shared.ptrs
        token = torch.zeros(1, dtype=torch.float32, device=dev)
    else:
        d_img = torch.zeros(B * npx, dtype=torch.float32, device=dev)
        out_ptrs = [C.c_void_p(d_img.data_ptr() + i * npx * 4) for i in range(B)]
    props_b = np.tile(scene.props, B).astype(np.float32)

    def frame_device(step, nb=1):
        """nb consecutive frames of the movie (camera +1 degree each) in ONE launch."""
        rots = np.concatenate([scene.rot(step + i).reshape(-1, order="F") for i in range(nb)]).astype(np.float32)
        outs = (C.c_void_p * nb)(*[p.value for p in out_ptrs[:nb]])
        _abi.check(lib.vr_render_batch_device(ctx, lights if nl else None, nl, C.byref(lutv) if nl else None, f32p(scene.factors), f32p(scene.es), resp,
                                              nb, f32p(rots), f32p(props_b), C.c_float(scene.thr), f32p(scene.color), outs,
                                              C.c_void_p(stream.cuda_stream)), ctx)
        if p2p:
            multigpu.frame_barrier(token)  # on the stream, behind the kernel: "batch complete in rank 0's images"
        elif world > 1:
            multigpu.combine_partials(d_img[: nb * npx], 0)

    def stats():
        s = _abi.VrStats()
        _abi.check(lib.vr_get_stats(ctx, C.byref(s)), ctx)
        return s

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident leg: `value` ----
    # (the sampler is started BEFORE the warm-up: nvidia-smi takes a few hundred ms to initialise NVML and holds driver locks
    # meanwhile - inside the timed region that stalls launches; its rows are filtered by time stamp to the timed region)
    sampler = ClockSampler(local) if rank == 0 else None
    for w in range(0, args.warmup, B):
        frame_device(-args.warmup + w, min(B, args.warmup - w))
    if sampler:
        time.sleep(0.3)  # nvidia-smi's start-up ends outside the timed region
    barrier()
    samples = shaded = 0
    kernel_ms = []
    launches = 0
    t0 = time.time()
    barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(stream)
    for k in range(0, args.steps, B):
        frame_device(k, min(B, args.steps - k))
        s = stats()  # synchronises this rank's stream; counters of the launch
        samples += s.samples_marched
        shaded += s.samples_shaded
        kernel_ms.append(s.kernel_ms)
        launches += s.kernel_launches
    ev1.record(stream)
    barrier()
    t1 = time.time()
    clocks = sampler.stop(t0, t1) if sampler else None
    ms = ev0.elapsed_time(ev1)
    tot = torch.tensor([ms, float(samples), float(shaded), float(launches), max(kernel_ms)], dtype=torch.float64, device=dev)
    if world > 1:
        mx = tot.clone()
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
        ms = float(mx[0])
        kmax = float(mx[4])
    else:
        kmax = max(kernel_ms)
    samples_all, shaded_all, launches_all = float(tot[1]), float(tot[2]), int(tot[3])
    value = samples_all / (ms * 1e-3) / 1e9

    # ---- end-to-end leg through the public C ABI call with HOST buffers ----
    # every step: 'sync_volumes' (nothing dirty -> nothing moves, the persistent-volume design of
    # the reference, example2.m) + 'render' with host-side parameters and a device->host read of
    # the frame.  N > 1: partials are reduced to rank 0 first, rank 0 reads the frame back.
    stamp_sync = 100
    # The frames land in ONE host buffer (a ring of two batches).  N = 1: pinned memory of this process.  N > 1: POSIX shared
    # memory mapped and page-locked by every rank - with the column partition (VR_OPT_PART_MODE = 1) each rank's device copies
    # exactly the pixel columns it rendered into the assembled frames over its own PCIe link, and the pipelined batches
    # (vr_render_batch_begin / _end) let those copies run while the next batch is rendered.
    host_frames = None
    Bd = B                                                   # frames per launch of the device-resident leg above
    B = frames_per_launch(args.steps, min(args.e2e_batch, args.batch))  # ... and of this leg's pipelined batches
    props_b = np.tile(scene.props, max(B, Bd)).astype(np.float32)
    if world > 1:
        host_frames = multigpu.SharedHostFrames(scene.H, scene.W, 2 * B, rank, world)
        ring = host_frames.ptrs
        multigpu.configure_partition(ctx, rank, world, args.e2e_tile, args.tile, mode=1)
    else:
        h_ring = torch.zeros(2 * B * npx, dtype=torch.float32).pin_memory()
        ring = [h_ring.data_ptr() + i * npx * 4 for i in range(2 * B)]

    def e2e_batch(step, nb, slot):
        _abi.check(lib.vr_sync_volumes(ctx, stamp_sync, C.byref(vE), C.byref(vR), C.byref(vA), *[C.byref(x) if x is not None else None for x in g]), ctx)
        rots = np.concatenate([scene.rot(step + i).reshape(-1, order="F") for i in range(nb)]).astype(np.float32)
        ptrs = (C.c_void_p * nb)(*ring[slot * B: slot * B + nb])
        _abi.check(lib.vr_render_batch_begin(ctx, lights if nl else None, nl, C.byref(lutv) if nl else None, f32p(scene.factors), f32p(scene.es), resp,
                                             nb, f32p(rots), f32p(props_b), C.c_float(scene.thr), f32p(scene.color), ptrs), ctx)

    def e2e_run(first, count):
        """`count` frames from `first` in pipelined batches of B; returns the lattice samples of this rank's launches."""
        for i, k in enumerate(range(0, count, B)):
            e2e_batch(first + k, min(B, count - k), i & 1)
        _abi.check(lib.vr_render_batch_end(ctx), ctx)  # every frame of this rank is in host memory
        return stats().samples_marched

    e2e_run(-2 * B, 2 * B)  # warm-up: both slots' device images, the pinned ring touched
    barrier()
    te = time.perf_counter()
    e_samples = e2e_run(0, args.steps)
    barrier()  # N > 1: every rank's columns have landed in the shared frames
    te = time.perf_counter() - te
    e = torch.tensor([te, float(e_samples)], dtype=torch.float64, device=dev)
    if world > 1:
        emx = e.clone()
        dist.all_reduce(emx, op=dist.ReduceOp.MAX)
        dist.all_reduce(e, op=dist.ReduceOp.SUM)
        te = float(emx[0])
    e2e_value = float(e[1]) / te / 1e9
    # the mex 'render' command itself (render.cpp:134-277): ONE frame per call into PAGEABLE host memory (an mxArray),
    # sync_volumes before every frame as VolumeRender.m:565-581 does
    single = None
    if world == 1:
        pageable = np.zeros(npx, dtype=np.float32)
        n1 = min(args.steps, 8)

        def one(step):
            _abi.check(lib.vr_sync_volumes(ctx, stamp_sync, C.byref(vE), C.byref(vR), C.byref(vA), *[C.byref(x) if x is not None else None for x in g]), ctx)
            _abi.check(lib.vr_render(ctx, lights if nl else None, nl, C.byref(lutv) if nl else None, f32p(scene.factors), f32p(scene.es), resp,
                                     f32p(scene.rot(step).reshape(-1, order="F").astype(np.float32)), f32p(scene.props), C.c_float(scene.thr), f32p(scene.color),
                                     pageable.ctypes.data_as(C.POINTER(C.c_float))), ctx)

        one(-1)
        s_samples, s_kernel = 0, 0.0
        per_frame = []
        ts = time.perf_counter()
        for k in range(n1):
            tk = time.perf_counter()
            one(k)
            per_frame.append(1e3 * (time.perf_counter() - tk))
            st = stats()
            s_samples += st.samples_marched
            s_kernel += st.kernel_ms
        ts = time.perf_counter() - ts
        med = float(np.median(per_frame))
        single = {"ms_per_frame": med, "ms_per_frame_mean": 1e3 * ts / n1, "value": (s_samples / n1) / (med * 1e-3) / 1e9, "unit": "Gsamples/s", "frames": n1,
                  "kernel_ms_per_frame": s_kernel / n1, "ms_each": [round(x, 2) for x in per_frame],
                  "note": "vr_sync_volumes + vr_render, one frame per call, output into pageable host memory (the MATLAB mex path)"}
    h2d = 9 * 4 + 3 * 4 * 4 + 8 * 2 + nl * 24 + 6 * 40  # rotation, factors/es/props/color, resolution, lights, volume descriptors
    d2h = scene.W * scene.H * 3 * 4

    def finish():
        if host_frames is not None:
            _abi.check(lib.vr_set_option(ctx, _abi.OPT_PART_MODE, 0), ctx)
            host_frames.close(world)
        if shared is not None:
            torch.cuda.synchronize()
            dist.barrier()
            if rank != 0:
                shared.close()  # unmap before the owner frees
            dist.barrier()
            if rank == 0:
                shared.close()
        if world > 1:
            dist.destroy_process_group()

    if rank != 0:
        finish()
        return 0

    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    hbm_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6.65 TB/s (B200_PROFILING.md)"
    # ---- roofline of the dominant kernel, k_raymarch ----
    # The path is an on-chip GATHER: its unit of memory work is one trilinear fetch = 8 taps x 4 B = 32
    # algorithmic bytes; a shaded sample issues FETCHES of them.  achieved = fetches of this run's launches x 32 B
    # / kernel time (CUDA events around each launch, on the launching stream); peak = the same 32 B x the fetch rate of a
    # kernel that ONLY fetches with the march's access pattern out of an L2-sized footprint, measured live below
    # (vr_probe_gather) - the ceiling whose L1 / L2 hit rates match the march's (profiles/r02_probe_hit_rates.md).
    nl_ = len(scene.lights)
    fetches_per_sample = (5 + nl_) if args.gradient == "lookup" else (2 + 6 + nl_)  # em, re (+ gx, gy, gz | 6 central-difference taps) + one LUT fetch per light
    n_launch = max(len(kernel_ms), 1)
    ksum_ms = float(np.sum(kernel_ms))
    kavg_ms = ksum_ms / n_launch
    frames_per_launch_avg = args.steps / float(n_launch)
    per_launch_shaded = shaded / n_launch  # rank 0's launches (a launch = a batch of frames)
    per_launch_marched = samples / n_launch
    fetch_rate = shaded * fetches_per_sample / (ksum_ms * 1e-3)  # trilinear fetches per second inside the kernel
    achieved = fetch_rate * 32.0 / 1e9
    # DRAM traffic of the kernel: not measurable live (needs ncu); taken from this round's committed capture of the same
    # kernel and scene, per frame, scaled to the frames of one launch of THIS run
    prof = DRAM_PROFILE.get((args.size, args.image, args.gradient, args.dtype, 1 if args.texture_unit else 0))
    traffic = prof["bytes_per_frame"] * frames_per_launch_avg if prof else None
    roofline = {
        "bound": "gather", "achieved": achieved, "peak": None, "unit": "GB/s", "frac": None, "traffic": traffic,
        "traffic_source": ("%s (%.3g B per frame x %.1f frames per launch of this run; dram__bytes_read.sum + dram__bytes_write.sum of an ncu --set full capture, not live)"
                           % (prof["profile"], prof["bytes_per_frame"], frames_per_launch_avg)) if prof else None,
        "kernel": "k_raymarch", "kernel_ms_avg": kavg_ms, "launches": n_launch, "frames_per_launch": frames_per_launch_avg,
        "algorithmic_bytes_per_fetch": 32, "fetches_per_shaded_sample": fetches_per_sample,
        "fetches_per_s": fetch_rate, "samples_shaded_per_launch": per_launch_shaded, "samples_marched_per_launch": per_launch_marched,
        "hbm": {"achieved": (traffic / (kavg_ms * 1e-3) / 1e9) if traffic else None, "peak": hbm_peak, "unit": "GB/s", "peak_source": hbm_src,
                "frac": (traffic / (kavg_ms * 1e-3) / 1e9 / hbm_peak) if traffic else None,
                "note": "DRAM bytes of the kernel (profile) / kernel time: the march touches each resident voxel of the frustum about once per frame and serves the rest from L1/L2 - HBM is not the bound"},
        "note": "bound = on-chip gather through the texture unit; achieved / peak in algorithmic bytes (32 B per trilinear fetch); peak filled from the live gather probe",
    }
    line = {
        "metric": "Gsamples/s (%d^2 image, %d^3 vol)" % (args.image, args.size), "value": value, "unit": "Gsamples/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": args.dtype, "data": "synthetic", "config": workload_config(args, scene),
        "frames_per_s": args.steps / (ms * 1e-3), "gsamples_shaded_per_s": shaded_all / (ms * 1e-3) / 1e9,
        "samples_per_frame": samples_all / args.steps, "upload_s": t_up,
        "e2e": {"value": e2e_value, "unit": "Gsamples/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "frames_per_s": args.steps / te,
                "note": ("vr_sync_volumes + vr_render_batch_begin/_end through the C ABI with host parameter and (page-locked) host image buffers, batches of %d frames, "
                         "copies of batch k overlapped with the launch of batch k+1; volumes persistent (uploaded once, %0.2f s)" % (B, t_up)) +
                        ("; N = %d: tile columns %d px wide dealt round-robin, each rank copies its own columns into frames in shared host memory" % (world, args.e2e_tile) if world > 1 else "")},
        "gpu_launches": launches_all, "clocks": clocks, "roofline": roofline,
        "fetch_mode": int(stats().fetch_mode), "arithmetic": int(stats().arithmetic),
    }
    if single is not None:
        line["e2e"]["single_frame"] = single

    # ---- the same frames on the load path (VR_OPT_TEXTURE_UNIT = 0: loads + fp32 lerps, the documented-formula mode) ----
    if world == 1 and args.texture_unit != 0 and not args.no_load_leg:
        try:
            _abi.check(lib.vr_set_option(ctx, _abi.OPT_TEXTURE_UNIT, 0), ctx)
            frame_device(-B, B)  # warm-up
            torch.cuda.synchronize()
            t_k, t_shaded = 0.0, 0
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            for k in range(0, args.steps, B):
                frame_device(k, min(B, args.steps - k))
                st = stats()
                t_shaded += st.samples_shaded
                t_k += st.kernel_ms
            e1.record(stream)
            torch.cuda.synchronize()
            t_s = e0.elapsed_time(e1)
            line["load_path"] = {
                "ms_per_step": t_s / args.steps, "value": samples_all / (t_s * 1e-3) / 1e9, "unit": "Gsamples/s", "kernel_ms_per_frame": t_k / args.steps,
                "gsamples_shaded_per_s": t_shaded / (t_s * 1e-3) / 1e9,
                "note": "VR_OPT_TEXTURE_UNIT = 0: every fetch as eight loads + fp32 lerps (the documented filter formula with fp32 weights, bit-level "
                        "with the CPU oracle's default); same frames, device-resident timing like `value`"}
            _abi.check(lib.vr_set_option(ctx, _abi.OPT_TEXTURE_UNIT, args.texture_unit), ctx)
        except Exception as ex:
            line["load_path"] = {"failed": repr(ex)[:300]}

    # ---- gather ceilings of this box with the march's access pattern (BW_gather, measured live) ----
    if world == 1 and not args.no_gather_probe:
        try:
            gp = gather_probe(lib, ctx, args, len(scene.lights), shaded / (float(np.sum(kernel_ms)) * 1e-3))
            roofline["gather"] = gp
            # peak: the fetch-only kernel with the march's mix of fetches, footprints L2-resident (the march's own L2 hit rate is 97 %)
            roofline["peak"] = gp["ceiling_l2_resident_gsamples_shaded_per_s"] * fetches_per_sample * 32.0
            roofline["frac"] = roofline["achieved"] / roofline["peak"]
            roofline["peak_source"] = "live vr_probe_gather: fetch-only kernels, march access pattern, L2-resident footprints (see roofline.gather; hit rates in profiles/r02_probe_hit_rates.md)"
            roofline["frac_vs_hbm_sized_footprint_ceiling"] = gp["frac"]
        except Exception as ex:
            roofline["gather"] = {"failed": repr(ex)}

    # ---- upload path: a dirty volume from host memory into resident bricks ----
    if world == 1 and not args.no_upload_probe:
        try:
            line["upload"] = upload_probe(args)
        except Exception as ex:
            line["upload"] = {"failed": repr(ex)}

    # ---- cpu_baseline: the oracle on host copies of the same volumes, bounded sample ----
    if world == 1 and not args.no_cpu_baseline:
        try:
            # frame 0 of the movie through the mex path in EVERY arithmetic flavour, for the full-size parity matrix
            one(0)  # (the last render may have been the load-path leg: the counters must be those of this run's own configuration)
            arith_run, fetch_run = int(stats().arithmetic), int(stats().fetch_mode)
            frames = {}
            for a_ in (0, 1, 2):
                _abi.check(lib.vr_set_option(ctx, _abi.OPT_ARITHMETIC, a_), ctx)
                one(0)
                frames[a_] = pageable.reshape((scene.H, scene.W, 3), order="F").copy()
            _abi.check(lib.vr_set_option(ctx, _abi.OPT_ARITHMETIC, arith_run), ctx)
            one(0)  # (leaves the counters of the run's own flavour behind)
            line["cpu_baseline"] = cpu_baseline(args, scene, keep, E, Rv, A, G if args.gradient == "lookup" else None, n, m, frames, fetch_run)
            cb = line["cpu_baseline"]
            if "parity" in cb:
                par = cb.pop("parity")
                names = {0: "exact", 1: "reference", 2: "fast"}
                mine = par.get("flavours", {}).get(names[arith_run], {})
                # headline figures: this run's flavour against the reference's own kernel where it ran, else against the CPU oracle
                head = mine.get("vs_reference_kernel") or mine.get("vs_oracle") or {}
                par.update({"max_abs_norm": head.get("max_abs_norm"), "psnr_db": head.get("psnr_db"), "frac_within_1e-3": head.get("frac_within_1e-3"),
                            "window": head.get("window", "whole frame"), "arithmetic": arith_run,
                            "against": "reference_kernel" if mine.get("vs_reference_kernel") else "oracle", "tolerance": 1e-3})
                line["parity"] = par
            rcb = line["cpu_baseline"].get("reference_cuda", {})
            if "ms_per_frame" in rcb:  # same frames, same lattice: samples per frame from this run's counters
                rcb["value"] = line["samples_per_frame"] / (rcb["ms_per_frame"] * 1e-3) / 1e9
                rcb["unit"] = "Gsamples/s"
                rcb["speedup_of_this_repo_e2e"] = line["e2e"]["value"] / rcb["value"]
        except Exception as ex:  # the bench line must still print
            line["cpu_baseline"] = {"value": None, "unit": "Gsamples/s", "cores": 0, "kind": "port", "sample": "failed: %r" % (ex,)}
    emit(line)
    finish()
    return 0


def gather_probe(lib, ctx, args, n_lights, shaded_per_s):
    """vr_probe_gather (csrc/probe_kernels.cu): trilinear fetches per second of kernels that ONLY gather, with the march's
    access pattern (8x4 bundles of parallel rays 0.7 voxels apart, 0.45-voxel steps), served from L1 / L2 / HBM-sized
    footprints; loads as the march issues them (8 x 4 B, 8 x 16 B) and the texture unit's filtered fetch (the reference's
    path) for comparison.  From those the time a shaded sample's fetches alone would take, i.e. the gather-only ceiling
    in shaded samples per second, and the march's fraction of it."""
    from volume_renderer_b200 import _abi

    def probe(kind, log2_edge):
        f, ms = C.c_double(), C.c_double()
        _abi.check(lib.vr_probe_gather(ctx, kind, log2_edge, C.byref(f), C.byref(ms)), ctx)
        return f.value

    out = {"pattern": "8x4 bundles of parallel rays 0.7 voxel apart, 0.45-voxel steps, one trilinear fetch (8 taps) per step, 2048 threads/SM",
           "unit": "G trilinear fetches/s"}
    for name, kind, e in (("ldg4_L1_32^3", 0, 5), ("ldg4_L2_256^3", 0, 8), ("ldg4_HBM_1024^3", 0, 10), ("ldg16_L2_128^3", 1, 7), ("ldg16_HBM_512^3", 1, 9),
                          ("tex_f32_L1_32^3", 2, 5), ("tex_f32_L2_256^3", 2, 8), ("tex_f32_HBM_1024^3", 2, 10), ("tex_f32x4_L2_128^3", 3, 7), ("tex_f32x4_HBM_512^3", 3, 9),
                          ("smem_taps_staging_free", 4, 8)):
        try:
            out[name] = probe(kind, e) / 1e9
        except Exception as ex:  # e.g. gather from layered arrays refused by the driver: the other probes still count
            out[name] = None
            out.setdefault("failed", {})[name] = repr(ex)[:200]
    # fetches of one shaded sample: gradient volumes -> emission + reflection (8 x 4 B each), one interleaved
    # gradient fetch (8 x 16 B), L LUT fetches (8 x 4 B, 1.1 MiB table: L1/L2); on the fly -> emission + reflection + 6 more + L LUT
    # emission, reflection (+ Gx, Gy, Gz) stream through the volume; the six central-difference fetches of computeGradient
    # re-read the texels of the emission fetch (within half a voxel): L1-sized rate
    nvol, nnear = (5.0, 0.0) if args.gradient == "lookup" else (2.0, 6.0)
    if args.texture_unit:
        # the launch fetched through the texture unit: one filtered fetch per volume and per light (LUT: 1 MiB, L2-sized)
        out["fetch"] = "texture unit"
        t = nvol / out["tex_f32_HBM_1024^3"] + nnear / out["tex_f32_L1_32^3"] + n_lights / out["tex_f32_L2_256^3"]
        t_l2 = nvol / out["tex_f32_L2_256^3"] + nnear / out["tex_f32_L1_32^3"] + n_lights / out["tex_f32_L2_256^3"]
    elif args.gradient == "lookup":
        out["fetch"] = "loads"
        t = 2.0 / out["ldg4_HBM_1024^3"] + 1.0 / out["ldg16_HBM_512^3"] + n_lights / out["ldg4_L1_32^3"]
        t_l2 = 2.0 / out["ldg4_L2_256^3"] + 1.0 / out["ldg16_L2_128^3"] + n_lights / out["ldg4_L1_32^3"]
    else:
        out["fetch"] = "loads"
        t = 2.0 / out["ldg4_HBM_1024^3"] + (6.0 + n_lights) / out["ldg4_L1_32^3"]
        t_l2 = 2.0 / out["ldg4_L2_256^3"] + (6.0 + n_lights) / out["ldg4_L1_32^3"]
    out["ceiling_gsamples_shaded_per_s"] = 1.0 / t  # volumes at their real size (HBM-sized footprints), LUT from L1
    out["ceiling_l2_resident_gsamples_shaded_per_s"] = 1.0 / t_l2
    out["achieved_gsamples_shaded_per_s"] = shaded_per_s / 1e9  # inside the kernel (kernel time, not step time)
    out["frac"] = out["achieved_gsamples_shaded_per_s"] * t
    out["note"] = ("ceiling = a kernel that issues ONLY the fetches of a shaded sample (no shading arithmetic, no empty-space test); "
                   "the march adds ~600-800 instructions of illumination per sample, so issue slots bind before the fetch path does")
    return out


def upload_probe(args):
    """vr_sync_volumes of one dirty N^3 fp32 volume from pageable and from pinned host memory, and vr_update_box of half of
    it (the example3.m:135-161 pattern), in a context of its own; wall clock around the call, best of 2 after a warm-up."""
    import torch

    from volume_renderer_b200 import _abi

    n = min(args.size, 1024)
    if mem_available_gib() < 3 * 4.0 * n ** 3 / 2 ** 30 + 8:
        return {"skipped": "host memory"}
    lib = _abi.lib()
    ctx = C.c_void_p()
    _abi.check(lib.vr_create(C.byref(ctx)))
    one = np.ones(1, dtype=np.float32)
    out = {"volume": "%d^3 fp32" % n, "gib": n ** 3 * 4 / 2 ** 30}

    def sync(ptr, stamp, T):
        E = _abi.volume_struct((ptr, (n, n, n), _abi.DTYPE_F32, _abi.MEM_HOST), stamp)
        R = _abi.volume_struct((one.ctypes.data, (1, 1, 1), _abi.DTYPE_F32, _abi.MEM_HOST), 1)
        t0 = time.perf_counter()
        _abi.check(lib.vr_sync_volumes(ctx, C.c_uint64(T), C.byref(E), C.byref(R), C.byref(E), None, None, None), ctx)
        torch.cuda.synchronize()
        return time.perf_counter() - t0

    pageable = np.full(n ** 3, 0.25, dtype=np.float32)
    pinned = torch.full((n ** 3,), 0.25, dtype=torch.float32).pin_memory()
    for name, ptr in (("pageable", pageable.ctypes.data), ("pinned", pinned.data_ptr())):
        stamp, T, ts = 10, 0, []
        for _ in range(3):
            ts.append(sync(ptr, stamp, T))
            T, stamp = stamp, stamp + 10  # the next stamp is newer than the last sync: dirty again
        out[name + "_GBps"] = n ** 3 * 4 / min(ts[1:]) / 1e9
    lo, hi = (C.c_uint64 * 3)(0, 0, 0), (C.c_uint64 * 3)(n, n, n // 2)
    E = _abi.volume_struct((pageable.ctypes.data, (n, n, n), _abi.DTYPE_F32, _abi.MEM_HOST), 1000)
    ts = []
    for _ in range(3):
        t0 = time.perf_counter()
        _abi.check(lib.vr_update_box(ctx, 0, C.byref(E), lo, hi), ctx)
        torch.cuda.synchronize()
        ts.append(time.perf_counter() - t0)
    out["half_volume_update_ms"] = min(ts[1:]) * 1e3
    out["whole_volume_pageable_ms"] = n ** 3 * 4 / out["pageable_GBps"] / 1e6
    _abi.check(lib.vr_destroy(ctx))
    return out


def cpu_baseline(args, scene, keep, E, Rv, A, G, n, m, gpu_frames=None, fetch_mode=1):
    import torch

    need = (5 if G is not None else (1 if scene.aliased else 2)) * 4.0 * n ** 3 / 2 ** 30
    if mem_available_gib() < need * 1.3 + 8 or E.dtype != torch.float32:
        # not enough host memory for copies of the full volumes: same scene on a 512^3 twin
        a2 = argparse.Namespace(**vars(args))
        a2.size = 512
        vols = host_volumes_cpu(a2)
        note = "512^3 twin of the volume (host memory too small for the %d^3 copies)" % n
    else:
        def host(t, d):  # the flat tensors are in MATLAB order (dims[0] fastest)
            return t.cpu().numpy().reshape(d, order="F")

        if scene.aliased:
            hE = host(E, (n, n, n))
            vols = {"E": hE, "R": np.ones((1, 1, 1), dtype=np.float32, order="F"), "A": hE}
        else:
            vols = {"E": host(E, (n, n, n)), "R": host(Rv, (n, n, n)), "A": host(A, (m, m, m))}
        if G is not None:
            vols.update({"Gx": host(G[0], (n, n, n)), "Gy": host(G[1], (n, n, n)), "Gz": host(G[2], (n, n, n))})
        note = "the same %d^3 volumes copied back from the GPU" % n
    cpu = CpuRender(scene, args, vols)
    half = cpu.calibrated_half(args.cpu_seconds)
    nsm, sec, threads = cpu.render_window(0, half)
    if sec < 0.6 * args.cpu_seconds and half < scene.W // 2:  # the small probe window over-estimates the cost per pixel: once more, rescaled
        half = int(min(half * math.sqrt(args.cpu_seconds / max(sec, 1e-3)), scene.W // 2))
        nsm, sec, threads = cpu.render_window(0, half)
    out = {"value": nsm / sec / 1e9, "unit": "Gsamples/s", "cores": int(threads), "kind": "port",
           "sample": "frame 0, centred %dx%d pixel window of the %dx%d image, %s; %.1f s of CPU time, %d samples; software tex3D by the documented filter formula (the model of the hardware filter is 30 %% slower on a CPU)" % (2 * half, 2 * half, scene.W, scene.H, note, sec, nsm)}
    # ---- full-size parity matrix (north_star's bar: 1e-3 of the image maximum).  Checkers: (i) the reference's OWN kernel on
    # this GPU (oracle/_ref/libref_cuda.so), the whole frame; (ii) the checker build of the oracle (strict IEEE), filtering the
    # way the product fetched, on a window of the very frame the GPU rendered.  Every arithmetic flavour against both, and the
    # two checkers against each other: what the reference's kernel differs from its own restated math by is the floor.
    # north_star: "a baseline from the reference's CUDA kernel rebuilt standalone via nvcc is added where it builds" -
    # oracle/_ref/libref_cuda.so = the reference's own d_render + upload code on texture objects (oracle/ref_cuda)
    ref_frame = None
    try:
        if note.startswith("512^3 twin"):
            raise RuntimeError("host memory too small for copies of the full volumes")
        out["reference_cuda"], ref_frame = reference_cuda_baseline(cpu, scene, vols)
    except Exception as ex:
        out["reference_cuda"] = {"unavailable": repr(ex)}
    if gpu_frames and not note.startswith("512^3 twin"):
        try:
            def metrics(a, b, peak, extra):
                d = np.abs(a.astype(np.float64) - b.astype(np.float64)) / peak
                mse = float(np.mean(d ** 2))
                worst = d.max(axis=2)
                m_ = {"max_abs_norm": float(d.max()), "psnr_db": (float("inf") if mse == 0 else 10.0 * math.log10(1.0 / mse)),
                      "frac_within_1e-3": float((worst <= 1e-3).mean()), "pixels_beyond_1e-3": int((worst > 1e-3).sum()), "pixels": int(worst.size)}
                m_.update(extra)
                return m_

            chk = CpuRender(scene, args, vols, fast=False, filter_bits=9 if fetch_mode == 1 else 0)
            ph = max(16, min(half, int(half * math.sqrt(8.0 / max(sec, 1e-3)))))
            t0 = time.perf_counter()
            chk.render_window(0, ph)
            secs = time.perf_counter() - t0
            x0, x1, y0, y1 = chk.window
            win = (slice(y0, y1), slice(x0, x1), slice(None))
            orc_w = chk.out[win]
            peak_o = float(chk.out.max()) or 1.0
            names = {0: "exact", 1: "reference", 2: "fast"}
            par = {"flavours": {}, "tolerance": 1e-3,
                   "checkers": {"oracle": "oracle/liboracle.so (strict IEEE build), %s, frame 0 of this run's movie, %s; window [x0,x1,y0,y1] = %s, normalised by the window's maximum %.6g, %.1f s"
                                          % ("texture filter as measured on the B200 texture unit (ORC_OPT_FILTER_FRAC_BITS = 9)" if fetch_mode == 1 else "fp32 filter weights (documented formula)",
                                             note, [x0, x1, y0, y1], peak_o, secs),
                                "reference_kernel": "the reference's own d_render (oracle/_ref/libref_cuda.so: volumeRender_kernel.cu on texture objects), same volumes and camera, whole frame, normalised by its maximum"}}
            peak_r = (float(ref_frame.max()) or 1.0) if ref_frame is not None else None
            for a_, fr in sorted(gpu_frames.items()):
                e_ = {"vs_oracle": metrics(fr[win], orc_w, peak_o, {"window": [x0, x1, y0, y1]})}
                if ref_frame is not None:
                    e_["vs_reference_kernel"] = metrics(fr, ref_frame, peak_r, {"window": "whole frame"})
                par["flavours"][names[a_]] = e_
            if ref_frame is not None:
                par["reference_kernel_vs_oracle"] = metrics(ref_frame[win], orc_w, peak_o, {"window": [x0, x1, y0, y1]})
            out["parity"] = par
        except Exception as ex:
            out["parity"] = {"failed": repr(ex)[:300]}
    return out


def reference_cuda_baseline(cpu, scene, vols):
    """Whole frames of the same scene through the reference's own vr::render (cudaMalloc + memset + d_render +
    sync + D2H, volumeRender.cpp:233-300) on this GPU; uploads (cudaMemcpy3D from pageable memory) not timed."""
    rc = cpu.vrt.ref_cuda()
    if rc is None:
        raise RuntimeError("oracle/_ref/libref_cuda.so not present")
    _abi = cpu.abi
    keep = []
    vs = {k: (_abi.volume_struct(v, 1, keep) if v is not None else None) for k, v in vols.items()}
    ptr = lambda k: C.byref(vs[k]) if vs.get(k) is not None else None
    assert rc.refcuda_sync_volumes(0, ptr("E"), ptr("R"), ptr("A"), ptr("Gx"), ptr("Gy"), ptr("Gz")) == 0
    n = len(scene.lights)
    out = np.zeros((scene.H, scene.W, 3), dtype=np.float32, order="F")
    times = []
    frame0 = None
    for k in range(9):  # one warm-up + 8 timed frames, median
        rot = scene.rot(k)
        t = time.perf_counter()
        assert rc.refcuda_render(cpu.lights if n else None, n, C.byref(cpu.lutv) if n else None, f32p(scene.factors), f32p(scene.es),
                                 scene.res.ctypes.data_as(C.POINTER(C.c_uint64)), f32p(rot), f32p(scene.props), C.c_float(scene.thr), f32p(scene.color),
                                 out.ctypes.data) == 0
        times.append(time.perf_counter() - t)
        if k == 0:
            frame0 = out.copy()  # frame 0 of the movie as the reference's own kernel renders it
    # the lattice of frame 0 is the one the oracle counted per pixel above; count it for the full frame with the same rule
    return {"kind": "shimmed reference: the reference's d_render/upload code on texture objects, nvcc sm_100a, unmodified launch shape (16x16 blocks)",
            "ms_per_frame": 1e3 * float(np.median(times[1:])), "ms_per_frame_min": 1e3 * min(times[1:]), "ms_per_frame_max": 1e3 * max(times[1:]),
            "frames": len(times) - 1, "first_frame_ms": 1e3 * times[0], "statistic": "median of %d frames (camera +1 degree each) after one warm-up frame" % (len(times) - 1)}, frame0


def run_example3(args):
    """BASELINE configs[3], examples/example3.m: a two-channel scene - the main channel E is faded on a half-space mask
    every frame (example3.m:135-161: the reference re-uploads the whole volume, here vr_update_box moves the changed half),
    the structure channel S (sparse tubes, green, fEm 0.5; :199-208) is rendered by a second renderer - as an off-axis stereo
    pair (CameraXOffset 0.06, RedCyan; :105-108) with three lights (:76 + example1.m:36), 2048 x 2048 per eye.  Per frame:
    update the half volume from host memory, render both eyes of both channels (one batch launch per channel), add the
    channels (example3.m:239), crop + compose the anaglyph (VolumeRender.m:286-306) on the device, read the frame back."""
    import torch

    from volume_renderer_b200 import _abi

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product has no CPU path")
    torch.cuda.set_device(0)
    dev = torch.device("cuda", 0)
    lib = _abi.lib()
    n, W = args.size, (args.image or 2048)
    H = W
    stream = torch.cuda.current_stream()
    ctxs = []
    for _ in range(2):
        c = C.c_void_p()
        _abi.check(lib.vr_create_on_device(C.byref(c), 0))
        _abi.check(lib.vr_set_stream(c, C.c_void_p(stream.cuda_stream)), c)  # one stream: the channels' launches and the epilogue are ordered
        if args.arith >= 0:
            _abi.check(lib.vr_set_option(c, _abi.OPT_ARITHMETIC, args.arith), c)
        ctxs.append(c)
    main, struct = ctxs
    dims = (C.c_uint64 * 3)(n, n, n)
    E = torch.empty(n ** 3, dtype=torch.float32, device=dev)
    S = torch.empty(n ** 3, dtype=torch.float32, device=dev)
    _abi.check(lib.vr_synth_volume(main, 0, SEED, dims, _abi.DTYPE_F32, C.c_void_p(E.data_ptr())), main)
    _abi.check(lib.vr_synth_volume(struct, 1, SEED, dims, _abi.DTYPE_F32, C.c_void_p(S.data_ptr())), struct)
    # the main channel lives in HOST memory (MATLAB rewrites it there): two page-locked versions with different fade factors
    # on the half space z < n/2 stand for the per-frame `Data(mask) = emission(mask) * factor` of example3.m:151
    half = n ** 3 // 2
    hosts = []
    for factor in (0.8, 0.4):
        h = E.cpu()
        h[:half] *= factor  # z is the slowest axis: the first half of the buffer is the half space
        hosts.append(h.pin_memory())
    del E
    one_host = np.ones(1, dtype=np.float32)

    def hostvol(i, stamp):
        return _abi.volume_struct((hosts[i].data_ptr(), (n, n, n), _abi.DTYPE_F32, _abi.MEM_HOST), stamp)

    vR = _abi.volume_struct((one_host.ctypes.data, (1, 1, 1), _abi.DTYPE_F32, _abi.MEM_HOST), 2)
    vS = _abi.volume_struct((S.data_ptr(), (n, n, n), _abi.DTYPE_F32, _abi.MEM_DEVICE), 3)
    vM = hostvol(0, 10)
    t_up = time.perf_counter()
    _abi.check(lib.vr_sync_volumes(main, 0, C.byref(vM), C.byref(vR), C.byref(vM), None, None, None), main)
    _abi.check(lib.vr_sync_volumes(struct, 0, C.byref(vS), C.byref(vR), C.byref(vS), None, None, None), struct)
    torch.cuda.synchronize()
    t_up = time.perf_counter() - t_up
    lut = np.zeros(64 ** 3, dtype=np.float32)
    _abi.check(lib.vr_henyey_greenstein(64, 0.8, lut.ctypes.data))
    lutv = _abi.volume_struct((lut.ctypes.data, (64, 64, 64), _abi.DTYPE_F32, _abi.MEM_HOST), 9)
    lights_spec = [([-15, 15, 0], [0.5, 0.5, 0.5])] + EX1_LIGHTS  # example3.m:76 + example1.m:36
    lights = (_abi.VrLight * 3)()
    for i, (p_, c_) in enumerate(lights_spec):
        lights[i].position[:] = p_
        lights[i].color[:] = c_
    F, D, thr, base = 4.5, 6.0, 0.95, 0.06 / 2  # example3.m:66-73,105
    delta = int(math.floor(base * W / 2.0 + 0.5))  # VolumeRender.m:277-282
    Wp = W + delta
    res_wide = np.array([H, Wp], dtype=np.uint64)
    es = np.array([1, 1, 1], dtype=np.float32)
    R0 = rotation(-15, 15, 15, rotation(90, 0, 0))
    props = np.array([base, F, D, -base, F, D], dtype=np.float32)  # right eye, left eye (VolumeRender.m:284-285)
    chan = [(np.array([1.0, 1.0, 1.0], dtype=np.float32), np.array([1, 1, 1], dtype=np.float32)),    # main: [fEm, fRe, fAb], color
            (np.array([0.5, 1.0, 1.0], dtype=np.float32), np.array([0, 1, 0], dtype=np.float32))]    # structure: example3.m:199-208
    npw = H * Wp * 3
    eyes = [torch.zeros(2 * npw, dtype=torch.float32, device=dev) for _ in ctxs]  # per channel: right eye, left eye
    summed = torch.zeros(2 * npw, dtype=torch.float32, device=dev)
    out_dev = torch.zeros(H * W * 3, dtype=torch.float32, device=dev)
    out_host = torch.zeros(H * W * 3, dtype=torch.float32).pin_memory()
    lo, hi = (C.c_uint64 * 3)(0, 0, 0), (C.c_uint64 * 3)(n, n, n // 2)
    resw = res_wide.ctypes.data_as(C.POINTER(C.c_uint64))
    counters = {"samples": 0, "shaded": 0, "upload_ms": 0.0, "kernel_ms": 0.0, "launches": 0}

    def frame(k, count=True):
        nonlocal vM
        vM = hostvol(k & 1, 100 + k)  # the other fade factor, a new timestamp
        _abi.check(lib.vr_update_box(main, 0, C.byref(vM), lo, hi), main)
        st = _abi.VrStats()
        _abi.check(lib.vr_get_stats(main, C.byref(st)), main)
        up_ms = st.upload_ms
        rot = np.asfortranarray(np.flipud(rotation(0, float(k), 0, R0)).astype(np.float32)).reshape(-1, order="F")
        rots = np.concatenate([rot, rot]).astype(np.float32)
        for ci, (c, v) in enumerate(((main, vM), (struct, vS))):
            fac, col = chan[ci]
            _abi.check(lib.vr_sync_volumes(c, 100 + k, C.byref(v), C.byref(vR), C.byref(v), None, None, None), c)  # nothing dirty: moves nothing
            outs = (C.c_void_p * 2)(eyes[ci].data_ptr(), eyes[ci].data_ptr() + 4 * npw)
            _abi.check(lib.vr_render_batch_device(c, lights, 3, C.byref(lutv), f32p(fac), f32p(es), resw, 2, f32p(rots), f32p(props), C.c_float(thr), f32p(col), outs,
                                                  C.c_void_p(stream.cuda_stream)), c)
        pair = (C.c_void_p * 2)(eyes[0].data_ptr(), eyes[1].data_ptr())
        _abi.check(lib.vr_compose_partials_device(main, 2, pair, 2 * npw, C.c_void_p(summed.data_ptr()), C.c_void_p(stream.cuda_stream)), main)  # main + structure
        _abi.check(lib.vr_compose_stereo_device(main, C.c_void_p(summed.data_ptr() + 4 * npw), C.c_void_p(summed.data_ptr()), resw, delta, 0,
                                                C.c_void_p(out_dev.data_ptr()), C.c_void_p(stream.cuda_stream)), main)  # (left, right) -> red-cyan
        out_host.copy_(out_dev, non_blocking=True)
        torch.cuda.synchronize()
        if count:
            for c in ctxs:
                _abi.check(lib.vr_get_stats(c, C.byref(st)), c)
                counters["samples"] += st.samples_marched
                counters["shaded"] += st.samples_shaded
                counters["kernel_ms"] += st.kernel_ms
                counters["launches"] += st.kernel_launches
            counters["upload_ms"] += up_ms

    sampler = ClockSampler(0)  # before the warm-up, see run_b200
    for w in range(args.warmup):
        frame(-1 - w, count=False)
    t0 = time.time()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(stream)
    tw = time.perf_counter()
    for k in range(args.steps):
        frame(k)
    tw = time.perf_counter() - tw
    ev1.record(stream)
    torch.cuda.synchronize()
    t1 = time.time()
    clocks = sampler.stop(t0, t1)
    K = args.steps
    line = {
        "metric": "Gsamples/s (two channels, stereo pair, %d^2 per eye, %d^3 vol)" % (W, n), "value": counters["samples"] / tw / 1e9, "unit": "Gsamples/s",
        "n_gpus": 1, "steps": K, "warmup": args.warmup, "ms_per_step": 1e3 * tw / K, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic", "frames_per_s": K / tw, "gsamples_shaded_per_s": counters["shaded"] / tw / 1e9,
        "kernel_ms_per_frame": counters["kernel_ms"] / K, "update_box_ms_per_frame": counters["upload_ms"] / K, "gpu_launches": counters["launches"] + 2 * K,
        "e2e": {"value": counters["samples"] / tw / 1e9, "unit": "Gsamples/s", "h2d_bytes_per_step": n ** 3 * 2, "d2h_bytes_per_step": H * W * 3 * 4,
                "note": "the step IS the end-to-end call sequence: vr_update_box from page-locked host memory, 2 x (vr_sync_volumes + vr_render_batch_device), "
                        "vr_compose_partials_device, vr_compose_stereo_device, read-back of the anaglyph; wall clock"},
        "config": {"workload": "example3-shaped: main channel %d^3 fp32 (host-resident, half of it rewritten every frame -> vr_update_box) + structure channel (sparse tubes, green), "
                               "3 HG lights, on-the-fly gradient, off-axis stereo pair CameraXOffset 0.06 (delta %d px), %dx%d per eye, F=4.5 D=6 thr=0.95, RedCyan compose on the device" % (n, delta, W, H),
                   "baseline_config": "example3", "volume": [n] * 3, "image": [W, H], "lights": 3, "gradient": "compute", "volume_upload_s": t_up,
                   "l2": "inputs (resident volumes, GiBs) are larger than the 126 MB L2; no flush between frames"},
        "clocks": clocks,
    }
    emit(line)
    for c in ctxs:
        _abi.check(lib.vr_destroy(c))
    return 0


def volume_cpu_baseline(args):
    """The oracle on the 512^3 analogue of the huge volume (config 5 is CPU-timed on the 512^3 analogue): same
    generator, same camera, 1 light, on-the-fly gradient, a centred window; all host cores."""
    a2 = argparse.Namespace(**vars(args))
    a2.size, a2.config, a2.gradient, a2.lights, a2.image = 512, "example1", "compute", 1, args.image
    scene = Scene(a2)
    cpu = CpuRender(scene, a2, host_volumes_cpu(a2))
    half = cpu.calibrated_half(args.cpu_seconds)
    nsm, sec, threads = cpu.render_window(0, half)
    return {"value": nsm / sec / 1e9, "unit": "Gsamples/s (lattice)", "cores": int(threads), "kind": "port",
            "sample": "512^3 fp32 analogue of the volume (same generator and camera), centred %dx%d window of the %dx%d image, %.1f s of CPU time, %d lattice samples" % (
                2 * half, 2 * half, scene.W, scene.H, sec, nsm)}


def run_volume(args):
    """BASELINE configs[4]: a volume too big to replicate, bricks sharded over the ranks (one brick per
    GPU, generated straight into its HBM), example1 camera, 1 light, on-the-fly gradient; per frame an
    opacity pass, an all-gather of H*W floats per brick, a shading pass whose kernel stores the brick's
    share into rank 0's HBM over NVLink, and the composite on rank 0."""
    import torch
    import torch.distributed as dist

    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from volume_renderer_b200 import _abi, api, multigpu

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    args.image = args.image or 2048
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    lib = _abi.lib()
    n = args.size
    dims = (n, n, n)
    bricks = multigpu.brick_grid(dims, multigpu.splits_for(world))
    lo, hi = bricks[rank]
    halo = 4
    dt = _abi.DTYPE_F32 if args.dtype == "f32" else _abi.DTYPE_BF16
    es = 4 if dt == _abi.DTYPE_F32 else 2
    r = api.VolumeRender()
    ctx = r.objectHandle
    _abi.check(lib.vr_set_stream(ctx, C.c_void_p(torch.cuda.current_stream().cuda_stream)), ctx)  # torch's stream, also for timing events
    if world > 1:
        r.setRegion(lo, hi, halo)
        if args.vs_slices:
            r.setOption(_abi.OPT_VS_SLICES, args.vs_slices)
    r.setOption(_abi.OPT_TEXTURE_UNIT, args.texture_unit)
    r.setOption(_abi.OPT_TILE_ORDER, args.order)
    if args.arith >= 0:
        r.setOption(_abi.OPT_ARITHMETIC, args.arith)
    for kv in args.opt:
        r.setOption(int(kv.split("=")[0]), int(kv.split("=")[1]))
    blo = [max(l - halo, 0) for l in lo] if world > 1 else [0, 0, 0]
    bhi = [min(h + halo, d) for h, d in zip(hi, dims)] if world > 1 else list(dims)
    nvox = int(np.prod([h - l for l, h in zip(blo, bhi)]))
    buf = C.c_void_p()
    _abi.check(lib.vr_device_alloc(ctx, nvox * es, C.byref(buf)), ctx)
    u3 = lambda v: (C.c_uint64 * 3)(*[int(x) for x in v])
    t_gen = time.perf_counter()
    _abi.check(lib.vr_synth_volume_region(ctx, 0, SEED, u3(dims), u3(blo), u3(bhi), dt, buf), ctx)
    t_gen = time.perf_counter() - t_gen
    E = api.DeviceVolume(buf.value, dims, dt, _abi.MEM_DEVICE_REGION if world > 1 else _abi.MEM_DEVICE)
    lut = api.HenyeyGreenstein(64)
    r.VolumeIllumination = api.Volume(lut)
    r.LightSources = [api.LightSource([500, 1000, 550], [0, 1, 1])]
    r.ElementSizeUm = [1, 1, 1]
    r.FocalLength, r.DistanceToObject, r.OpacityThreshold = 3.0, 6.0, 0.9
    r.ImageResolution = [args.image, args.image]
    r.VolumeEmission = E
    r.VolumeAbsorption = E
    r.FactorEmission, r.FactorAbsorption, r.FactorReflection = 1.0, 0.6, 0.4
    r.Color = [1, 1, 0]
    R0 = rotation(125, 25, 0)
    t_up = time.perf_counter()
    r.syncVolumes()
    torch.cuda.synchronize()
    t_up = time.perf_counter() - t_up
    _abi.check(lib.vr_device_free(ctx, buf), ctx)  # only the resident bricks stay
    r.VolumeEmission.TimeLastUpdate = 1  # keep the (now dangling) source clean: nothing is re-read
    r.VolumeAbsorption.TimeLastUpdate = 1
    if world > 1:
        vs = multigpu.VolumeSpaceRank(r, bricks, rank, exchange=args.exchange)
        frame = lambda k: (setattr(r, "RotationMatrix", rotation(0, float(k), 0, R0)), vs.render(download=False))
    else:
        d_img = torch.zeros(args.image * args.image * 3, dtype=torch.float32, device=dev)

        def frame(k):
            r.RotationMatrix = rotation(0, float(k), 0, R0)
            a, keep = r._packed_mono_args()
            _abi.check(lib.vr_render_device(ctx, *a, C.c_void_p(d_img.data_ptr()), None), ctx)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    sampler = ClockSampler(local) if rank == 0 else None  # before the warm-up, see run_b200
    for w in range(args.warmup):
        frame(-1 - w)
    barrier()
    if sampler:
        time.sleep(0.3)  # let nvidia-smi finish its start-up outside the timed region (three warm-up frames are 70 ms)
    barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    shaded = marched = 0
    kernel_ms = 0.0
    if world > 1:
        vs.timing = []
    t0 = time.time()
    barrier()
    ev0.record()
    dbg = os.environ.get("BENCH_DEBUG_TIMES")
    for k in range(args.steps):
        tf0 = time.perf_counter()
        frame(k)
        tf1 = time.perf_counter()
        st = r.stats()  # counters of this rank's shading pass
        if dbg:
            print("[times] rank %d frame %d: frame() %.2f ms, stats() %.2f ms, kernel %.2f ms, slices %d" % (rank, k, (tf1 - tf0) * 1e3, (time.perf_counter() - tf1) * 1e3, st["kernel_ms"], st["vs_slices"]), file=sys.stderr, flush=True)
        shaded += st["samples_shaded"]
        marched += st["samples_marched"]
        kernel_ms += st["kernel_ms"]
        vs_slices = int(st["vs_slices"])
    ev1.record()
    barrier()
    t1 = time.time()
    clocks = sampler.stop(t0, t1) if sampler else None
    ms = ev0.elapsed_time(ev1)
    phases = vs.phase_ms() if world > 1 else {"shading_pass": kernel_ms}
    if world > 1:
        vs.timing = None
    fetch_mode, arith = r.stats()["fetch_mode"], r.stats()["arithmetic"]
    names = ["opacity_pass", "all_gather", "shading_pass", "barrier_and_compose"]
    tot = torch.tensor([ms, float(shaded), float(marched), kernel_ms] + [phases.get(p_, 0.0) for p_ in names], dtype=torch.float64, device=dev)
    per_rank = None
    if world > 1:
        mx = tot.clone()
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        allr = [torch.zeros_like(tot) for _ in range(world)]
        dist.all_gather(allr, tot)
        per_rank = [{"rank": i, "shaded_per_frame": float(t[1]) / args.steps, "shading_kernel_ms_per_frame": float(t[3]) / args.steps,
                     "opacity_pass_ms_per_frame": float(t[4]) / args.steps} for i, t in enumerate(allr)]
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
        ms = float(mx[0])
        kmax = float(mx[3])
        phase_max = {p_: float(mx[4 + i]) / args.steps for i, p_ in enumerate(names)}
    else:
        kmax = kernel_ms
        phase_max = {"shading_pass": kernel_ms / args.steps}
    # ---- end to end: the same frames, each read back to host memory on rank 0 (wall clock, max over ranks) ----
    barrier()
    te = time.perf_counter()
    e_shaded = 0
    for k in range(args.steps):
        if world > 1:
            r.RotationMatrix = rotation(0, float(k), 0, R0)
            img = vs.render(download=True)
        else:
            frame(k)
            img = d_img.cpu()
        e_shaded += r.stats()["samples_shaded"]
    barrier()
    te = time.perf_counter() - te
    e = torch.tensor([te, float(e_shaded)], dtype=torch.float64, device=dev)
    if world > 1:
        emx = e.clone()
        dist.all_reduce(emx, op=dist.ReduceOp.MAX)
        dist.all_reduce(e, op=dist.ReduceOp.SUM)
        te = float(emx[0])
    res_b = r.stats()["bytes_resident"]
    if rank == 0:
        # roofline of the shading pass (the dominant kernel): trilinear fetches x 32 algorithmic bytes / the slowest rank's kernel time,
        # against the fetch-only ceiling of the fetch path that ran (loads for bf16 / region contexts, else the texture unit)
        fetches = 2 + 6 + 1  # emission, reflection, six central-difference taps, one LUT fetch (1 light)
        achieved = float(tot[1]) * fetches * 32.0 / (kmax * 1e-3) / 1e9 / world  # per GPU: the slowest rank bounds the frame
        roofline = {"bound": "gather", "achieved": achieved, "peak": None, "unit": "GB/s", "frac": None, "traffic": None, "kernel": "k_raymarch (shading pass)",
                    "kernel_ms_avg": kmax / args.steps, "fetches_per_shaded_sample": fetches, "algorithmic_bytes_per_fetch": 32,
                    "note": "per GPU: all ranks' shaded samples / ranks, over the SLOWEST rank's shading-pass time (ranks behind the opacity threshold idle)"}
        try:
            f_, m_ = C.c_double(), C.c_double()
            _abi.check(lib.vr_probe_gather(ctx, 2 if fetch_mode == 1 else 0, 8, C.byref(f_), C.byref(m_)), ctx)
            roofline["peak"] = f_.value * 32.0 / 1e9
            roofline["frac"] = achieved / roofline["peak"]
            roofline["peak_source"] = "live vr_probe_gather, %s, L2-resident footprint" % ("tex3D<float>" if fetch_mode == 1 else "eight 4-byte loads per fetch")
        except Exception as ex:
            roofline["peak_source"] = "probe failed: %r" % (ex,)
        line = {
            "metric": "Gsamples/s shaded (%d^2 image, %d^3 %s vol sharded over %d GPU(s))" % (args.image, n, args.dtype, world),
            "value": float(tot[1]) / (ms * 1e-3) / 1e9, "unit": "Gsamples/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": args.dtype, "data": "synthetic",
            "frames_per_s": args.steps / (ms * 1e-3), "samples_shaded_per_frame": float(tot[1]) / args.steps,
            "lattice_samples_visited_per_frame_all_ranks": float(tot[2]) / args.steps,
            "phase_ms_per_frame_max_over_ranks": phase_max, "per_rank": per_rank, "fetch_mode": int(fetch_mode), "arithmetic": int(arith),
            "e2e": {"value": float(e[1]) / te / 1e9, "unit": "Gsamples/s", "frames_per_s": args.steps / te, "h2d_bytes_per_step": 388, "d2h_bytes_per_step": args.image * args.image * 3 * 4,
                    "note": "every frame composited on rank 0 and read back to host memory; wall clock, max over ranks"},
            "roofline": roofline,
            "config": {"workload": "huge-volume-shaped: %d^3 %s synthetic ViBE-Z-like E (Em=Ab), bricks %s (+%d voxel halo) one per GPU, 1 HG light, on-the-fly gradient, %dx%d mono, F=3 D=6 thr=0.9, +1deg/step"
                       % (n, args.dtype, "x".join(str(x) for x in multigpu.splits_for(world)), halo, args.image, args.image), "baseline_config": "huge volume (configs[4])",
                       "multi_gpu": "volume-space sort-last: opacity pass, all-gather of H*W floats per brick (NCCL), shading pass storing shares into rank 0 over NVLink (%s), composite kernel" % args.exchange,
                       "depth_slices": "%d per brick (VR_OPT_VS_SLICES = %d%s): a ray's stretch inside a brick is marched by that many warps" % (vs_slices, args.vs_slices, ", automatic" if args.vs_slices == 0 else ""),
                       "brick_generate_s": t_gen, "brick_upload_s": t_up, "bytes_resident_rank0": res_b,
                       "l2": "inputs (resident bricks, GiBs) are larger than the 126 MB L2; no flush between frames"},
            "clocks": clocks,
            # per rank and frame: two march kernels; with depth slices also the slice-start kernel and a fold kernel per pass
            "gpu_launches": (5 if vs_slices > 1 else 2) * args.steps * world + args.steps,
        }
        if not args.no_cpu_baseline:
            try:
                line["cpu_baseline"] = volume_cpu_baseline(args)
            except Exception as ex:
                line["cpu_baseline"] = {"value": None, "unit": "Gsamples/s", "cores": 0, "kind": "port", "sample": "failed: %r" % (ex,)}
        emit(line)
    if world > 1:
        torch.cuda.synchronize()
        dist.barrier()
        if rank != 0:
            vs.close()
        dist.barrier()
        if rank == 0:
            vs.close()
        dist.destroy_process_group()
    return 0


_JSON_FD = None


def emit(line):
    """The ONE JSON line of the contract, on the process's real stdout."""
    data = (json.dumps(line) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_JSON_FD, data)


if __name__ == "__main__":
    # Libraries print on stdout (NCCL's version banner, for one): everything but the JSON line goes to stderr.
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)
    a = parse()
    if a.impl == "reference":
        os.environ["OMP_NUM_THREADS"] = str(host_cores())  # before libgomp loads: torchrun hands its workers OMP_NUM_THREADS=1
        sys.exit(run_reference(a))
    if a.config == "example3":
        sys.exit(run_example3(a))
    sys.exit(run_volume(a) if a.mode == "volume" else run_b200(a))
