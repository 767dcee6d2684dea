"""Shared test helpers: oracle binding, host emulation binding, synthetic volumes, scenes, metrics.

TEST INFRASTRUCTURE.  The oracle (oracle/liboracle.so) is the checker; it is only ever loaded
from here, from __graft_entry__.smoke() and from bench.py's cpu_baseline / --impl reference.
"""
import ctypes as C
import math
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from volume_renderer_b200 import _abi, api  # noqa: E402

ORACLE_DIR = os.path.join(ROOT, "oracle")
HOSTCXX = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
SEED = 0x5EED01BE


def _run(cmd, cwd):
    r = subprocess.run(cmd, cwd=cwd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        raise RuntimeError("%s failed:\n%s" % (" ".join(cmd), r.stdout))


def build_oracle():
    """Build the CPU oracle and, when the reference tree is present, the reference's own CPU
    sources (oracle/_ref).  On the GPU box only the prebuilt files are used."""
    _run(["make", "-s", "all"], ORACLE_DIR)
    if os.path.isdir("/root/reference/src/C"):
        _run(["make", "-s", "ref"], ORACLE_DIR)


_oracle = {}


def oracle(fast=False):
    name = "liboracle_fast.so" if fast else "liboracle.so"
    if name not in _oracle:
        path = os.path.join(ORACLE_DIR, name)
        if not os.path.exists(path):
            build_oracle()
        lib = C.CDLL(path)
        P = C.POINTER
        vp, f3, u64p = C.c_void_p, P(C.c_float), P(C.c_uint64)
        lib.orc_create.restype = vp
        lib.orc_destroy.argtypes = [vp]
        lib.orc_set_option.argtypes = [vp, C.c_int, C.c_int64]
        lib.orc_sync_volumes.argtypes = [vp, C.c_uint64] + [P(_abi.VrVolume)] * 6
        lib.orc_render.argtypes = [vp, P(_abi.VrLight), C.c_size_t, P(_abi.VrVolume), f3, f3, u64p, f3, f3, C.c_float, f3, u64p, vp]
        lib.orc_get_state.argtypes = [vp, C.c_void_p]
        lib.orc_get_stats.argtypes = [vp, C.c_void_p]
        lib.orc_init_render.argtypes = [u64p, f3, f3, f3, f3]
        lib.orc_henyey_greenstein.argtypes = [C.c_uint32, C.c_float, vp]
        lib.orc_tex3d.argtypes = [vp, C.c_uint64, C.c_uint64, C.c_uint64, C.c_float, C.c_float, C.c_float, C.c_int]
        lib.orc_tex3d.restype = C.c_float
        lib.orc_last_error.argtypes = [vp]
        lib.orc_last_error.restype = C.c_char_p
        _oracle[name] = lib
    return _oracle[name]


class OrcState(C.Structure):
    _fields_ = [("idx_emission", C.c_int32), ("idx_absorption", C.c_int32), ("idx_reflection", C.c_int32),
                ("gradient_method", C.c_int32), ("bound", C.c_int32 * 7), ("uploads", C.c_uint64 * 7),
                ("num_lights", C.c_uint64)]


class OrcStats(C.Structure):
    _fields_ = [("samples_marched", C.c_uint64), ("rays_hit", C.c_uint64), ("seconds", C.c_double), ("threads", C.c_int32)]


ORC_OPT_FILTER_FRAC_BITS = 0
ORC_OPT_ABSORPTION_ROLE_FIX = 3
ORC_OPT_BORROW_HOST_DATA = 100
ORC_OPT_THREADS = 101
ORC_OPT_FAST_EXP = 102


def ref_cpu():
    """The reference's own CPU sources compiled unmodified (oracle/_ref/libref_cpu.so) or None."""
    path = os.path.join(ORACLE_DIR, "_ref", "libref_cpu.so")
    if not os.path.exists(path):
        if not os.path.isdir("/root/reference/src/C"):
            return None
        build_oracle()
    lib = C.CDLL(path)
    P = C.POINTER
    lib.ref_henyey_greenstein.argtypes = [C.c_uint32, C.c_int, C.c_float, C.c_void_p, C.c_char_p, C.c_size_t]
    lib.ref_timestamp.restype = C.c_uint64
    lib.ref_init_render.argtypes = [C.c_void_p, P(C.c_uint64), C.c_int, C.c_uint64, P(C.c_float), C.c_uint64, C.c_uint64,
                                    P(C.c_float), C.c_float, P(C.c_uint64), P(C.c_uint64), P(C.c_float), P(C.c_float),
                                    P(C.c_float), P(C.c_float)]
    lib.ref_volume_equal.argtypes = [C.c_void_p, C.c_uint64, P(C.c_uint64), C.c_void_p, C.c_uint64, P(C.c_uint64)]
    lib.ref_idivup.argtypes = [C.c_size_t, C.c_size_t]
    lib.ref_idivup.restype = C.c_size_t
    return lib


class OracleBackend:
    """The mex command protocol on top of the oracle: same surface as api._CudaBackend."""

    fast = False
    options = {}

    def __init__(self):
        self.lib = oracle(self.fast)

    def new(self):
        h = C.c_void_p(self.lib.orc_create())
        for k, v in self.options.items():
            self.lib.orc_set_option(h, k, v)
        return h

    def delete(self, h):
        self.lib.orc_destroy(h)
        return 0

    def mem_info(self, h):
        return "oracle\n"

    def sync_volumes(self, h, T, em, re, ab, gx=None, gy=None, gz=None):
        keep = []
        vols = [_abi.volume_struct(v.Data, v.TimeLastUpdate, keep) if v is not None else None for v in (em, re, ab, gx, gy, gz)]
        ptrs = [C.byref(v) if v is not None else None for v in vols]
        if self.lib.orc_sync_volumes(h, int(T), *ptrs) != 0:
            raise RuntimeError(self.lib.orc_last_error(h).decode())
        # the oracle keeps copies unless ORC_OPT_BORROW_HOST_DATA is set; keep arrays alive anyway
        self._keep = keep

    def render(self, h, lights, lut, factors, es, res, rot, props, thr, color, window=None):
        H, W = int(res[0]), int(res[1])
        out = np.zeros((H, W, 3), dtype=np.float32, order="F")
        args, keep = api._render_args(lights, lut, factors, es, res, rot, props, thr, color)
        win = None
        if window is not None:
            win = (C.c_uint64 * 4)(*[int(x) for x in window])
        if self.lib.orc_render(h, *args, win, out.ctypes.data) != 0:
            raise RuntimeError(self.lib.orc_last_error(h).decode())
        return out

    def state(self, h):
        s = OrcState()
        self.lib.orc_get_state(h, C.byref(s))
        return s

    def stats(self, h):
        s = OrcStats()
        self.lib.orc_get_stats(h, C.byref(s))
        return s


class OracleVolumeRender(api.VolumeRender):
    """VolumeRender whose mex calls land in the oracle."""

    _backend_cls = OracleBackend


class FastOracleBackend(OracleBackend):
    fast = True


class FastOracleVolumeRender(api.VolumeRender):
    _backend_cls = FastOracleBackend


# ---- host emulation of the device code ---------------------------------------------------------
_emul = None


def host_emul():
    global _emul
    if _emul is None:
        src = os.path.join(ROOT, "tests", "host_emul.cpp")
        out = os.path.join(ROOT, "tests", "_build", "libhost_emul.so")
        os.makedirs(os.path.dirname(out), exist_ok=True)
        deps = [src] + [os.path.join(ROOT, "volume_renderer_b200", "csrc", f) for f in ("raymarch_core.h", "render_setup.h", "vr_types.h")]
        if not os.path.exists(out) or any(os.path.getmtime(d) > os.path.getmtime(out) for d in deps):
            _run([HOSTCXX, "-O2", "-std=c++17", "-fPIC", "-fopenmp", "-ffp-contract=off", "-shared", "-o", out, src], ROOT)
        lib = C.CDLL(out)
        P = C.POINTER
        f3, u64p = P(C.c_float), P(C.c_uint64)
        lib.emul_render.argtypes = [P(_abi.VrVolume)] * 6 + [P(_abi.VrLight), C.c_size_t, f3, f3, u64p, f3, f3, C.c_float, f3,
                                                             C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, u64p, C.c_void_p, u64p]
        _emul = lib
    return _emul


def emul_render(r, layout=1, skip=2, frac_bits=0, resolution=None, xoffset=0.0, variant=0, bricks=None, slices=1):
    """Run a configured api.VolumeRender-like scene `r` (Em = Ab aliased) through the host
    emulation of the device code.  Returns (image HxWx3, stats[3]).  `slices` > 1 (with bricks): the volume-space passes
    cut into depth slices (VR_OPT_VS_SLICES)."""
    lib = host_emul()
    lib.emul_set_slices.argtypes = [C.c_int]
    lib.emul_set_slices.restype = None
    lib.emul_set_slices(int(slices))
    keep = []

    def vs(v):
        return C.byref(_abi.volume_struct(v.Data, v.TimeLastUpdate, keep)) if v is not None and v is not False else None

    g = [r.VolumeGradientX, r.VolumeGradientY, r.VolumeGradientZ]
    g = g if all(v is not False for v in g) else [None, None, None]
    res = resolution or [int(r.ImageResolution[1]), int(r.ImageResolution[0])]
    factors = [r.FactorEmission, r.FactorReflection, r.FactorAbsorption]
    props = [xoffset, r.FocalLength, r.DistanceToObject]
    lights = r.LightSources if r.LightSources is not False else None
    lut = r.VolumeIllumination if r.VolumeIllumination is not False else None
    args, keep2 = api._render_args(lights, lut, factors, r.ElementSizeUm, res, np.flipud(np.asarray(r.RotationMatrix)), props,
                                   r.OpacityThreshold, r.Color)
    H, W = res
    out = np.zeros((H, W, 3), dtype=np.float32, order="F")
    stats = (C.c_uint64 * 3)()
    nb = len(bricks) if bricks else 0
    bv = (C.c_uint64 * (6 * max(nb, 1)))(*[int(v) for box in (bricks or []) for v in (list(box[0]) + list(box[1]))])
    rc = lib.emul_render(vs(r.VolumeEmission), vs(r.VolumeReflection), vs(g[0]), vs(g[1]), vs(g[2]), args[2], args[0], args[1],
                         *args[3:], layout, skip, frac_bits, variant, nb, bv, out.ctypes.data, stats)
    lib.emul_set_slices(1)
    assert rc == 0, rc
    return out, list(stats)


# ---- synthetic ViBE-Z-shaped volumes (numpy twin of resident_kernels.cu: k_synth) ---------------
def _hash_u32(x):
    x = x.astype(np.uint32)
    x ^= x >> np.uint32(16)
    x *= np.uint32(0x7FEB352D)
    x ^= x >> np.uint32(15)
    x *= np.uint32(0x846CA68B)
    x ^= x >> np.uint32(16)
    return x


def _hash3(i, j, k, seed):
    with np.errstate(over="ignore"):
        a = _hash_u32(k.astype(np.uint32) * np.uint32(0xC2B2AE3D) ^ np.uint32(seed & 0xFFFFFFFF))
        b = _hash_u32(j.astype(np.uint32) * np.uint32(0x85EBCA77) ^ a)
        return _hash_u32(i.astype(np.uint32) * np.uint32(0x9E3779B1) ^ b)


def _u01(h):
    return (h >> np.uint32(8)).astype(np.float32) * np.float32(1.0 / 16777216.0)


def _smooth(f):
    return (f * f) * (np.float32(3.0) - np.float32(2.0) * f)


def _lerp(a, b, f):
    return a + f * (b - a)


def _value_noise(px, py, pz, seed):
    fx, fy, fz = np.floor(px), np.floor(py), np.floor(pz)
    i, j, k = fx.astype(np.int32).astype(np.uint32), fy.astype(np.int32).astype(np.uint32), fz.astype(np.int32).astype(np.uint32)
    wx, wy, wz = _smooth(px - fx), _smooth(py - fy), _smooth(pz - fz)
    one = np.uint32(1)
    c = {}
    with np.errstate(over="ignore"):
        for dz in (0, 1):
            for dy in (0, 1):
                for dx in (0, 1):
                    c[(dx, dy, dz)] = _u01(_hash3(i + one * np.uint32(dx), j + one * np.uint32(dy), k + one * np.uint32(dz), seed))
    c00 = _lerp(c[(0, 0, 0)], c[(1, 0, 0)], wx)
    c10 = _lerp(c[(0, 1, 0)], c[(1, 1, 0)], wx)
    c01 = _lerp(c[(0, 0, 1)], c[(1, 0, 1)], wx)
    c11 = _lerp(c[(0, 1, 1)], c[(1, 1, 1)], wx)
    return _lerp(_lerp(c00, c10, wy), _lerp(c01, c11, wy), wz)


def _ncoord(n):
    i = np.arange(n, dtype=np.float32)
    return ((i + np.float32(0.5)) / np.float32(n)) * np.float32(2.0) - np.float32(1.0)


def synth_brain(dims, seed=SEED):
    """kind 0 of vr_synth_volume: array of shape dims (dims[0] fastest = MATLAB dim 1), float32."""
    w, h, d = dims
    f = np.float32
    u = _ncoord(w)[:, None, None]
    v = _ncoord(h)[None, :, None]
    s = _ncoord(d)[None, None, :]
    a, b, c = u / f(0.8), v / f(0.55), s / f(0.4)
    r2 = ((a * a + b * b) + c * c).astype(np.float32)
    inside = r2 < f(1.0)
    m = _smooth(np.minimum((f(1.0) - r2) * f(4.0), f(1.0)))
    qx = np.broadcast_to((u + f(1.0)) * f(0.5), r2.shape).astype(np.float32)
    qy = np.broadcast_to((v + f(1.0)) * f(0.5), r2.shape).astype(np.float32)
    qz = np.broadcast_to((s + f(1.0)) * f(0.5), r2.shape).astype(np.float32)
    n0 = _value_noise(qx * f(8), qy * f(8), qz * f(8), seed)
    n1 = _value_noise(qx * f(16), qy * f(16), qz * f(16), seed + 1)
    n2 = _value_noise(qx * f(32), qy * f(32), qz * f(32), seed + 2)
    nz = (f(0.5) * n0 + f(0.3) * n1) + f(0.2) * n2
    val = m * (f(0.6) + f(0.4) * nz)
    out = np.where(inside, np.minimum(np.maximum(val, f(0.0)), f(1.0)), f(0.0)).astype(np.float32)
    return np.asfortranarray(out)


def synth_tubes(dims, seed=SEED, ntubes=200):
    """kind 1 of vr_synth_volume."""
    w, h, d = dims
    f = np.float32
    px = np.broadcast_to(_ncoord(w)[:, None, None], (w, h, d)).astype(np.float32)
    py = np.broadcast_to(_ncoord(h)[None, :, None], (w, h, d)).astype(np.float32)
    pz = np.broadcast_to(_ncoord(d)[None, None, :], (w, h, d)).astype(np.float32)
    vox = f(2.0) / f(w)
    best = np.zeros((w, h, d), dtype=np.float32)
    for t in range(ntubes):
        with np.errstate(over="ignore"):
            s = np.uint32(seed & 0xFFFFFFFF) ^ (np.uint32(t) * np.uint32(0x632BE5AB))
            hs = [_u01(_hash_u32(np.array([s + np.uint32(k)], dtype=np.uint32)))[0] for k in range(1, 8)]
        ax, ay, az = hs[0] * f(1.6) - f(0.8), hs[1] * f(1.6) - f(0.8), hs[2] * f(1.6) - f(0.8)
        dx, dy, dz = hs[3] * f(0.8) - f(0.4), hs[4] * f(0.8) - f(0.4), hs[5] * f(0.8) - f(0.4)
        rad = (f(3.0) + hs[6] * f(3.0)) * vox
        ex, ey, ez = px - ax, py - ay, pz - az
        dd = (dx * dx + dy * dy) + dz * dz
        ed = (ex * dx + ey * dy) + ez * dz
        tt = np.minimum(np.maximum(ed / np.maximum(dd, f(1e-12)), f(0.0)), f(1.0))
        cx, cy, cz = ex - tt * dx, ey - tt * dy, ez - tt * dz
        dist = np.sqrt((cx * cx + cy * cy) + cz * cz)
        edge = f(1.5) * vox
        vv = np.minimum(np.maximum((rad - dist) / edge, f(0.0)), f(1.0))
        best = np.maximum(best, vv.astype(np.float32))
    return np.asfortranarray(best)


# ---- scenes shaped like the reference's examples ------------------------------------------------
def example1_scene(cls, emission, resolution=(128, 128), lights=1, lut=None, n_lut=64):
    """examples/example1.m:28-60 with a synthetic volume.  `cls` is api.VolumeRender or a subclass
    bound to another backend."""
    r = cls()
    if lut is None:
        lut = oracle_lut(n_lut)
    r.VolumeIllumination = api.Volume(lut)
    all_lights = [api.LightSource([500, 1000, 550], [0, 1, 1]), api.LightSource([550, 1000, 500], [1, 0, 1])]
    r.LightSources = all_lights[:lights]
    r.ElementSizeUm = [1, 1, 1]  # synthetic isotropic volume
    r.FocalLength = 3.0
    r.DistanceToObject = 6.0
    r.rotate(125, 25, 0)
    r.OpacityThreshold = 0.9
    r.ImageResolution = list(resolution)
    vol = emission if isinstance(emission, api.Volume) else api.Volume(emission)
    r.VolumeEmission = vol
    r.VolumeAbsorption = vol
    r.FactorEmission = 1
    r.FactorAbsorption = 0.6
    r.FactorReflection = 0.4
    r.Color = [1, 1, 0]
    return r


def oracle_lut(n=64, g=0.8):
    out = np.zeros(n * n * n, dtype=np.float32)
    oracle().orc_henyey_greenstein(n, g, out.ctypes.data)
    return out.reshape((n, n, n), order="F")


# ---- image metrics --------------------------------------------------------------------------------
def compare(img, ref):
    """max-abs error and PSNR on RGB normalised to [0,1] by the oracle image's own max."""
    img = np.asarray(img, dtype=np.float64)
    ref = np.asarray(ref, dtype=np.float64)
    assert img.shape == ref.shape, (img.shape, ref.shape)
    peak = float(np.max(ref))
    if not peak > 0:
        peak = 1.0
    diff = np.abs(img - ref)
    raw = float(np.nanmax(diff)) if diff.size else 0.0
    if np.isnan(img).any() != np.isnan(ref).any():
        raw = float("inf")
    mse = float(np.mean((diff / peak) ** 2))
    psnr = float("inf") if mse == 0 else 10.0 * math.log10(1.0 / mse)
    return {"max_abs_norm": raw / peak, "max_abs_raw": raw, "psnr_db": psnr, "peak": peak}


TOL = 1e-3  # BASELINE.json north_star: max abs error <= 1e-3 on [0,1]-normalised RGB


def exact(r):
    """Select the "exact" arithmetic flavour (VR_OPT_ARITHMETIC = 0: no FMA contraction, IEEE sqrt / divide) on a product
    renderer: the flavour that agrees with the CPU oracle bit for bit up to acosf, and is therefore held to TOL on every
    scene.  (The default flavour reproduces what nvcc makes of the reference's own source and is pinned by the reference
    kernel's golden images instead, tests/test_golden_refcuda.py.)  No-op on oracle-backed renderers."""
    if type(r._backend).__name__ == "_CudaBackend":
        r.setOption(_abi.OPT_ARITHMETIC, _abi.ARITH_EXACT)
    return r


def check_parity(img, ref, exact_flavour=True):
    """The bar of a product image against the oracle's.  Exact flavour: max-abs <= TOL of the oracle image's maximum.
    Other flavours differ from the oracle exactly as the reference's own CUDA kernel does (FMA contraction, rsqrtf; isolated
    samples on the illumination's NaN discontinuity flip sides): they are held to the envelope the ORACLE is held to against
    the reference kernel's golden images (tests/test_golden_refcuda.py: check_envelope) - PSNR >= 70 dB, >= 99.5 % of the
    pixels within 1e-3, median <= 1e-4, isolated pixels <= 3e-2 - on the small test scenes, whose few voxels make single
    samples weigh percent of a pixel, and to TOL itself at full size (bench.py `parity`)."""
    m = compare(img, ref)
    if exact_flavour:
        assert m["max_abs_norm"] <= TOL, m
        return m
    d = np.abs(np.asarray(img, np.float64) - np.asarray(ref, np.float64)).max(axis=2) / m["peak"]
    m["frac_within_1e-3"] = float((d <= 1e-3).mean())
    m["median"] = float(np.median(d))
    assert m["psnr_db"] >= 70.0 and m["frac_within_1e-3"] >= 0.995 and m["median"] <= 1e-4 and m["max_abs_norm"] <= 3e-2, m
    return m


# ---- the shimmed reference CUDA kernel (oracle/_ref/libref_cuda.so) -----------------------------------
def ref_cuda():
    """The reference's own d_render + upload code on texture objects (oracle/ref_cuda), or None."""
    path = os.path.join(ORACLE_DIR, "_ref", "libref_cuda.so")
    if not os.path.exists(path):
        return None
    lib = C.CDLL(path)
    P = C.POINTER
    f3, u64p = P(C.c_float), P(C.c_uint64)
    lib.refcuda_sync_volumes.argtypes = [C.c_uint64] + [P(_abi.VrVolume)] * 6
    lib.refcuda_render.argtypes = [P(_abi.VrLight), C.c_size_t, P(_abi.VrVolume), f3, f3, u64p, f3, f3, C.c_float, f3, C.c_void_p]
    return lib


class RefCudaBackend:
    """The mex command protocol on top of the shimmed reference.  Its state is process-global (the
    reference's texture references and d_idx* are), so use ONE renderer per process."""

    def __init__(self):
        self.lib = ref_cuda()
        if self.lib is None:
            raise RuntimeError("oracle/_ref/libref_cuda.so is not built (make -C oracle refcuda)")

    def new(self):
        return C.c_void_p(1)

    def delete(self, h):
        return 0

    def mem_info(self, h):
        return "shimmed reference\n"

    def sync_volumes(self, h, T, em, re, ab, gx=None, gy=None, gz=None):
        keep = []
        vols = [_abi.volume_struct(v.Data, v.TimeLastUpdate, keep) if v is not None else None for v in (em, re, ab, gx, gy, gz)]
        ptrs = [C.byref(v) if v is not None else None for v in vols]
        rc = self.lib.refcuda_sync_volumes(int(T), *ptrs)
        assert rc == 0, rc
        self._keep = keep

    def render(self, h, lights, lut, factors, es, res, rot, props, thr, color):
        H, W = int(res[0]), int(res[1])
        out = np.zeros((H, W, 3), dtype=np.float32, order="F")
        args, keep = api._render_args(lights, lut, factors, es, res, rot, props, thr, color)
        rc = self.lib.refcuda_render(*args, out.ctypes.data)
        assert rc == 0, rc
        return out


class RefCudaVolumeRender(api.VolumeRender):
    _backend_cls = RefCudaBackend


def golden_scenes():
    """name -> function(cls) building a configured renderer: the scenes whose images, rendered on a
    B200 by the shimmed reference kernel, are committed under tests/golden/refcuda_*.npy."""
    res = (96, 80)

    def base(cls, vol, lights, **kw):
        return example1_scene(cls, vol, res, lights=lights, lut=oracle_lut(64))

    def s_example1(cls):
        return base(cls, synth_brain((48, 48, 48)), 1)

    def s_two_lights_dense(cls):
        r = base(cls, synth_brain((48, 48, 48)), 2)
        r.FactorAbsorption = 30.0  # rays terminate at the 0.9 threshold
        return r

    def s_gradient_volumes(cls):
        r = base(cls, synth_brain((48, 48, 48)), 2)
        gx, gy, gz = r.VolumeEmission.grad()
        r.VolumeGradientX, r.VolumeGradientY, r.VolumeGradientZ = gx, gy, gz
        r.FactorEmission, r.FactorAbsorption, r.FactorReflection = 0.1, 0.4, 0.1
        r.Color = [1, 1, 1]
        return r

    def s_anisotropic(cls):
        r = base(cls, synth_brain((56, 40, 28)), 1)
        r.ElementSizeUm = [1.0, 0.8, 2.5]
        return r

    def s_distinct_roles(cls):
        v = synth_brain((40, 40, 40))
        r = base(cls, v, 1)
        r.VolumeAbsorption = api.Volume(np.asfortranarray(np.sqrt(v)))
        r.VolumeReflection = api.Volume(np.asfortranarray(v * v))
        return r

    def s_constant_volume_nan_normal(cls):
        return base(cls, np.full((16, 16, 16), 0.5, dtype=np.float32, order="F"), 1)

    def s_stereo_pass(cls):
        r = base(cls, synth_brain((40, 40, 40)), 1)
        r.CameraXOffset = 0.06
        r.FocalLength = 4.5
        r.StereoOutput = api.StereoRenderMode.LeftRightHorizontal
        return r

    return {"example1": s_example1, "two_lights_dense": s_two_lights_dense, "gradient_volumes": s_gradient_volumes,
            "anisotropic": s_anisotropic, "distinct_roles": s_distinct_roles, "constant_volume_nan_normal": s_constant_volume_nan_normal,
            "stereo_pass": s_stereo_pass}
