"""Single-frame latency of the headline scene (bench.py's configs[1]) on G GPUs behind one context (vr_set_devices), plain march
against VR_OPT_RAY_SLICES = S.  `python tools/latency_mode.py --gpus 8 [--size 1024]`.  Times: vr_stats.kernel_ms (the slowest
member's launches, CUDA events) and the host's wall clock around vr_render_device + vr_get_stats, median of the frames."""
import ctypes as C
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import bench  # noqa: E402
from volume_renderer_b200 import _abi  # noqa: E402

argv = sys.argv[1:]
gpus = int(argv[argv.index("--gpus") + 1]) if "--gpus" in argv else 1
slices = [int(x) for x in argv[argv.index("--slices") + 1].split(",")] if "--slices" in argv else [1, 4, 8, 16]
for flag in ("--gpus", "--slices"):
    if flag in argv:
        i = argv.index(flag)
        del argv[i : i + 2]
sys.argv = [sys.argv[0]] + argv
args = bench.parse()
lib = _abi.lib()
ctx = C.c_void_p()
_abi.check(lib.vr_create(C.byref(ctx)))
if gpus > 1:
    ids = (C.c_int * gpus)(*range(gpus))
    _abi.check(lib.vr_set_option(ctx, _abi.OPT_TILE_W, 32), ctx)
    _abi.check(lib.vr_set_devices(ctx, gpus, ids), ctx)
scene = bench.Scene(args)
n = args.size
dims = (C.c_uint64 * 3)(n, n, n)
E = torch.empty(n ** 3, dtype=torch.float32, device="cuda")
_abi.check(lib.vr_synth_volume(ctx, 0, bench.SEED, dims, 0, C.c_void_p(E.data_ptr())), ctx)


def vol(t, st):
    v = _abi.VrVolume()
    v.data = t.data_ptr()
    v.dims[0] = v.dims[1] = v.dims[2] = n
    v.last_update, v.dtype, v.memspace = st, 0, 1
    return v


vE = vol(E, 1)
R = E * E
vR = vol(R, 2)
G = [torch.empty(n ** 3, dtype=torch.float32, device="cuda") for _ in range(3)]
_abi.check(lib.vr_volume_gradient(ctx, C.byref(vE), 1, *[C.c_void_p(t.data_ptr()) for t in G]), ctx)
g = [vol(G[i], 4 + i) for i in range(3)]
_abi.check(lib.vr_sync_volumes(ctx, 0, C.byref(vE), C.byref(vR), C.byref(vE), *[C.byref(x) for x in g]), ctx)
lut = np.zeros(64 ** 3, dtype=np.float32)
_abi.check(lib.vr_henyey_greenstein(64, 0.8, lut.ctypes.data))
lutv = _abi.VrVolume()
lutv.data = lut.ctypes.data
lutv.dims[0] = lutv.dims[1] = lutv.dims[2] = 64
lutv.last_update = 9
lights = (_abi.VrLight * 2)()
for i, (p, c) in enumerate(scene.lights):
    lights[i].position[:] = p
    lights[i].color[:] = c
img = torch.zeros(scene.W * scene.H * 3, device="cuda")
ref = None
resp = scene.res.ctypes.data_as(C.POINTER(C.c_uint64))
f = bench.f32p
for S in slices:
    _abi.check(lib.vr_set_option(ctx, _abi.OPT_RAY_SLICES, S), ctx)
    kms, wall = [], []
    for it in range(8):
        rot = scene.rot(it % 4)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        _abi.check(lib.vr_render_device(ctx, lights, 2, C.byref(lutv), f(scene.factors), f(scene.es), resp, f(rot), f(scene.props), C.c_float(scene.thr), f(scene.color),
                                        C.c_void_p(img.data_ptr()), None), ctx)
        s = _abi.VrStats()
        _abi.check(lib.vr_get_stats(ctx, C.byref(s)), ctx)
        wall.append((time.perf_counter() - t0) * 1e3)
        kms.append(s.kernel_ms)
        if it == 3:
            frame = img.cpu().numpy()
    if ref is None:
        ref = frame
    err = float(np.abs(frame - ref).max() / ref.max())
    print("gpus %d ray slices %2d: kernel %.2f ms (median of frames 3..7; slowest member), wall %.2f ms, launches %d, slices used %d, max |diff| to the plain frame %.2e of the maximum" % (
        gpus, S, float(np.median(kms[3:])), float(np.median(wall[3:])), s.kernel_launches, s.vs_slices, err), flush=True)
