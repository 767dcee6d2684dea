"""One rank of a volume-space (sort-last) render, alone on one GPU: the region context of brick `--rank` of `--world` bricks of a
`--size`^3 volume generated straight into HBM, its opacity pass and its shading pass (the opacities of the other bricks taken
as 0, i.e. the work of a front brick).  For ncu: `ncu -k regex:k_raymarch -s 2 -c 2 python tools/vs_one_rank.py ...` captures one
opacity-pass and one shading-pass launch of exactly the kernels a rank of `bench.py --mode volume` runs."""
import argparse
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import bench  # noqa: E402
from volume_renderer_b200 import _abi, api, multigpu  # noqa: E402

p = argparse.ArgumentParser()
p.add_argument("--size", type=int, default=4096)
p.add_argument("--world", type=int, default=8)
p.add_argument("--rank", type=int, default=1)
p.add_argument("--dtype", default="bf16")
p.add_argument("--image", type=int, default=2048)
p.add_argument("--frames", type=int, default=3)
p.add_argument("--texture-unit", type=int, default=2)
p.add_argument("--slices", type=int, default=1)
p.add_argument("--part", type=int, nargs=2, default=None, metavar=("RANK", "NRANKS"), help="render only the tiles RANK + j * NRANKS (debugging)")
p.add_argument("--tile", type=int, default=64)
p.add_argument("--first", type=int, default=0, help="index of the first frame (the camera turns 1 degree per frame)")
a = p.parse_args()
lib = _abi.lib()
n = a.size
dims = (n, n, n)
bricks = multigpu.brick_grid(dims, multigpu.splits_for(a.world))
lo, hi = bricks[a.rank]
halo = 4
dt = _abi.DTYPE_F32 if a.dtype == "f32" else _abi.DTYPE_BF16
es = 4 if dt == _abi.DTYPE_F32 else 2
r = api.VolumeRender()
ctx = r.objectHandle
r.setRegion(lo, hi, halo)
r.setOption(_abi.OPT_TEXTURE_UNIT, a.texture_unit)
r.setOption(_abi.OPT_VS_SLICES, a.slices)
if a.part:
    multigpu.configure_partition(ctx, a.part[0], a.part[1], a.tile, a.tile)
blo = [max(l - halo, 0) for l in lo]
bhi = [min(h + halo, d) for h, d in zip(hi, dims)]
nvox = int(np.prod([h - l for l, h in zip(blo, bhi)]))
buf = C.c_void_p()
_abi.check(lib.vr_device_alloc(ctx, nvox * es, C.byref(buf)), ctx)
u3 = lambda v: (C.c_uint64 * 3)(*[int(x) for x in v])
_abi.check(lib.vr_synth_volume_region(ctx, 0, bench.SEED, u3(dims), u3(blo), u3(bhi), dt, buf), ctx)
E = api.DeviceVolume(buf.value, dims, dt, _abi.MEM_DEVICE_REGION)
r.VolumeIllumination = api.Volume(api.HenyeyGreenstein(64))
r.LightSources = [api.LightSource([500, 1000, 550], [0, 1, 1])]
r.ElementSizeUm = [1, 1, 1]
r.FocalLength, r.DistanceToObject, r.OpacityThreshold = 3.0, 6.0, 0.9
r.ImageResolution = [a.image, a.image]
r.VolumeEmission = E
r.VolumeAbsorption = E
r.FactorEmission, r.FactorAbsorption, r.FactorReflection = 1.0, 0.6, 0.4
r.Color = [1, 1, 0]
R0 = bench.rotation(125, 25, 0)
r.syncVolumes()
_abi.check(lib.vr_device_free(ctx, buf), ctx)
r.VolumeEmission.TimeLastUpdate = 1
r.VolumeAbsorption.TimeLastUpdate = 1
npx = a.image * a.image
opac_all = torch.zeros(a.world * npx, dtype=torch.float32, device="cuda")
mine = torch.zeros(npx, dtype=torch.float32, device="cuda")
out = torch.zeros(3 * npx, dtype=torch.float32, device="cuda")
def dump_trace():
    full = np.zeros(8 * 8192, dtype=np.uint64)
    _abi.check(lib.vr_debug_warp_times(ctx, full.ctypes.data, 4 * 8192), ctx)
    f = lambda u: float(np.array([u & 0xffffffff], dtype=np.uint32).view(np.float32)[0])
    for slot in range(1, 65):
        e = [int(v) for v in full[8 * slot : 8 * slot + 6]]
        print("px +%d,+%d stage %d n %d t %.7g tfar %.7g pos %.7g %.7g %.7g marched %d a %d b %d" % (
            (slot - 1) % 8, (slot - 1) // 8, e[0] & 0xff, e[0] >> 8, f(e[1]), f(e[1] >> 32), f(e[2]), f(e[2] >> 32), f(e[3]), e[3] >> 32, e[4], e[5]))


if os.environ.get("VR_TRACE"):
    import atexit
    atexit.register(dump_trace)
for k in range(a.first, a.first + a.frames):
    r.RotationMatrix = bench.rotation(0, float(k), 0, R0)
    packed = r._packed_mono_args()
    r.renderOpacity(mine.data_ptr(), None, packed)
    torch.cuda.synchronize()
    op_ms = r.stats()["kernel_ms"]
    r.renderPartial(bricks, a.rank, opac_all.data_ptr(), out.data_ptr(), None, packed)
    if os.environ.get("VR_DEBUG_BOUNDS"):
        torch.cuda.synchronize()
        full = np.zeros(8 * 8192, dtype=np.uint64)
        _abi.check(lib.vr_debug_warp_times(ctx, full.ctypes.data, 4 * 8192), ctx)
        nlog = int(full[2 * 8192])
        print("out-of-range cells logged:", nlog)
        for i in range(min(nlog, 12)):
            e = full[2 * 8192 + 8 + 6 * i : 2 * 8192 + 8 + 6 * i + 6]
            f = lambda u: float(np.array([u & 0xffffffff], dtype=np.uint32).view(np.float32)[0])
            print("   cell %d t %.6f tfar %.6f pos (%.6f %.6f %.6f) cx,cy,cz %d %d %d marched %d" % (
                int(e[0]), f(int(e[1])), f(int(e[1]) >> 32), f(int(e[2])), f(int(e[2]) >> 32), f(int(e[3])), int(e[4]) & 0xfffff, (int(e[4]) >> 20) & 0xfffff, int(e[4]) >> 40, int(e[5])))
    st = r.stats()
    print("frame %d rank %d/%d slices %d -> %d: opacity pass %.2f ms, shading pass %.2f ms, %d shaded samples (%.1f G/s), %d lattice samples visited, fetch_mode %d, resident %.1f GB" % (
        k, a.rank, a.world, a.slices, st["vs_slices"], op_ms, st["kernel_ms"], st["samples_shaded"], st["samples_shaded"] / st["kernel_ms"] / 1e6, st["samples_marched"], st["fetch_mode"], st["bytes_resident"] / 1e9), flush=True)
    if os.environ.get("VR_WARP_TIMES"):
        NW = 148 * 32
        full = np.zeros(8 * 8192, dtype=np.uint64)
        _abi.check(lib.vr_debug_warp_times(ctx, full.ctypes.data, 4 * 8192), ctx)
        wt = full[: 2 * NW]
        rec = full[2 * 8192 : 2 * 8192 + 6 * NW].reshape(NW, 6)
        if rec[:, 0].any():  # a VR_ITEM_PROFILE build: the longest item of every warp
            for i in np.argsort(rec[:, 0])[::-1][:6]:
                d, xy, sh, mm, loops, jumps = [int(v) for v in rec[i]]
                print("   item of %.2f ms at pixel (%d, %d) slice %d: longest ray %d shaded / %d visited, %d shaded in the warp, %d loop trips, %d jump calls" % (
                    d / 1e6, xy & 0xffff, (xy >> 16) & 0xffff, xy >> 32, sh & 0xffffffff, mm, sh >> 32, loops, jumps))
            print("   items: median of the per-warp longest %.3f ms" % (np.median(rec[:, 0]) / 1e6))
        st0, en = wt[0::2].astype(np.int64), wt[1::2].astype(np.int64)
        end = (en - st0.min()) / 1e6
        print("   shading-pass warps end at ms: p10 %.2f p50 %.2f p90 %.2f p99 %.2f max %.2f mean %.2f; still running at T-4,-2,-1,-0.5 ms: %s" % (
            *np.percentile(end, [10, 50, 90, 99]), end.max(), end.mean(), [int((end > end.max() - d).sum()) for d in (4, 2, 1, 0.5)]), flush=True)
