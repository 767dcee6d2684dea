"""Volume-space (sort-last) mode on the GPU with virtual shards: every brick is its own context on
cuda:0 holding only its voxel box plus halo; opacity pass, shading pass, composite.  Compared with
the same library's single-pass render (bit-identical storage contents, different association of the
opacity product) and with the oracle."""
import ctypes as C

import numpy as np
import pytest

import vrt
from volume_renderer_b200 import _abi, api, multigpu

pytestmark = pytest.mark.gpu
RES = (160, 128)


def _scene(vol, lut, grad, bricks=None, b=None, halo=0):
    r = api.VolumeRender()
    if bricks is not None:
        r.setRegion(bricks[b][0], bricks[b][1], halo)  # before the volumes are assigned / synced
    r = vrt.example1_scene(lambda: r, vol, RES, lights=2, lut=lut)
    r.setOption(_abi.OPT_TEXTURE_UNIT, 0)  # region contexts fetch with loads: compare like with like
    vrt.exact(r)  # held to TOL against the oracle
    r.FactorAbsorption = 30.0  # rays reach the 0.9 threshold inside the volume
    if grad:
        gx, gy, gz = api.Volume(vol).grad()
        r.VolumeGradientX, r.VolumeGradientY, r.VolumeGradientZ = gx, gy, gz
    return r


@pytest.mark.parametrize("splits", [(1, 1, 2), (2, 2, 2), (3, 1, 2)])
@pytest.mark.parametrize("grad", [0, 1])
def test_bricks_on_one_gpu_match_single_pass(splits, grad):
    vol = vrt.synth_brain((72, 64, 56))
    lut = api.HenyeyGreenstein(64)
    full = _scene(vol, lut, grad).render()
    bricks = multigpu.brick_grid(vol.shape, splits)
    rs = [_scene(vol, lut, grad, bricks, b) for b in range(len(bricks))]
    img = multigpu.render_bricks_one_gpu(rs, bricks)
    m = vrt.compare(img, full)
    print(splits, grad, m)
    assert m["max_abs_norm"] < 2e-5, m
    # only the owned box (+ halo) is resident
    whole = _scene(vol, lut, grad)
    whole.render()
    assert rs[0].stats()["bytes_resident"] < 0.75 * whole.stats()["bytes_resident"] or len(bricks) < 3
    orc = vrt.example1_scene(vrt.OracleVolumeRender, vol, RES, lights=2, lut=lut)
    orc.FactorAbsorption = 30.0
    if grad:
        gx, gy, gz = api.Volume(vol).grad()
        orc.VolumeGradientX, orc.VolumeGradientY, orc.VolumeGradientZ = gx, gy, gz
    assert vrt.compare(img, orc.render())["max_abs_norm"] <= vrt.TOL


@pytest.mark.parametrize("splits,slices", [((1, 1, 1), 2), ((1, 1, 2), 3), ((2, 2, 2), 4), ((1, 2, 1), 16)])
@pytest.mark.parametrize("grad", [0, 1])
def test_depth_slices_shade_every_sample_once(splits, slices, grad):
    """VR_OPT_VS_SLICES: the stretch of a ray inside a brick is marched by `slices` different warps.  Every lattice sample is
    still shaded exactly once and at its bit-exact position - the counters say so - and the image agrees with whole rays to
    float rounding (a slice's colour share is summed afterwards instead of in the loop)."""
    vol = vrt.synth_brain((72, 64, 56))
    lut = api.HenyeyGreenstein(64)
    bricks = multigpu.brick_grid(vol.shape, splits)
    imgs, shaded, marched = {}, {}, {}
    for S in (1, slices):
        rs = [_scene(vol, lut, grad, bricks, b) for b in range(len(bricks))]
        for r in rs:
            r.FactorAbsorption = 4.0  # translucent: no ray ends at the threshold, both passes shade the same samples
            r.setOption(_abi.OPT_VS_SLICES, S)
        imgs[S] = multigpu.render_bricks_one_gpu(rs, bricks)
        shaded[S] = sum(r.stats()["samples_shaded"] for r in rs)  # (the shading pass: the last launch of each context)
        marched[S] = sum(r.stats()["samples_marched"] for r in rs)
    m = vrt.compare(imgs[slices], imgs[1])
    print(splits, slices, grad, m, shaded)
    assert m["max_abs_norm"] < 2e-6, m
    assert shaded[slices] == shaded[1] and shaded[1] > 0
    # ... and visited by one worker only: the fast-forward to a slice's start is not counted, a jump stops at its end
    # (the two real steps a jump starts with may cross the slice's end: at most two samples per ray and border are counted twice)
    assert 0 <= marched[slices] - marched[1] <= 2 * (slices - 1) * RES[0] * RES[1] * len(bricks), marched
    full = _scene(vol, lut, grad)
    full.FactorAbsorption = 4.0
    assert vrt.compare(imgs[slices], full.render())["max_abs_norm"] < 2e-5
    # opaque scene: rays end at the 0.9 threshold inside the volume - a slice behind the break must add nothing
    for S in (1, slices):
        rs = [_scene(vol, lut, grad, bricks, b) for b in range(len(bricks))]
        for r in rs:
            r.setOption(_abi.OPT_VS_SLICES, S)
        imgs[S] = multigpu.render_bricks_one_gpu(rs, bricks)
    m = vrt.compare(imgs[slices], imgs[1])
    print("opaque", m)
    assert m["max_abs_norm"] < 2e-5, m
    assert vrt.compare(imgs[slices], _scene(vol, lut, grad).render())["max_abs_norm"] < 2e-5


def test_big_brick_whole_rays_and_depth_slices_agree():
    """One brick of a 1024^3 volume at the benchmark's camera and image size (bench.py --mode volume, rank 0 of 2), whole rays
    against five depth slices: the same opacity plane and colour share.  Also the frame on which a build of
    this library once faulted: keep it at this size and camera."""
    import os
    import sys

    import torch

    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import bench

    n, image, halo = 1024, 2048, 4
    dims = (n, n, n)
    bricks = multigpu.brick_grid(dims, multigpu.splits_for(2))
    lo, hi = bricks[0]
    lib = _abi.lib()
    npx = image * image
    got = {}
    for slices in (1, 5):
        r = api.VolumeRender()
        ctx = r.objectHandle
        r.setRegion(lo, hi, halo)
        r.setOption(_abi.OPT_VS_SLICES, slices)
        blo = [max(l - halo, 0) for l in lo]
        bhi = [min(h + halo, d) for h, d in zip(hi, dims)]
        nvox = int(np.prod([h - l for l, h in zip(blo, bhi)]))
        buf = C.c_void_p()
        _abi.check(lib.vr_device_alloc(ctx, nvox * 4, C.byref(buf)), ctx)
        u3 = lambda v: (C.c_uint64 * 3)(*[int(x) for x in v])
        _abi.check(lib.vr_synth_volume_region(ctx, 0, bench.SEED, u3(dims), u3(blo), u3(bhi), _abi.DTYPE_F32, buf), ctx)
        E = api.DeviceVolume(buf.value, dims, _abi.DTYPE_F32, _abi.MEM_DEVICE_REGION)
        r.VolumeIllumination = api.Volume(api.HenyeyGreenstein(64))
        r.LightSources = [api.LightSource([500, 1000, 550], [0, 1, 1])]
        r.ElementSizeUm = [1, 1, 1]
        r.FocalLength, r.DistanceToObject, r.OpacityThreshold = 3.0, 6.0, 0.9
        r.ImageResolution = [image, image]
        r.VolumeEmission = E
        r.VolumeAbsorption = E
        r.FactorEmission, r.FactorAbsorption, r.FactorReflection = 1.0, 0.6, 0.4
        r.Color = [1, 1, 0]
        r.syncVolumes()
        _abi.check(lib.vr_device_free(ctx, buf), ctx)
        r.VolumeEmission.TimeLastUpdate = 1
        r.VolumeAbsorption.TimeLastUpdate = 1
        opac_all = torch.zeros(2 * npx, dtype=torch.float32, device="cuda")
        out = torch.zeros(3 * npx, dtype=torch.float32, device="cuda")
        R0 = bench.rotation(125, 25, 0)
        for k in (0, 1):
            r.RotationMatrix = bench.rotation(0, float(k), 0, R0)
            packed = r._packed_mono_args()
            r.renderOpacity(opac_all.data_ptr(), None, packed)  # plane 0 = this brick's; the other brick's stays 0
            r.renderPartial(bricks, 0, opac_all.data_ptr(), out.data_ptr(), None, packed)
            st = r.stats()
        assert st["fetch_mode"] == 1 and st["vs_slices"] == slices, st
        got[slices] = (opac_all[:npx].cpu().numpy(), out.cpu().numpy(), st["samples_shaded"])
        del r
    peak = got[1][1].max()
    assert peak > 0 and got[1][2] == got[5][2]
    assert np.abs(got[1][0] - got[5][0]).max() < 2e-6
    assert np.abs(got[1][1] - got[5][1]).max() / peak < 2e-5


def test_depth_slices_need_the_same_cut_in_both_passes():
    vol = vrt.synth_brain((40, 40, 40))
    lut = api.HenyeyGreenstein(64)
    bricks = multigpu.brick_grid(vol.shape, (1, 1, 1))
    r = _scene(vol, lut, 0, bricks, 0)
    r.setOption(_abi.OPT_VS_SLICES, 4)
    import torch

    op = torch.zeros(RES[0] * RES[1], dtype=torch.float32, device="cuda")
    out = torch.zeros(RES[0] * RES[1] * 3, dtype=torch.float32, device="cuda")
    r.renderOpacity(op.data_ptr())
    r.setOption(_abi.OPT_VS_SLICES, 2)
    with pytest.raises(_abi.VrError, match="depth slices"):
        r.renderPartial(bricks, 0, op.data_ptr(), out.data_ptr())
    with pytest.raises(_abi.VrError):
        r.setOption(_abi.OPT_VS_SLICES, 17)


def test_region_buffer_generated_on_the_device():
    """config 5 shape: each brick's voxels (+ halo) are generated straight into HBM and handed over
    as a VR_MEM_DEVICE_REGION source - the whole volume never exists anywhere."""
    dims = (64, 48, 80)
    vol = vrt.synth_brain(dims)
    lut = api.HenyeyGreenstein(64)
    full = _scene(vol, lut, 0).render()
    bricks = multigpu.brick_grid(dims, (2, 1, 2))
    lib = _abi.lib()
    rs, bufs = [], []
    halo = 3
    for b, (lo, hi) in enumerate(bricks):
        r = api.VolumeRender()
        r.setRegion(lo, hi, halo)
        blo = [max(l - halo, 0) for l in lo]
        bhi = [min(h + halo, d) for h, d in zip(hi, dims)]
        n = int(np.prod([h - l for l, h in zip(blo, bhi)]))
        ptr = C.c_void_p()
        _abi.check(lib.vr_device_alloc(r.objectHandle, n * 4, C.byref(ptr)), r.objectHandle)
        u3 = lambda v: (C.c_uint64 * 3)(*[int(x) for x in v])
        _abi.check(lib.vr_synth_volume_region(r.objectHandle, 0, vrt.SEED, u3(dims), u3(blo), u3(bhi), _abi.DTYPE_F32, ptr), r.objectHandle)
        dv = api.DeviceVolume(ptr.value, dims, _abi.DTYPE_F32, _abi.MEM_DEVICE_REGION)
        r = vrt.exact(vrt.example1_scene(lambda r=r: r, dv, RES, lights=2, lut=lut))  # like _scene: compared with its single-pass render
        r.setOption(_abi.OPT_TEXTURE_UNIT, 0)
        r.FactorAbsorption = 30.0
        rs.append(r)
        bufs.append((r, ptr))
    img = multigpu.render_bricks_one_gpu(rs, bricks)
    assert vrt.compare(img, full)["max_abs_norm"] < 2e-5
    for r, ptr in bufs:
        _abi.check(lib.vr_device_free(r.objectHandle, ptr), r.objectHandle)


def test_region_contexts_refuse_plain_render_and_small_halos():
    vol = vrt.synth_brain((32, 32, 32))
    lut = api.HenyeyGreenstein(64)
    bricks = multigpu.brick_grid(vol.shape, (1, 1, 2))
    r = _scene(vol, lut, 0, bricks, 0)
    with pytest.raises(_abi.VrError):
        r.render()
    r2 = api.VolumeRender()
    r2.setRegion(*bricks[0], 2)
    r2 = vrt.example1_scene(lambda: r2, vol, RES, lights=1, lut=lut)
    r2.ElementSizeUm = [1.0, 0.2, 1.0]  # boxmax.y = 0.2: the on-the-fly gradient taps reach 4 voxels out
    import torch

    buf = torch.zeros(RES[0] * RES[1], device="cuda")
    with pytest.raises(_abi.VrError):
        r2.renderOpacity(buf.data_ptr())


@pytest.mark.parametrize("dtype", ["f32", "bf16"])
@pytest.mark.parametrize("splits", [(1, 1, 2), (2, 2, 2)])
@pytest.mark.parametrize("grad", [0, 1])
def test_bricks_fetch_through_the_texture_unit(splits, grad, dtype):
    """Region contexts in the default fetch mode: each brick's resident box (fp32 or bf16 storage) gets fp32 cudaArrays of
    its own and the texture-unit kernels remap the volume coordinate into them (k_raymarch<.., TEXHW, .., UNI = 1>).  Against
    the oracle with the hardware filter model: the 1e-3 bar; against the single-pass texture-unit render of the whole volume:
    the same bar with PSNR >= 95 dB - the remapped coordinate meets the filter's 8-bit weights a hair differently now and then,
    and at a rare pixel that moves the `sum.w > threshold` break (kernel.cu:485) by one sample (isolated pixels)."""
    from test_gpu_bf16 import host_bf16_volume

    vol = vrt.synth_brain((72, 64, 56))
    lut = api.HenyeyGreenstein(64)
    keep = []

    def wrap(a):
        if dtype == "f32":
            return api.Volume(a), a
        v16, v32 = host_bf16_volume(a)
        keep.append(v16)
        return v16, v32

    E, E32 = wrap(vol)
    G = G32 = None
    if grad:
        pairs = [wrap(np.asarray(g.Data)) for g in api.Volume(E32).grad()]
        G, G32 = [p[0] for p in pairs], [p[1] for p in pairs]

    def scene(cls, em, gr, region=None):
        r = cls()
        if region is not None:
            r.setRegion(region[0], region[1], 4)
        r = vrt.exact(vrt.example1_scene(lambda: r, em, RES, lights=2, lut=lut))
        r.FactorAbsorption = 30.0
        if gr is not None:
            r.VolumeGradientX, r.VolumeGradientY, r.VolumeGradientZ = gr
        return r

    whole = scene(api.VolumeRender, E, G)
    full = whole.render()
    assert whole.stats()["fetch_mode"] == 1  # bf16 volumes fetch through the texture unit too (fp32 arrays)
    bricks = multigpu.brick_grid(vol.shape, splits)
    rs = [scene(api.VolumeRender, E, G, b) for b in bricks]
    img = multigpu.render_bricks_one_gpu(rs, bricks)
    assert all(r.stats()["fetch_mode"] == 1 for r in rs)
    m = vrt.compare(img, full)
    print(splits, grad, dtype, "bricks vs single pass, both through the texture unit:", m)
    assert m["max_abs_norm"] <= vrt.TOL and m["psnr_db"] >= 95.0, m
    orc = scene(vrt.OracleVolumeRender, api.Volume(E32), [api.Volume(g) for g in G32] if grad else None)
    orc._backend.lib.orc_set_option(orc.objectHandle, vrt.ORC_OPT_FILTER_FRAC_BITS, 9)
    mo = vrt.compare(img, orc.render())
    print("bricks vs oracle:", mo)
    assert mo["max_abs_norm"] <= vrt.TOL, mo
