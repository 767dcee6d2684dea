"""VR_OPT_RAY_SLICES: one plain frame rendered as an opacity pass and a shading pass, both cut into depth slices marched by
different warps - what a single frame spread over several GPUs needs.  No counterpart in the reference (its kernel
marches one thread per pixel, volumeRender_kernel.cu:382-508); the bar is the plain march of this library: equal up to the
rounding of the slices' folds, every sample shaded exactly once, bit-identical under any image-space partition."""
import ctypes as C

import numpy as np
import pytest

import vrt
from volume_renderer_b200 import _abi, api, multigpu

pytestmark = pytest.mark.gpu
RES = (203, 160)  # a clipped last tile column


@pytest.fixture(scope="module")
def lut():
    return api.HenyeyGreenstein(64)


@pytest.fixture(scope="module")
def brain():
    return vrt.synth_brain((72, 64, 56))


def _scene(vol, lut, tu, grad, slices, absorption=None):
    r = vrt.example1_scene(api.VolumeRender, vol, RES, lights=2, lut=lut)
    r.setOption(_abi.OPT_TEXTURE_UNIT, tu)
    if tu == 0:
        vrt.exact(r)
    if grad:
        gx, gy, gz = api.Volume(vol).grad()
        r.VolumeGradientX, r.VolumeGradientY, r.VolumeGradientZ = gx, gy, gz
    if absorption is not None:
        r.FactorAbsorption = absorption
    r.setOption(_abi.OPT_RAY_SLICES, slices)
    return r


@pytest.mark.parametrize("tu", [0, 2])
@pytest.mark.parametrize("grad", [0, 1])
@pytest.mark.parametrize("slices", [2, 5, 16])
def test_sliced_frame_equals_the_plain_march(brain, lut, tu, grad, slices):
    for absorption in (4.0, 30.0):  # translucent / rays that end at the 0.9 threshold inside the volume
        plain = _scene(brain, lut, tu, grad, 1, absorption)
        want = plain.render()
        sp = plain.stats()
        r = _scene(brain, lut, tu, grad, slices, absorption)
        got = r.render()
        st = r.stats()
        m = vrt.compare(got, want)
        print(tu, grad, slices, absorption, m, st["samples_shaded"], sp["samples_shaded"])
        assert m["max_abs_norm"] < 2e-5, m
        assert st["vs_slices"] == slices and st["kernel_launches"] == 5 and st["fetch_mode"] == sp["fetch_mode"]
        assert st["rays_hit"] == sp["rays_hit"]
        if absorption == 4.0:
            assert st["samples_shaded"] == sp["samples_shaded"] > 0
        assert st["kernel_ms"] > 0
        # a second frame reuses every buffer; a movie (batches of several frames) is rendered plainly
        r.rotate(0, 20, 0)
        plain.rotate(0, 20, 0)
        assert vrt.compare(r.render(), plain.render())["max_abs_norm"] < 2e-5


@pytest.mark.parametrize("mode", [0, 1])
@pytest.mark.parametrize("tu", [0, 2])
def test_partitioned_sliced_frames_assemble_bit_identically(brain, lut, mode, tu):
    import torch

    full = _scene(brain, lut, tu, 0, 4).render()
    buf = torch.full((full.size,), 7.0, dtype=torch.float32, device="cuda")  # poisoned: every pixel is written by exactly one rank
    torch.cuda.synchronize()
    for rank in range(3):
        r = _scene(brain, lut, tu, 0, 4)
        r.syncVolumes()
        multigpu.configure_partition(r.objectHandle, rank, 3, 32, 32, mode=mode)
        r.setOption(_abi.OPT_CLEAR_IMAGE, 0)
        args, keep = r._packed_mono_args()
        _abi.check(_abi.lib().vr_render_device(r.objectHandle, *args, C.c_void_p(buf.data_ptr()), None), r.objectHandle)
        assert r.stats()["vs_slices"] == 4
    out = buf.cpu().numpy().reshape(full.shape, order="F")
    assert np.array_equal(out, full)


def test_device_group_renders_sliced_frames(brain, lut):
    import torch

    have = torch.cuda.device_count()
    want = _scene(brain, lut, 2, 1, 6).render()
    g = api.VolumeRender()
    g.setOption(_abi.OPT_TILE_W, 32)
    g.setDevices([i % have for i in range(3)])
    g = vrt.example1_scene(lambda: g, brain, RES, lights=2, lut=lut)
    gx, gy, gz = api.Volume(brain).grad()
    g.VolumeGradientX, g.VolumeGradientY, g.VolumeGradientZ = gx, gy, gz
    g.setOption(_abi.OPT_RAY_SLICES, 6)
    assert np.array_equal(g.render(), want)
    assert g.stats()["kernel_launches"] == 15


def test_scenes_the_passes_do_not_take_are_rendered_plainly(brain, lut):
    # an absorption volume of another size: the volume-space passes want one geometry
    def scene(slices):
        r = _scene(brain, lut, 2, 0, slices)
        r.VolumeAbsorption = api.Volume(np.ascontiguousarray(brain[::2, ::2, ::2]))
        r.setOption(_abi.OPT_ABSORPTION_ROLE_FIX, 1)
        return r

    want = scene(1).render()
    r = scene(8)
    assert np.array_equal(r.render(), want)
    st = r.stats()
    assert st["vs_slices"] == 1 and st["kernel_launches"] == 1
    with pytest.raises(_abi.VrError):
        r.setOption(_abi.OPT_RAY_SLICES, 17)
