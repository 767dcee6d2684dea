"""Several GPUs behind the reference's own boundary: the column partition (VR_OPT_PART_MODE = 1), the pipelined batches
(vr_render_batch_begin / _end) and device groups (vr_set_devices).  Everything here must be BIT-IDENTICAL to one
unpartitioned context (image-space mode renders the same per-pixel code).  With one GPU visible the group
members share it (a device may be listed more than once); with two or more the group spans real devices."""
import ctypes as C

import numpy as np
import pytest

import vrt
from volume_renderer_b200 import _abi, api, multigpu

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def lut():
    return api.HenyeyGreenstein(64)


@pytest.fixture(scope="module")
def brain():
    return vrt.synth_brain((64, 56, 48))


def _device_ids(n):
    import torch

    have = torch.cuda.device_count()
    return [i % have for i in range(n)]


@pytest.mark.parametrize("res", [(192, 160), (203, 97)])  # the second: a clipped last tile column, odd height
@pytest.mark.parametrize("nranks", [2, 3, 8])
def test_column_partition_writes_only_owned_pixels_of_a_host_image(brain, lut, res, nranks):
    full = vrt.example1_scene(api.VolumeRender, brain, res, lights=2, lut=lut).render()
    shared = np.full(full.shape, 7.0, dtype=np.float32, order="F")  # poisoned: every pixel must be written by exactly one rank
    lib = _abi.lib()
    for rank in range(nranks):
        r = vrt.example1_scene(api.VolumeRender, brain, res, lights=2, lut=lut)
        multigpu.configure_partition(r.objectHandle, rank, nranks, 32, 32, mode=1)
        before = shared.copy()
        args, keep = r._packed_mono_args()
        _abi.check(lib.vr_render(r.objectHandle, *args, C.c_void_p(shared.ctypes.data)), r.objectHandle)
        mask = multigpu.owned_mask(res[1], res[0], 32, 32, rank, nranks, mode=1)
        assert np.array_equal(shared[~mask], before[~mask])  # nothing outside the owned columns was touched
        assert np.array_equal(shared[mask], full[mask])
    assert np.array_equal(shared, full)


def test_pipelined_batches_equal_single_frames_and_sum_their_counters(brain, lut):
    import torch

    r = vrt.example1_scene(api.VolumeRender, brain, (160, 128), lights=1, lut=lut)
    rots, singles, marched = [], [], 0
    for _ in range(7):
        r.rotate(0, 11, 0)
        rots.append(r.RotationMatrix.copy())
        singles.append(r.render())
        marched += r.stats()["samples_marched"]
    pinned = [torch.zeros(128 * 160 * 3, dtype=torch.float32).pin_memory() for _ in rots]
    outs = [p.numpy().reshape((128, 160, 3), order="F") for p in pinned]
    for batch in (3, 1, 7):
        for o in outs:
            o[...] = -1.0
        r.renderMoviePipelined(rots, batch=batch, out=outs)
        for a, b in zip(singles, outs):
            assert np.array_equal(a, b)
        st = r.stats()
        assert st["samples_marched"] == marched and st["kernel_launches"] == -(-7 // batch) and st["kernel_ms"] > 0
    # pageable destinations work too (the copies are synchronous then)
    for a, b in zip(singles, r.renderMoviePipelined(rots, batch=4)):
        assert np.array_equal(a, b)
    # a synchronous render right after a pipeline reports its own counters again
    r.RotationMatrix = rots[0]
    assert np.array_equal(r.render(), singles[0]) and r.stats()["kernel_launches"] == 1


@pytest.mark.parametrize("n", [2, 3])
def test_device_group_host_and_device_images_are_bit_identical(brain, lut, n):
    import torch

    res = (203, 160)
    single = vrt.example1_scene(api.VolumeRender, brain, res, lights=2, lut=lut)
    gx, gy, gz = single.VolumeEmission.grad()
    single.VolumeGradientX, single.VolumeGradientY, single.VolumeGradientZ = gx, gy, gz
    want = single.render()
    s1 = single.stats()

    g = api.VolumeRender()
    g.setOption(_abi.OPT_TILE_W, 32)
    g.setDevices(_device_ids(n))
    g = vrt.example1_scene(lambda: g, brain, res, lights=2, lut=lut)
    g.VolumeGradientX, g.VolumeGradientY, g.VolumeGradientZ = [api.Volume(np.asarray(v.Data)) for v in (gx, gy, gz)]
    got = g.render()  # volumeRender('render', ...): the assembled H x W x 3 image, as render.cpp:259-271 returns it
    assert np.array_equal(got, want)
    st = g.stats()
    assert st["samples_marched"] == s1["samples_marched"] and st["rays_hit"] == s1["rays_hit"] and st["kernel_launches"] == n
    assert st["bytes_resident"] > 1.5 * s1["bytes_resident"]
    assert "device group: %d members" % n in g._backend.mem_info(g.objectHandle)
    # a dirty volume reaches every member; a partial update too
    for r in (single, g):
        d = r.VolumeEmission.Data
        d[:, :, :20] = d[:, :, :20] * np.float32(0.5)
    assert np.array_equal(g.render(), single.render())
    for r in (single, g):
        r.VolumeEmission.Data.view(np.ndarray)[:, :10, :] = 0.25  # in place, no stamp
        r.updateVolumeBox("emission", (0, 0, 0), (64, 10, 48))
    want2 = single.render()
    assert np.array_equal(g.render(), want2)
    # device image: the kernels of all members store into the caller's buffer
    lib, h = _abi.lib(), g.objectHandle
    buf = torch.full((want2.size,), 7.0, dtype=torch.float32, device="cuda:0")
    torch.cuda.synchronize()
    args, keep = g._packed_mono_args()
    _abi.check(lib.vr_render_device(h, *args, C.c_void_p(buf.data_ptr()), None), h)
    g.stats()  # waits for every member
    assert np.array_equal(buf.cpu().numpy().reshape(want2.shape, order="F"), want2)
    # a movie through the group, pipelined
    rots = []
    for _ in range(4):
        single.rotate(0, 15, 0)
        rots.append(single.RotationMatrix.copy())
    for a, b in zip(single.renderMovie(rots), g.renderMoviePipelined(rots, batch=2)):
        assert np.array_equal(a, b)
    # options set on the leader reach the members; the partition itself is the group's
    g.setOption(_abi.OPT_TEXTURE_UNIT, 0)
    single.setOption(_abi.OPT_TEXTURE_UNIT, 0)
    g.RotationMatrix = single.RotationMatrix
    assert np.array_equal(g.render(), single.render())
    with pytest.raises(_abi.VrError):
        g.setOption(_abi.OPT_PART_RANK, 1)
    # dissolving the group gives the plain context back
    g.setDevices([_device_ids(1)[0]])
    assert np.array_equal(g.render(), single.render()) and g.stats()["kernel_launches"] == 1


def test_two_real_devices_when_the_box_has_them(brain, lut):
    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("one GPU visible: the group tests above share it")
    want = vrt.example1_scene(api.VolumeRender, brain, (256, 192), lights=2, lut=lut).render()
    n = torch.cuda.device_count()
    g = api.VolumeRender()
    g.setDevices(list(range(n)))
    g = vrt.example1_scene(lambda: g, brain, (256, 192), lights=2, lut=lut)
    assert np.array_equal(g.render(), want)
    assert "peer access yes" in g._backend.mem_info(g.objectHandle)
