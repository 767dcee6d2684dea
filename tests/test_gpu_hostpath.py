"""Host-side paths of the C ABI on the GPU that the parity scenes do not reach: the upload staging ring with several large
pageable volumes in one sync, vr_update_box on a role whose volume is not the one its redirect index names, renders on a
caller's stream right after a LUT upload, a negative opacity threshold, and the plain-C harness (north_star's "plain C
harness so the path can also be driven without a MATLAB licence") against the Python mirror, bit for bit."""
import ctypes as C
import math
import os
import subprocess

import numpy as np
import pytest

import vrt
from volume_renderer_b200 import _abi, api

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lut():
    return api.HenyeyGreenstein(64)


def test_six_large_pageable_volumes_in_one_sync(lut):
    """Six distinct 256^3 fp32 volumes from PAGEABLE host memory in one vr_sync_volumes: each is exactly one 64 MiB slab, so
    every upload starts on pinned staging buffer 0 while the previous volume's DMA out of it may still be in flight.
    The image must equal the one of a context that received the same volumes from device memory (no staging)."""
    import torch

    n = 256
    E = vrt.synth_brain((n, n, n))
    vols = {"E": E, "A": np.asfortranarray(np.sqrt(E)), "R": np.asfortranarray(E * E)}
    g = [np.asarray(v.Data) for v in api.Volume(E).grad()]

    def scene(wrap):
        r = vrt.example1_scene(api.VolumeRender, wrap(vols["E"]), (256, 192), lights=2, lut=lut)
        r.VolumeAbsorption = wrap(vols["A"])
        r.VolumeReflection = wrap(vols["R"])
        r.VolumeGradientX, r.VolumeGradientY, r.VolumeGradientZ = [wrap(x) for x in g]
        r.setOption(_abi.OPT_ABSORPTION_ROLE_FIX, 1)  # every one of the six volumes is sampled
        return r

    keep = []

    def on_device(a):
        t = torch.from_numpy(np.ascontiguousarray(a.reshape(-1, order="F"))).cuda()
        keep.append(t)
        return api.DeviceVolume(t.data_ptr(), a.shape, _abi.DTYPE_F32, _abi.MEM_DEVICE, keepalive=t)

    want = scene(on_device).render()
    torch.cuda.synchronize()
    for attempt in range(3):  # a race needs a few tries to show
        r = scene(lambda a: api.Volume(a))
        got = r.render()
        assert r.stats()["uploads"] >= 6
        assert np.array_equal(got, want), "attempt %d: max diff %g" % (attempt, np.abs(got - want).max())


@pytest.mark.parametrize("fix", [0, 1])
def test_update_box_on_a_distinct_absorption_volume(lut, fix):
    """Distinct Em and Ab volumes: as written d_idxAbsorption keeps naming the emission texture (kernel.cu:80), but the
    absorption VOLUME lives in its own slot.  vr_update_box(role = absorption) must rewrite that slot - and leave the
    emission texture alone - exactly like a full sync of the dirty volume would."""
    E = vrt.synth_brain((56, 48, 40))
    A0 = np.asfortranarray(np.sqrt(E))

    def scene(ab):
        r = vrt.example1_scene(api.VolumeRender, np.array(E, order="F"), (128, 96), lights=1, lut=lut)
        r.VolumeAbsorption = api.Volume(np.array(ab, order="F"))
        r.setOption(_abi.OPT_ABSORPTION_ROLE_FIX, fix)
        return r

    r = scene(A0)
    before = r.render()
    A1 = np.array(A0, order="F")
    A1[:, :, :20] *= np.float32(0.25)
    r.VolumeAbsorption.Data[:, :, :20] = A1[:, :, :20]
    r.updateVolumeBox("absorption", (0, 0, 0), (56, 48, 20))
    after = r.render()
    assert r.stats()["uploads"] == 0
    want = scene(A1).render()
    assert np.array_equal(after, want)
    if fix:
        assert not np.array_equal(after, before)  # the absorption role reads the updated slot
    else:
        assert np.array_equal(after, before)  # as written absorption is sampled from emission, which must be untouched
    # and a full re-sync afterwards changes nothing
    r.VolumeAbsorption = api.Volume(A1)
    assert np.array_equal(r.render(), want)


def test_render_on_a_caller_stream_sees_the_lut_it_uploads(lut):
    """vr_render_device on a stream that is not the context's: the first frame uploads the LUT on the context's stream and
    must still be ordered behind it."""
    import torch

    vol = vrt.synth_brain((48, 48, 48))
    want = vrt.example1_scene(api.VolumeRender, vol, (128, 96), lights=2, lut=lut).render()
    side = torch.cuda.Stream()
    lib = _abi.lib()
    for attempt in range(3):
        r = vrt.example1_scene(api.VolumeRender, vol, (128, 96), lights=2, lut=np.array(lut, order="F"))
        h = r.objectHandle
        out = torch.empty(want.size, dtype=torch.float32, device="cuda")
        args, keep = r._packed_mono_args()
        _abi.check(lib.vr_render_device(h, *args, C.c_void_p(out.data_ptr()), C.c_void_p(side.cuda_stream)), h)
        side.synchronize()
        assert np.array_equal(out.cpu().numpy().reshape(want.shape, order="F"), want)


@pytest.mark.parametrize("texture_unit", [2, 0])
def test_negative_opacity_threshold_stops_after_the_first_sample(lut, texture_unit):
    """sum.w > threshold (kernel.cu:485) with a negative threshold ends every ray after its first sample, wherever that
    sample is - also inside empty space, which the skipping must therefore not jump over."""
    vol = vrt.synth_brain((40, 40, 40))
    gpu = vrt.exact(vrt.example1_scene(api.VolumeRender, vol, (96, 96), lights=1, lut=lut))
    orc = vrt.example1_scene(vrt.OracleVolumeRender, vol, (96, 96), lights=1, lut=lut)
    for r in (gpu, orc):
        r.OpacityThreshold = -0.5
        r.DistanceToObject = 0.3  # the camera sits inside the volume: first samples are not all empty
    gpu.setOption(_abi.OPT_TEXTURE_UNIT, texture_unit)
    if texture_unit:
        orc._backend.lib.orc_set_option(orc.objectHandle, vrt.ORC_OPT_FILTER_FRAC_BITS, 9)
    img, ref = gpu.render(), orc.render()
    m = vrt.compare(img, ref)
    assert m["max_abs_norm"] <= vrt.TOL, m
    st = gpu.stats()
    assert st["samples_marched"] == st["rays_hit"] == orc._backend.stats(orc.objectHandle).samples_marched


# ---- the plain-C harness --------------------------------------------------------------------------------
def _read_pfm(path):
    with open(path, "rb") as f:
        assert f.readline().strip() == b"PF"
        W, H = [int(x) for x in f.readline().split()]
        assert float(f.readline()) < 0  # little endian
        data = np.frombuffer(f.read(), dtype="<f4").reshape(H, W, 3)
    return np.asfortranarray(data[::-1])  # PFM rows go bottom-up


def _rotation_like_the_harness(al, be, ga):
    """harness/vrbench.c: rotate() - R = I * Rx * Ry * Rz with the products summed in the harness's order (doubles)."""
    k = 3.14159265358979323846 / 180.0
    ca, sa, cb, sb, cg, sg = math.cos(al * k), math.sin(al * k), math.cos(be * k), math.sin(be * k), math.cos(ga * k), math.sin(ga * k)
    R = [1.0, 0, 0, 0, 1.0, 0, 0, 0, 1.0]
    for B in ([1, 0, 0, 0, ca, -sa, 0, sa, ca], [cb, 0, sb, 0, 1, 0, -sb, 0, cb], [cg, -sg, 0, sg, cg, 0, 0, 0, 1]):
        R = [R[3 * i] * B[j] + R[3 * i + 1] * B[3 + j] + R[3 * i + 2] * B[6 + j] for i in range(3) for j in range(3)]
    return np.array(R).reshape(3, 3)


@pytest.mark.parametrize("flags,lights,grad", [([], 1, False), (["--grad", "--lights", "2"], 2, True)])
def test_plain_c_harness_equals_the_python_mirror(tmp_path, lut, flags, lights, grad):
    exe = os.path.join(ROOT, "harness", "vrbench")
    if not os.path.exists(exe):
        pytest.fail("harness/vrbench is not built (python -c 'import __graft_entry__ as g; g.build()')")
    out = str(tmp_path / "frame.pfm")
    n, W, H = 64, 160, 96
    p = subprocess.run([exe, "--size", str(n), "--image", str(W), str(H), "--frames", "1", "--out", out] + (flags or ["--lights", "1"]),
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300)
    print(p.stdout)
    assert p.returncode == 0, p.stdout
    assert "Gsamples/s" in p.stdout and "Memory Information" in p.stdout
    img = _read_pfm(out)
    assert img.shape == (H, W, 3)
    vol = vrt.synth_brain((n, n, n))
    r = vrt.example1_scene(api.VolumeRender, vol, (W, H), lights=lights, lut=lut)
    r.RotationMatrix = _rotation_like_the_harness(125, 25, 0)
    if grad:
        # the harness takes Gx, Gy, Gz from vr_volume_gradient: bit-equal to the twin of MATLAB's gradient (test_gpu_gradient_producer)
        from test_gpu_gradient_producer import cpu_twin

        r.VolumeGradientX, r.VolumeGradientY, r.VolumeGradientZ = [api.Volume(g) for g in cpu_twin(vol)]
    want = r.render()
    assert np.array_equal(img, want), np.abs(img - want).max()
    kernel_samples = int(p.stdout.split("samples marched")[0].split(",")[-1])
    assert kernel_samples == r.stats()["samples_marched"]


def test_texture_unit_mode_keeps_one_copy_per_volume(lut):
    """VR_OPT_TEXTURE_SINGLE_COPY (default on): once a texture-unit render has its cudaArrays and skip grids, the padded brick
    storages they were built from are freed - bytes_resident ~ 1x the voxels, not 2x - and everything that needs a storage
    again (the load path, vr_update_box, new skip grids) gets the bits an upload would have written."""
    import torch

    n = 128
    E = vrt.synth_brain((n, n, n))
    g = [np.asarray(v.Data) for v in api.Volume(E).grad()]
    keep = []

    def on_device(a):
        t = torch.from_numpy(np.ascontiguousarray(a.reshape(-1, order="F"))).cuda()
        keep.append(t)
        return api.DeviceVolume(t.data_ptr(), a.shape, _abi.DTYPE_F32, _abi.MEM_DEVICE, keepalive=t)

    def scene(single_copy=1, texture_unit=2):
        r = vrt.example1_scene(api.VolumeRender, on_device(E), (160, 128), lights=2, lut=lut)  # device sources: no staging buffers
        r.VolumeGradientX, r.VolumeGradientY, r.VolumeGradientZ = [on_device(x) for x in g]
        r.setOption(_abi.OPT_TEXTURE_SINGLE_COPY, single_copy)
        r.setOption(_abi.OPT_TEXTURE_UNIT, texture_unit)
        return r

    voxels = 4 * n ** 3 * 4  # E, Gx, Gy, Gz
    one, two = scene(1), scene(0)
    img = one.render()
    assert np.array_equal(two.render(), img)
    b1, b2 = one.stats()["bytes_resident"], two.stats()["bytes_resident"]
    print("resident bytes / voxel bytes: single copy %.3f, both copies %.3f" % (b1 / voxels, b2 / voxels))
    assert b1 < 1.15 * voxels and b2 > 2.2 * voxels  # (both copies: the interleaved gradient storage is 4 elements per voxel)
    assert one.stats()["fetch_mode"] == 1
    # frames after the first change nothing (nothing left to drop, nothing rebuilt)
    one.rotate(0, 10, 0), two.rotate(0, 10, 0)
    assert np.array_equal(one.render(), two.render()) and one.stats()["bytes_resident"] == b1
    # the load path needs the storages: rebuilt from the arrays, same bits as a context that never dropped them
    loads = scene(1, texture_unit=0)
    loads.rotate(0, 10, 0)
    want_loads = loads.render()
    one.setOption(_abi.OPT_TEXTURE_UNIT, 0)
    assert np.array_equal(one.render(), want_loads) and one.stats()["fetch_mode"] == 0
    assert one.stats()["bytes_resident"] > 2.2 * voxels
    # ... and back: dropped again
    one.setOption(_abi.OPT_TEXTURE_UNIT, 2)
    assert np.array_equal(one.render(), two.render()) and one.stats()["bytes_resident"] == b1
    # the skip grids are dropped by a partial update and rebuilt from a storage that was itself rebuilt from the array
    for r in (one, two):
        r.updateVolumeBox("emission", (0, 0, 0), (n, n, 8))  # same data: the image must not change
    assert np.array_equal(one.render(), two.render())
    assert one.stats()["bytes_resident"] > b1  # a volume that is updated in place keeps both copies from now on


def test_plain_c_harness_on_several_gpus_equals_one_gpu(tmp_path):
    """harness/vrbench --gpus G: vr_set_devices behind the same 'render' call, driven from plain C (no torch, no MPI)."""
    import torch

    g = torch.cuda.device_count()
    if g < 2:
        pytest.skip("one GPU visible (tests/test_gpu_group.py covers device groups whose members share a GPU)")
    exe = os.path.join(ROOT, "harness", "vrbench")
    imgs = []
    for gpus in (1, g):
        out = str(tmp_path / ("frame_%d.pfm" % gpus))
        p = subprocess.run([exe, "--size", "96", "--image", "200", "120", "--frames", "2", "--grad", "--lights", "2", "--gpus", str(gpus), "--out", out],
                           stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
        print(p.stdout)
        assert p.returncode == 0, p.stdout
        imgs.append(_read_pfm(out))
        if gpus > 1:
            assert "device group: %d members" % gpus in p.stdout and "peer access yes" in p.stdout
    assert np.array_equal(imgs[0], imgs[1])


def test_plain_c_harness_latency_mode(tmp_path):
    """harness/vrbench --ray-slices S: VR_OPT_RAY_SLICES from plain C - every frame of the harness is one 'render' call."""
    exe = os.path.join(ROOT, "harness", "vrbench")
    imgs = []
    for extra in ([], ["--ray-slices", "4"]):
        out = str(tmp_path / ("frame_%d.pfm" % len(extra)))
        p = subprocess.run([exe, "--size", "96", "--image", "200", "120", "--frames", "2", "--grad", "--lights", "2", "--out", out] + extra,
                           stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
        assert p.returncode == 0, p.stdout
        imgs.append(_read_pfm(out))
    assert imgs[0].max() > 0 and not np.array_equal(imgs[0], imgs[1])  # (the folds round differently: the option did something)
    assert np.abs(imgs[0] - imgs[1]).max() / imgs[0].max() < 2e-5
