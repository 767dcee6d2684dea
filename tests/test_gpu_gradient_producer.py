"""vr_volume_gradient: Volume.grad on the GPU - MATLAB's [gx, gy, gz] = gradient(Data)
(src/matlab/VolumeRender/Volume.m:181-205: unit spacing, central differences inside, one-sided at the ends, 0 along a
singleton axis; gx is the derivative along dim 2, gy along dim 1, gz along dim 3) - against its CPU twin
orc_volume_gradient, BIT FOR BIT.  bench.py's CUDA arm builds Gx, Gy, Gz of the headline workload with this kernel while
the reference arm builds them with the twin: this file is what makes "same inputs on both arms" a tested statement."""
import ctypes as C

import numpy as np
import pytest

import vrt
from test_gpu_bf16 import bf16_bits, bf16_to_f32
from volume_renderer_b200 import _abi, api

pytestmark = pytest.mark.gpu


def cpu_twin(a):
    lib = vrt.oracle()
    lib.orc_volume_gradient.argtypes = [C.c_void_p, C.POINTER(C.c_uint64)] + [C.c_void_p] * 3
    a = np.asfortranarray(a, dtype=np.float32)
    dims = tuple(a.shape) + (1,) * (3 - a.ndim)
    g = [np.zeros(dims, dtype=np.float32, order="F") for _ in range(3)]
    lib.orc_volume_gradient(a.ctypes.data, (C.c_uint64 * 3)(*dims), g[0].ctypes.data, g[1].ctypes.data, g[2].ctypes.data)
    return g


def gpu_gradient(r, a, dtype, in_space, out_space):
    """a: host ndarray (fp32 values, or uint16 bf16 bits) in MATLAB layout.  Returns three host arrays of a's dtype."""
    lib, h = _abi.lib(), r.objectHandle
    dims = tuple(a.shape) + (1,) * (3 - a.ndim)
    nbytes = a.nbytes
    d_in = C.c_void_p()
    v = _abi.VrVolume()
    v.dims[0], v.dims[1], v.dims[2] = dims
    v.dtype, v.last_update = dtype, 1
    if in_space == _abi.MEM_DEVICE:
        _abi.check(lib.vr_device_alloc(h, nbytes, C.byref(d_in)), h)
        _abi.check(lib.vr_device_upload(h, d_in, a.ctypes.data, nbytes), h)
        v.data, v.memspace = d_in.value, _abi.MEM_DEVICE
    else:
        v.data, v.memspace = a.ctypes.data, _abi.MEM_HOST
    outs = [np.zeros(dims, dtype=a.dtype, order="F") for _ in range(3)]
    if out_space == _abi.MEM_DEVICE:
        d_out = [C.c_void_p() for _ in range(3)]
        for p in d_out:
            _abi.check(lib.vr_device_alloc(h, nbytes, C.byref(p)), h)
        _abi.check(lib.vr_volume_gradient(h, C.byref(v), _abi.MEM_DEVICE, *d_out), h)
        for p, o in zip(d_out, outs):
            _abi.check(lib.vr_device_download(h, o.ctypes.data, p, nbytes), h)
            _abi.check(lib.vr_device_free(h, p), h)
    else:
        _abi.check(lib.vr_volume_gradient(h, C.byref(v), _abi.MEM_HOST, *[C.c_void_p(o.ctypes.data) for o in outs]), h)
    if d_in:
        _abi.check(lib.vr_device_free(h, d_in), h)
    return outs


SHAPES = [(33, 27, 19), (64, 64, 64), (40, 24), (1, 17, 5), (2, 2, 2), (7, 1, 1)]  # odd dims, a 2-D "volume", singleton axes


@pytest.mark.parametrize("shape", SHAPES)
@pytest.mark.parametrize("spaces", [(_abi.MEM_HOST, _abi.MEM_HOST), (_abi.MEM_DEVICE, _abi.MEM_DEVICE), (_abi.MEM_HOST, _abi.MEM_DEVICE)])
def test_gradient_f32_is_bit_equal_to_the_cpu_twin(shape, spaces):
    rng = np.random.default_rng(sum(shape))
    a = np.asfortranarray(rng.random(shape, dtype=np.float32) * np.float32(3.0) - np.float32(1.0))
    r = api.VolumeRender()
    got = gpu_gradient(r, a, _abi.DTYPE_F32, *spaces)
    want = cpu_twin(a)
    for name, g, w in zip("xyz", got, want):
        assert np.array_equal(g.view(np.uint32), w.reshape(g.shape).view(np.uint32)), (shape, name, np.abs(g - w.reshape(g.shape)).max())


def test_gradient_matches_matlab_axis_order_and_numpy():
    """gx = d/d dim2, gy = d/d dim1, gz = d/d dim3 (Volume.m:193); values of np.gradient (same stencil) to fp32 rounding."""
    a = vrt.synth_brain((48, 40, 36))
    r = api.VolumeRender()
    gx, gy, gz = gpu_gradient(r, a, _abi.DTYPE_F32, _abi.MEM_HOST, _abi.MEM_HOST)
    n1, n2, n3 = np.gradient(a.astype(np.float64))
    for g, n in ((gx, n2), (gy, n1), (gz, n3)):
        assert np.abs(g - n).max() < 1e-6
    # the Python mirror of Volume.grad hands out the same three volumes
    for g, v in zip((gx, gy, gz), api.Volume(a).grad()):
        assert np.abs(g - np.asarray(v.Data)).max() < 1e-6


@pytest.mark.parametrize("shape", [(33, 27, 19), (40, 24), (64, 48, 32)])
@pytest.mark.parametrize("in_space", [_abi.MEM_HOST, _abi.MEM_DEVICE])
def test_gradient_bf16(shape, in_space):
    """bf16 in, bf16 out: differences are taken in fp32 on the exact bf16 values and rounded to nearest even once."""
    rng = np.random.default_rng(7)
    bits = np.asfortranarray(bf16_bits(rng.random(shape, dtype=np.float32)))
    r = api.VolumeRender()
    got = gpu_gradient(r, bits, _abi.DTYPE_BF16, in_space, _abi.MEM_DEVICE)
    want = cpu_twin(bf16_to_f32(bits))
    for name, g, w in zip("xyz", got, want):
        assert np.array_equal(g, bf16_bits(w).reshape(g.shape)), (shape, name)


def test_headline_workload_gradients_equal_on_both_bench_arms():
    """The exact producer chain of bench.py at a reduced size: vr_synth_volume -> vr_volume_gradient on the device (CUDA arm)
    against orc_synth_volume -> orc_volume_gradient (reference arm): identical bits, volume and gradients."""
    lib = _abi.lib()
    r = api.VolumeRender()
    h = r.objectHandle
    n = 96
    dims = (C.c_uint64 * 3)(n, n, n)
    ptrs = [C.c_void_p() for _ in range(4)]
    for p in ptrs:
        _abi.check(lib.vr_device_alloc(h, n ** 3 * 4, C.byref(p)), h)
    _abi.check(lib.vr_synth_volume(h, 0, vrt.SEED, dims, _abi.DTYPE_F32, ptrs[0]), h)
    v = _abi.VrVolume()
    v.data, v.dtype, v.memspace, v.last_update = ptrs[0].value, _abi.DTYPE_F32, _abi.MEM_DEVICE, 1
    v.dims[0] = v.dims[1] = v.dims[2] = n
    _abi.check(lib.vr_volume_gradient(h, C.byref(v), _abi.MEM_DEVICE, *ptrs[1:]), h)
    dev = []
    for p in ptrs:
        a = np.zeros((n, n, n), dtype=np.float32, order="F")
        _abi.check(lib.vr_device_download(h, a.ctypes.data, p, a.nbytes), h)
        _abi.check(lib.vr_device_free(h, p), h)
        dev.append(a)
    orc = vrt.oracle()
    orc.orc_synth_volume.argtypes = [C.c_int, C.c_uint32, C.POINTER(C.c_uint64), C.c_void_p]
    E = np.zeros((n, n, n), dtype=np.float32, order="F")
    orc.orc_synth_volume(0, vrt.SEED, dims, E.ctypes.data)
    assert np.array_equal(dev[0], E)
    for g, w in zip(dev[1:], cpu_twin(E)):
        assert np.array_equal(g.view(np.uint32), w.view(np.uint32))
