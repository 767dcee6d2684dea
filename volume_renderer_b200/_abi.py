"""ctypes binding of the C ABI in include/vr_abi.h (libvr_b200.so).

There is no CPU fallback: if the CUDA library is missing or no B200 is visible the calls raise.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("VR_B200_LIB") or os.path.join(_HERE, "csrc", "libvr_b200.so")  # env: developer A/B builds

VR_OK = 0
STATUS_NAMES = {
    0: "VR_OK",
    1: "VR_ERR_INVALID_ARGUMENT",
    2: "VR_ERR_INVALID_HANDLE",
    3: "VR_ERR_INSUFFICIENT_VRAM",
    4: "VR_ERR_CUDA",
    5: "VR_ERR_NO_DEVICE",
    6: "VR_ERR_UNSUPPORTED",
}
DTYPE_F32, DTYPE_BF16 = 0, 1
MEM_HOST, MEM_DEVICE, MEM_DEVICE_REGION = 0, 1, 2

# vr_option
OPT_FILTER_FRAC_BITS = 0
OPT_EMPTY_SPACE_SKIP = 1
OPT_KERNEL_VARIANT = 2
OPT_ABSORPTION_ROLE_FIX = 3
OPT_PART_RANK = 4
OPT_PART_NRANKS = 5
OPT_TILE_W = 6
OPT_TILE_H = 7
OPT_BRICK_LAYOUT = 8
OPT_COUNT_SAMPLES = 9
OPT_CLEAR_IMAGE = 10
OPT_TILE_ORDER = 11
OPT_WARP_SHAPE = 12
OPT_GRADIENT_INTERLEAVE = 13
OPT_TEXTURE_UNIT = 14
OPT_ARITHMETIC = 15
OPT_PART_MODE = 16
OPT_TEXTURE_SINGLE_COPY = 17
OPT_VS_SLICES = 18
OPT_RAY_SLICES = 19
ARITH_EXACT, ARITH_REFERENCE, ARITH_FAST = 0, 1, 2
IPC_HANDLE_BYTES = 64
HG_STATIC_MAX_N = 64


class VrVolume(C.Structure):
    _fields_ = [
        ("data", C.c_void_p),
        ("dims", C.c_uint64 * 3),
        ("last_update", C.c_uint64),
        ("dtype", C.c_int32),
        ("memspace", C.c_int32),
    ]


class VrLight(C.Structure):
    _fields_ = [("position", C.c_float * 3), ("color", C.c_float * 3)]


class VrStats(C.Structure):
    _fields_ = [
        ("kernel_ms", C.c_double),
        ("upload_ms", C.c_double),
        ("samples_marched", C.c_uint64),
        ("samples_shaded", C.c_uint64),
        ("rays_hit", C.c_uint64),
        ("bytes_resident", C.c_uint64),
        ("bytes_uploaded", C.c_uint64),
        ("kernel_launches", C.c_uint32),
        ("uploads", C.c_uint32),
        ("fetch_mode", C.c_uint32),
        ("arithmetic", C.c_uint32),
        ("vs_slices", C.c_uint32),
        ("reserved_", C.c_uint32),
    ]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


class VrError(RuntimeError):
    def __init__(self, status, message):
        super().__init__("%s: %s" % (STATUS_NAMES.get(status, status), message))
        self.status = status
        self.message = message


_lib = None


def declare(lib):
    """Attach argtypes/restypes for every symbol of include/vr_abi.h."""
    P = C.POINTER
    vp, f3, u64p = C.c_void_p, P(C.c_float), P(C.c_uint64)
    render_args = [vp, P(VrLight), C.c_size_t, P(VrVolume), f3, f3, u64p, f3, f3, C.c_float, f3]
    sig = {
        "vr_create": ([P(vp)], C.c_int),
        "vr_create_on_device": ([P(vp), C.c_int], C.c_int),
        "vr_destroy": ([vp], C.c_int),
        "vr_mem_info": ([vp, C.c_char_p, C.c_size_t, P(C.c_size_t)], C.c_int),
        "vr_sync_volumes": ([vp, C.c_uint64] + [P(VrVolume)] * 6, C.c_int),
        "vr_render": (render_args + [vp], C.c_int),
        "vr_render_device": (render_args + [vp, vp], C.c_int),
        "vr_render_batch": ([vp, P(VrLight), C.c_size_t, P(VrVolume), f3, f3, u64p, C.c_size_t, f3, f3, C.c_float, f3, P(vp)], C.c_int),
        "vr_render_batch_device": ([vp, P(VrLight), C.c_size_t, P(VrVolume), f3, f3, u64p, C.c_size_t, f3, f3, C.c_float, f3, P(vp), vp], C.c_int),
        "vr_render_batch_begin": ([vp, P(VrLight), C.c_size_t, P(VrVolume), f3, f3, u64p, C.c_size_t, f3, f3, C.c_float, f3, P(vp)], C.c_int),
        "vr_render_batch_end": ([vp], C.c_int),
        "vr_set_devices": ([vp, C.c_uint32, P(C.c_int)], C.c_int),
        "vr_henyey_greenstein": ([C.c_uint32, C.c_float, vp], C.c_int),
        "vr_henyey_greenstein_static": ([C.c_uint32, C.c_float, vp, vp], C.c_int),
        "vr_timestamp_ms": ([C.c_int], C.c_uint64),
        "vr_last_error": ([vp], C.c_char_p),
        "vr_status_string": ([C.c_int], C.c_char_p),
        "vr_abi_version": ([], C.c_int),
        "vr_set_option": ([vp, C.c_int, C.c_int64], C.c_int),
        "vr_get_option": ([vp, C.c_int, P(C.c_int64)], C.c_int),
        "vr_get_stats": ([vp, P(VrStats)], C.c_int),
        "vr_set_stream": ([vp, vp], C.c_int),
        "vr_get_device": ([vp, P(C.c_int)], C.c_int),
        "vr_synth_volume": ([vp, C.c_int, C.c_uint32, u64p, C.c_int32, vp], C.c_int),
        "vr_volume_gradient": ([vp, P(VrVolume), C.c_int32, vp, vp, vp], C.c_int),
        "vr_set_region": ([vp, u64p, u64p, C.c_uint32], C.c_int),
        "vr_update_box": ([vp, C.c_int, P(VrVolume), u64p, u64p], C.c_int),
        "vr_render_opacity_device": (render_args + [vp, vp], C.c_int),
        "vr_render_partial_device": (render_args + [C.c_uint32, C.c_uint32, u64p, vp, vp, vp], C.c_int),
        "vr_compose_partials_device": ([vp, C.c_uint32, P(vp), C.c_uint64, vp, vp], C.c_int),
        "vr_synth_volume_region": ([vp, C.c_int, C.c_uint32, u64p, u64p, u64p, C.c_int32, vp], C.c_int),
        "vr_compose_stereo_device": ([vp, vp, vp, u64p, C.c_uint64, C.c_int, vp, vp], C.c_int),
        "vr_normalize_sequence_device": ([vp, P(vp), C.c_uint32, C.c_uint64, vp], C.c_int),
        "vr_device_upload": ([vp, vp, vp, C.c_size_t], C.c_int),
        "vr_debug_warp_times": ([vp, vp, C.c_size_t], C.c_int),
        "vr_ipc_export": ([vp, vp, C.c_char_p], C.c_int),
        "vr_ipc_open": ([vp, C.c_char_p, P(vp)], C.c_int),
        "vr_ipc_close": ([vp, vp], C.c_int),
        "vr_device_alloc": ([vp, C.c_size_t, P(vp)], C.c_int),
        "vr_device_free": ([vp, vp], C.c_int),
        "vr_device_download": ([vp, vp, vp, C.c_size_t], C.c_int),
        "vr_probe_gather": ([vp, C.c_int, C.c_int, P(C.c_double), P(C.c_double)], C.c_int),
    }
    for name, (args, res) in sig.items():
        fn = getattr(lib, name)
        fn.argtypes = args
        fn.restype = res
    return sig


def lib():
    """Load libvr_b200.so (built in-tree by __graft_entry__.build()).  Fails loudly."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                "CUDA library %s is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                "(there is no CPU fallback)" % LIB_PATH
            )
        l = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
        declare(l)
        _lib = l
    return _lib


def check(status, ctx=None):
    if status != VR_OK:
        msg = lib().vr_last_error(ctx)
        raise VrError(status, msg.decode() if msg else "")


def volume_struct(data, last_update, keepalive=None):
    """vr_volume for a host numpy array in MATLAB (column-major) layout or a device buffer.

    `data` is a float32 ndarray with up to 3 dims (Fortran order is passed zero-copy; anything
    else is copied) or a tuple (device_ptr, dims, dtype) for memory already on the GPU.
    """
    v = VrVolume()
    if isinstance(data, tuple):
        ptr, dims, dtype = data[:3]
        v.data = ptr
        v.memspace = data[3] if len(data) > 3 else MEM_DEVICE
        v.dtype = dtype
    else:
        a = np.asarray(data)
        if a.dtype != np.float32 or not a.flags.f_contiguous:
            a = np.asfortranarray(a, dtype=np.float32)
        if keepalive is not None:
            keepalive.append(a)
        dims = tuple(a.shape) + (1,) * (3 - a.ndim)
        # mxMake_volume: only a 3-D array has depth, anything else is depth 1 (volumeRender.cpp:322-339)
        v.data = a.ctypes.data
        v.memspace = MEM_HOST
        v.dtype = DTYPE_F32
    v.dims[0], v.dims[1], v.dims[2] = int(dims[0]), int(dims[1]), int(dims[2])
    v.last_update = int(last_update)
    return v
