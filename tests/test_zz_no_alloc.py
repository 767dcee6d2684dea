"""vr_henyey_greenstein_static: the Henyey-Greenstein LUT (src/C/mex/HenyeyGreenstein.cc:29-96) through the entry point that
allocates nothing - its device scratch is a file-scope `static __device__` array of the library (64^3 floats), for callers
that cannot call cudaMalloc.  No counterpart in the reference.

Checked through the C ABI against the existing entry point vr_henyey_greenstein (same kernel, its scratch from cudaMalloc):
bit-equal - the tolerance of this repo's bit-level tests - on several calls in a row with different sizes, so that the stale
contents the shared array keeps from the previous call are seen not to leak; and against the CPU oracle's LUT with the bar
of tests/test_gpu_parity.py::test_hg_lut_matches_reference_lut.  This file is collected last on purpose;
it passes alone in a fresh process and after the other tests.
"""
import ctypes as C

import numpy as np
import pytest

import vrt
from volume_renderer_b200 import _abi

pytestmark = pytest.mark.gpu


def _existing(n, g):
    out = np.zeros(n * n * n, dtype=np.float32)
    _abi.check(_abi.lib().vr_henyey_greenstein(n, float(g), out.ctypes.data))
    return out


def _static(n, g, out_ptr, stream=None):
    return _abi.lib().vr_henyey_greenstein_static(n, float(g), C.c_void_p(out_ptr), C.c_void_p(stream or 0))


def test_static_scratch_lut_equals_the_allocating_entry_point_on_calls_in_a_row():
    # 64 fills the whole array; 32 and 7 then reuse its front while the rest keeps the previous call's values; 64 again
    for n, g in ((64, 0.8), (64, 0.8), (32, -0.35), (7, 0.0), (64, 0.5)):
        got = np.full(n * n * n, np.nan, dtype=np.float32)
        assert _static(n, g, got.ctypes.data) == _abi.VR_OK, _abi.lib().vr_last_error(None)
        want = _existing(n, g)
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), (n, g, np.nanmax(np.abs(got - want)))
        assert np.isfinite(got).all() and got.min() > 0
    # the bar the allocating entry point is held to against the reference-bit-equal CPU LUT
    ref = vrt.oracle_lut(64).reshape(-1, order="F")
    got = np.zeros(64 ** 3, dtype=np.float32)
    assert _static(64, 0.8, got.ctypes.data) == _abi.VR_OK
    assert (np.abs(got - ref) / np.abs(ref)).max() < 2e-4 and np.abs(got - ref).max() / ref.max() < 5e-5


def test_static_scratch_lut_into_device_memory_on_a_caller_stream():
    import torch

    stream = torch.cuda.Stream()
    out = torch.full((64 ** 3,), float("nan"), dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    for n, g in ((48, 0.8), (64, 0.8)):
        before = torch.cuda.memory_allocated()
        assert _static(n, g, out.data_ptr(), stream.cuda_stream) == _abi.VR_OK, _abi.lib().vr_last_error(None)
        assert torch.cuda.memory_allocated() == before
        assert np.array_equal(out[: n ** 3].cpu().numpy().view(np.uint32), _existing(n, g).view(np.uint32))


def test_sizes_beyond_the_static_scratch_are_rejected_with_a_clear_error():
    lib = _abi.lib()
    out = np.zeros(8, dtype=np.float32)
    assert _static(_abi.HG_STATIC_MAX_N + 1, 0.8, out.ctypes.data) == 1  # VR_ERR_INVALID_ARGUMENT
    msg = lib.vr_last_error(None).decode()
    assert "exceeds the static scratch" in msg and "64" in msg, msg
    assert _static(0, 0.8, out.ctypes.data) == 1
    assert _static(8, 1.5, out.ctypes.data) == 1 and "[-1,1]" in lib.vr_last_error(None).decode()  # HenyeyGreenstein.cc:44-45
    assert lib.vr_henyey_greenstein_static(8, C.c_float(0.8), None, None) == 1
    assert not out.any()  # a rejected call writes nothing
