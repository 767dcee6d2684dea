"""Host-side logic that needs no GPU: the C-ABI library loads and exports every symbol the header
declares, fails loudly without a device, and the Python mirror of the MATLAB classes behaves
like src/matlab/VolumeRender/*.m (dirty tracking, rotate, stereo geometry, normalisation)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import vrt
from volume_renderer_b200 import _abi, api, multigpu

ROOT = vrt.ROOT


def _header_symbols():
    text = open(os.path.join(ROOT, "include", "vr_abi.h")).read()
    return sorted(set(re.findall(r"VR_API[^;(]*?\b(vr_[a-z_0-9]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    lib = _abi.lib()  # raises ImportError when libvr_b200.so is missing: there is no fallback
    names = _header_symbols()
    assert len(names) >= 22, names
    declared = _abi.declare(lib)
    for n in names:
        assert hasattr(lib, n), n
        assert n in declared, "ctypes signature missing for %s" % n
    assert lib.vr_abi_version() == 1
    assert lib.vr_status_string(3) == b"VR_ERR_INSUFFICIENT_VRAM"
    assert lib.vr_timestamp_ms(0) > 1_600_000_000_000 and lib.vr_timestamp_ms(1) < 2 ** 32


def test_no_cpu_fallback_without_a_device():
    import torch

    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    lib = _abi.lib()
    h = C.c_void_p()
    assert lib.vr_create(C.byref(h)) == 5  # VR_ERR_NO_DEVICE
    assert b"no CPU fallback" in lib.vr_last_error(None)
    out = np.zeros(8, dtype=np.float32)
    assert lib.vr_henyey_greenstein(2, 0.8, out.ctypes.data) == 5
    with pytest.raises(_abi.VrError):
        api.VolumeRender()
    assert lib.vr_destroy(C.c_void_p(0x1234)) == 2  # VR_ERR_INVALID_HANDLE, no crash


def test_product_never_touches_the_oracle():
    """Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may load anything under oracle/."""
    bad = []
    for base in ("volume_renderer_b200", "harness", "mex", "include"):
        for dp, _, files in os.walk(os.path.join(ROOT, base)):
            for f in files:
                if f.endswith((".py", ".cu", ".h", ".c", ".cpp", ".hpp")) or f == "Makefile":
                    if re.search(r"oracle/|liboracle|orc_", open(os.path.join(dp, f), errors="ignore").read()):
                        bad.append(os.path.join(dp, f))
    assert not bad, bad


# ---- MATLAB class mirror ---------------------------------------------------------------------------
def test_volume_stamps_on_assignment_and_indexed_write():
    t = [1000]
    api._ts_clock = lambda: t[0]
    try:
        v = api.Volume(np.zeros((4, 4, 4)))
        t0 = v.TimeLastUpdate
        assert v.Data.dtype == np.float32 and v.Data.flags.f_contiguous  # set.Data -> single (Volume.m:89)
        t[0] += 10
        v.Data[1, :, :] = 3.0  # indexed assignment fires the PostSet listener (Volume.m:225-238)
        assert v.TimeLastUpdate > t0
        t1 = v.TimeLastUpdate
        t[0] += 10
        v.Data = np.ones((2, 2, 2))
        assert v.TimeLastUpdate > t1
    finally:
        api._ts_clock = None


def test_timestamp_is_strictly_increasing():
    a = [api.timestamp() for _ in range(100)]
    assert all(y > x for x, y in zip(a, a[1:]))


def test_renderer_property_assignment_stamps_volume():
    r = vrt.OracleVolumeRender()
    v = api.Volume(np.zeros((2, 2, 2)))
    t0 = v.TimeLastUpdate
    r.VolumeEmission = v  # VolumeRender.m:723-740
    assert v.TimeLastUpdate > t0
    with pytest.raises(TypeError):
        r.VolumeAbsorption = False  # VolumeRender.m:491
    with pytest.raises(ValueError):
        r.ImageResolution = [1, 2, 3]
    with pytest.raises(ValueError):
        api.LightSource([1, 2], [1, 1, 1])


def test_rotate_matches_matlab_composition():
    r = vrt.OracleVolumeRender()
    r.rotate(90, 0, 0)
    assert np.allclose(r.RotationMatrix, [[1, 0, 0], [0, 0, -1], [0, 1, 0]], atol=1e-12)
    r.rotate(0, 0, 90)
    assert np.allclose(r.RotationMatrix @ r.RotationMatrix.T, np.eye(3), atol=1e-12)


def test_stereo_delta_example3():
    r = vrt.OracleVolumeRender()
    r.FocalLength = 4.5
    r.CameraXOffset = 0.06
    r.ImageResolution = [2048, 2048]
    base, delta = r.stereo_delta()
    assert base == 0.03 and delta == 31  # round(0.03 * 2048 / 2), SURVEY 8(d) config 4


def test_grad_follows_matlab_gradient_order():
    a = np.zeros((5, 6, 7), dtype=np.float32)
    a += np.arange(5, dtype=np.float32)[:, None, None] * 2      # slope 2 along dim 1
    a += np.arange(6, dtype=np.float32)[None, :, None] * 3      # slope 3 along dim 2
    a += np.arange(7, dtype=np.float32)[None, None, :] * 5      # slope 5 along dim 3
    gx, gy, gz = api.Volume(a).grad()
    assert np.allclose(gx.Data, 3) and np.allclose(gy.Data, 2) and np.allclose(gz.Data, 5)


def test_normalize_image_and_sequence():
    img = np.random.default_rng(0).random((4, 5, 3))
    n = api.VolumeRender.normalizeImage(img)
    assert np.isclose(n.max(), 1.0)
    seq = np.stack([img, 2 * img], axis=3)
    ns = api.VolumeRender.normalizeSequence(seq)
    assert np.isclose(ns.max(), 1.0) and np.isclose(ns[..., 0].max(), 0.5)
    with pytest.raises(ValueError):
        api.VolumeRender.normalizeSequence(img)


def test_multi_pass_render_then_combine_is_image_addition():
    """examples/example1.m:60-85: two passes with different factors/colours, added in MATLAB."""
    vol = vrt.synth_brain((24, 24, 24))
    r = vrt.example1_scene(vrt.OracleVolumeRender, vol, (40, 40), lights=1)
    first = r.render()
    r.FactorEmission, r.FactorAbsorption, r.FactorReflection, r.Color = 0.1, 0.4, 0.1, [1, 1, 1]
    second = r.render()
    both = first + second
    assert both.shape == first.shape and (both >= first).all()


# ---- image-space partition (host twin of the kernel's tile ownership) --------------------------------
@pytest.mark.parametrize("W,H,tile,n", [(192, 160, 32, 2), (2048, 2048, 64, 8), (100, 70, 16, 3), (33, 17, 8, 4)])
def test_owned_masks_partition_the_image(W, H, tile, n):
    masks = [multigpu.owned_mask(H, W, tile, tile, r, n) for r in range(n)]
    total = np.zeros((H, W), dtype=np.int32)
    for m in masks:
        total += m
    assert (total == 1).all()
    counts = [int(m.sum()) for m in masks]
    if W * H >= 2048 * 2048:
        assert max(counts) - min(counts) <= tile * tile


def test_div_pi_is_exact():
    """raymarch_core.h: div_pi(x) must equal the IEEE quotient x / pi on the whole range of angles
    (0 and [3e-4, pi]); checked here on every 61st float of (1e-30, 4] plus 0 (the full exhaustive
    sweep, 1.08e9 values, was run once during development: 0 mismatches)."""
    import ctypes as C
    import struct

    import vrt

    lib = vrt.host_emul()
    lib.emul_div_pi_mismatches.restype = C.c_uint64
    lib.emul_div_pi_mismatches.argtypes = [C.c_uint32] * 3
    lo = struct.unpack("<I", struct.pack("<f", 1e-30))[0] + 1
    hi = struct.unpack("<I", struct.pack("<f", 4.0))[0]
    assert lib.emul_div_pi_mismatches(lo, hi, 61) == 0
    assert lib.emul_div_pi_mismatches(0, 0, 1) == 0


def test_acos_polynomial_accuracy_and_nan_rule():
    """raymarch_core.h: acos_poly (the device's acos) stays within 5e-7 (two ulps of pi) of acos on [-1,1], is
    exact at +-1 and is NaN exactly where acosf is (|x| > 1, NaN)."""
    import ctypes as C

    import numpy as np
    import vrt

    lib = vrt.host_emul()
    lib.emul_acos_poly.restype = C.c_float
    lib.emul_acos_poly.argtypes = [C.c_float]
    xs = np.concatenate([np.linspace(-1, 1, 20001), 1 - np.logspace(-8, -1, 200), -1 + np.logspace(-8, -1, 200)]).astype(np.float32)
    got = np.array([lib.emul_acos_poly(float(x)) for x in xs], dtype=np.float64)
    assert np.abs(got - np.arccos(xs.astype(np.float64))).max() < 5e-7
    assert lib.emul_acos_poly(1.0) == 0.0 and lib.emul_acos_poly(-1.0) == np.float32(np.pi)
    for bad in (1.0000001, -1.0000001, 2.0, float("nan"), float("inf")):
        assert np.isnan(lib.emul_acos_poly(bad))


def test_python_constants_and_structs_match_the_header(tmp_path):
    """volume_renderer_b200/_abi.py restates include/vr_abi.h by hand: every VR_OPT_* value, and vr_stats field by field
    (size and offsets from a C program compiled against the header itself)."""
    import ctypes
    import re
    import subprocess

    from volume_renderer_b200 import _abi

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    hdr = open(os.path.join(root, "include", "vr_abi.h")).read()
    opts = dict(re.findall(r"\bVR_OPT_([A-Z0-9_]+)\s*=\s*(\d+)", hdr))
    assert len(opts) >= 20
    for name, value in opts.items():
        assert getattr(_abi, "OPT_" + name) == int(value), name
    fields = [f for f, _ in _abi.VrStats._fields_]
    src = tmp_path / "layout.c"
    src.write_text('#include <stddef.h>\n#include <stdio.h>\n#include "vr_abi.h"\nint main(void) {\n  printf("%zu\\n", sizeof(vr_stats));\n'
                   + "".join('  printf("%%zu\\n", offsetof(vr_stats, %s));\n' % f for f in fields) + "  return 0;\n}\n")
    exe = tmp_path / "layout"
    subprocess.run(["gcc", "-std=c11", "-I", os.path.join(root, "include"), "-o", str(exe), str(src)], check=True)
    out = [int(x) for x in subprocess.run([str(exe)], check=True, stdout=subprocess.PIPE, text=True).stdout.split()]
    assert out[0] == ctypes.sizeof(_abi.VrStats)
    assert out[1:] == [getattr(_abi.VrStats, f).offset for f in fields]


def test_slice_starts_are_the_serial_recurrences_bit_for_bit():
    """Depth slices: the exact fast-forward that finds where a slice starts (raymarch_core.h: fast_forward_to,
    run here on the host) lands on the very float state the reference's `t += tstep; pos += step` loop reaches, on 20 000 random
    rays with coordinates that cross zero, tiny and zero direction components and ends before / behind the ray's end; and
    float_below - which turns a slice's end into the ray's tfar - is nextafter(x, -inf)."""
    import ctypes as C

    import vrt

    lib = vrt.host_emul()
    lib.emul_fast_forward_mismatches.restype = C.c_uint64
    lib.emul_fast_forward_mismatches.argtypes = [C.c_uint32] * 2
    assert lib.emul_fast_forward_mismatches(vrt.SEED, 20000) == 0
    lib.emul_float_below_mismatches.restype = C.c_uint64
    lib.emul_float_below_mismatches.argtypes = [C.c_uint32] * 2
    assert lib.emul_float_below_mismatches(vrt.SEED, 200000) == 0
