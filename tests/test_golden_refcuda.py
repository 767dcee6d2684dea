"""The pin of the oracle: images rendered on a B200 by the reference's OWN CUDA kernel and upload code
(oracle/ref_cuda: src/C/vr/volumeRender_kernel.cu compiled through a texture-object shim, see
tests/golden/make_golden_refcuda.py) are committed under tests/golden/refcuda_*.npy.

The hardware path differs from "the render math" in ways no CPU restatement can reproduce bit for bit:
the texture unit filters with 1.8 fixed-point weights and its own internal arithmetic, nvcc contracts
the reference's arithmetic to FMAs at its own discretion, rsqrtf/__expf are approximations - and the
illumination has a discontinuity (an acos argument that rounds outside [-1,1] is NaN -> LUT coordinate
0 instead of 1), so single samples flip sides.  Measured: median deviation 1e-4 of the image maximum,
98-99.9 % of the pixels within 1e-3, isolated pixels up to 1.5e-2, PSNR 74-85 dB; the scene without a
discontinuity (constant volume, NaN normals everywhere: the NaN -> 0 coordinate rule itself) agrees to
6e-6.  The bars below hold the oracle - with the texture unit's 8-bit weights and with fp32 weights -
to that envelope."""
import os

import numpy as np
import pytest

import vrt

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
# stereo_pass: pixels on the illumination's NaN discontinuity that may flip against the reference kernel's image
# (measured: one, 2.5e-2 of the image maximum); everything else of that scene is held to the tight bar
STEREO_DISCONTINUITY_PIXELS = 2


def envelope(img, gold):
    img, gold = np.asarray(img, np.float64), np.asarray(gold, np.float64)
    assert img.shape == gold.shape
    peak = gold.max()
    d = np.abs(img - gold).max(axis=2) / peak
    mse = np.mean(((img - gold) / peak) ** 2)
    return {"max": float(d.max()), "frac_within_1e-3": float((d <= 1e-3).mean()), "median": float(np.median(d)),
            "psnr_db": float("inf") if mse == 0 else float(10 * np.log10(1.0 / mse))}


def check_envelope(e, name):
    assert np.isfinite(e["max"]), (name, e)
    if name == "constant_volume_nan_normal":
        assert e["max"] < 5e-5, (name, e)  # no discontinuity in play: the NaN-coordinate rule itself
        return
    assert e["psnr_db"] >= 68.0, (name, e)
    assert e["frac_within_1e-3"] >= 0.97, (name, e)
    assert e["max"] <= 3e-2, (name, e)
    assert e["median"] <= 5e-4, (name, e)


@pytest.mark.parametrize("name", sorted(vrt.golden_scenes()))
@pytest.mark.parametrize("frac_bits", [9, 8, 0])
def test_oracle_against_the_reference_kernel_on_b200(name, frac_bits):
    gold = np.load(os.path.join(GOLD, "refcuda_%s.npy" % name))
    r = vrt.golden_scenes()[name](vrt.OracleVolumeRender)
    r._backend.lib.orc_set_option(r.objectHandle, vrt.ORC_OPT_FILTER_FRAC_BITS, frac_bits)
    e = envelope(r.render(), gold)
    print(name, frac_bits, e)
    check_envelope(e, name)
    if frac_bits == 9 and name != "stereo_pass":
        # with the filter as measured on the texture unit the checker is 20 dB closer to the reference's own output
        # (one pixel of the stereo scene sits on the illumination's NaN discontinuity: 2.5e-2)
        assert e["psnr_db"] >= 95.0 and e["max"] <= 2e-3 and e["frac_within_1e-3"] >= 0.9995, (name, e)


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(vrt.golden_scenes()))
def test_product_against_the_reference_kernel_on_b200(name):
    from volume_renderer_b200 import _abi, api

    gold = np.load(os.path.join(GOLD, "refcuda_%s.npy" % name))
    r = vrt.golden_scenes()[name](api.VolumeRender)
    r.setOption(_abi.OPT_TEXTURE_UNIT, 0)  # the load path with the documented 8-bit formula
    r.setOption(_abi.OPT_FILTER_FRAC_BITS, 8)
    r.setOption(_abi.OPT_ARITHMETIC, _abi.ARITH_EXACT)
    e = envelope(r.render(), gold)
    print(name, e)
    check_envelope(e, name)


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(vrt.golden_scenes()))
@pytest.mark.parametrize("arithmetic", [2, 1, 0])
def test_texture_unit_mode_against_the_reference_kernel_on_b200(name, arithmetic):
    """VR_OPT_TEXTURE_UNIT (the default fetch): every fetch through the texture unit, configured like the reference's
    textures - the same hardware filter the golden images were rendered with - in each arithmetic flavour
    (VR_OPT_ARITHMETIC), held to what was MEASURED against the reference kernel's own images (profiles/r02_arith_ab.md):
      * flavours 2 (the default) and 1 - nvcc's FMA contraction and rsqrtf, as the reference's own source compiles:
        PSNR >= 110 dB, maximum <= 5e-4 of the image maximum, every pixel within 1e-3 (measured 112-125 dB, 7.8e-5 ... 3.2e-4);
      * flavour 0 - individually rounded IEEE operations, the CPU checker's arithmetic: PSNR >= 95 dB, maximum <= 2e-3,
        >= 99.95 % of the pixels within 1e-3 - the bar the oracle with the measured filter is held to above - and ONE
        exception listed by scene: a single pixel of the stereo pass sits on the illumination's NaN discontinuity (an acos
        argument rounding to either side of -1), 2.5e-2."""
    from volume_renderer_b200 import _abi, api

    gold = np.load(os.path.join(GOLD, "refcuda_%s.npy" % name))
    r = vrt.golden_scenes()[name](api.VolumeRender)
    r.setOption(_abi.OPT_TEXTURE_UNIT, 1)
    r.setOption(_abi.OPT_ARITHMETIC, arithmetic)
    img = np.asarray(r.render(), np.float64)
    assert r.stats()["arithmetic"] == arithmetic and r.stats()["fetch_mode"] == 1
    e = envelope(img, gold)
    print(name, "arithmetic", arithmetic, e)
    check_envelope(e, name)
    d = np.abs(img - np.asarray(gold, np.float64)).max(axis=2) / gold.max()
    outliers = [(int(y), int(x), float(d[y, x])) for y, x in zip(*np.where(d > 2e-3))]
    print(name, "pixels beyond 2e-3:", outliers)
    if name == "constant_volume_nan_normal":
        return
    if arithmetic >= 1:
        assert e["psnr_db"] >= 110.0 and e["max"] <= 5e-4 and e["frac_within_1e-3"] == 1.0, (name, arithmetic, e)
    elif name == "stereo_pass":
        assert len(outliers) <= STEREO_DISCONTINUITY_PIXELS, outliers
        d_rest = d.copy()
        for y, x, _ in outliers:
            d_rest[y, x] = 0.0
        assert d_rest.max() <= 2e-3 and (d <= 1e-3).mean() >= 0.9995, e
    else:
        assert not outliers, outliers
        assert e["psnr_db"] >= 95.0 and e["max"] <= 2e-3 and e["frac_within_1e-3"] >= 0.9995, (name, e)


@pytest.mark.gpu
def test_shimmed_reference_still_reproduces_its_golden():
    """When oracle/_ref/libref_cuda.so travelled to the box: the reference kernel, run again in a fresh
    process, reproduces the committed image (same GPU generation, same toolchain)."""
    import subprocess
    import sys
    import tempfile

    if vrt.ref_cuda() is None:
        pytest.skip("oracle/_ref/libref_cuda.so not built")
    with tempfile.TemporaryDirectory() as d:
        subprocess.run([sys.executable, os.path.join(GOLD, "make_golden_refcuda.py"), d, "example1"], check=True)
        now = np.load(os.path.join(d, "refcuda_example1.npy"))
    gold = np.load(os.path.join(GOLD, "refcuda_example1.npy"))
    assert vrt.compare(now, gold)["max_abs_norm"] < 1e-5
