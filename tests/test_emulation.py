"""The product's per-lane device code (raymarch_core.h is __host__ __device__) run on the CPU,
lane-serially, against the oracle: proves the kernel arithmetic - padded/bricked addressing,
float clamp instead of integer clamps, exact empty-space skipping, template gradient variants -
before any GPU time is spent.  Both sides are built without FMA contraction, so images must be
bit-identical; the CUDA build (FMA on) is held to the 1e-3 tolerance in test_gpu_parity.py."""
import numpy as np
import pytest

import vrt
from volume_renderer_b200 import api


@pytest.fixture(scope="module")
def brain():
    return vrt.synth_brain((40, 40, 40))


@pytest.mark.parametrize("layout", [0, 1])
@pytest.mark.parametrize("skip", [0, 1, 2])
@pytest.mark.parametrize("lights", [0, 1, 2])
def test_on_the_fly_gradient(brain, layout, skip, lights):
    r = vrt.example1_scene(vrt.OracleVolumeRender, brain, (80, 64), lights=max(lights, 1))
    if lights == 0:
        r.LightSources = False
    ref = r.render()
    img, st = vrt.emul_render(r, layout=layout, skip=skip)
    assert np.array_equal(img, ref)
    o = r._backend.stats(r.objectHandle)
    assert st[0] == o.samples_marched and st[2] == o.rays_hit
    assert st[1] < st[0]  # most marched samples are empty space


@pytest.mark.parametrize("layout", [0, 1])
def test_gradient_lookup(brain, layout):
    r = vrt.example1_scene(vrt.OracleVolumeRender, brain, (64, 64), lights=2)
    gx, gy, gz = r.VolumeEmission.grad()
    r.VolumeGradientX, r.VolumeGradientY, r.VolumeGradientZ = gx, gy, gz
    ref = r.render()
    o = r._backend.stats(r.objectHandle)
    for skip, variant in ((1, 0), (2, 0), (2, 1)):
        img, st = vrt.emul_render(r, layout=layout, skip=skip, variant=variant)
        assert np.array_equal(img, ref)
        assert st[0] == o.samples_marched


def test_anisotropic_and_frac_bits():
    vol = vrt.synth_brain((50, 34, 22))
    r = vrt.example1_scene(vrt.OracleVolumeRender, vol, (72, 56), lights=1)
    r.ElementSizeUm = [1.0, 0.8, 2.5]
    r._backend.lib.orc_set_option(r.objectHandle, vrt.ORC_OPT_FILTER_FRAC_BITS, 8)
    ref = r.render()
    img, _ = vrt.emul_render(r, layout=1, skip=1, frac_bits=8)
    assert np.array_equal(img, ref)


def test_skipping_with_tubes_and_inside_camera():
    vol = vrt.synth_tubes((64, 64, 64), ntubes=200)
    vol[vol < 0.999] = 0  # thin the union out so that there is empty space to skip
    r = vrt.example1_scene(vrt.OracleVolumeRender, np.asfortranarray(vol), (64, 64), lights=1)
    r.DistanceToObject = 0.7  # camera inside the box
    r.OpacityThreshold = 0.5
    r.FactorAbsorption = 30.0  # early ray termination does happen
    ref = r.render()
    for skip in (0, 1, 2):
        img, st = vrt.emul_render(r, layout=1, skip=skip)
        assert np.array_equal(img, ref)
        assert st[0] == r._backend.stats(r.objectHandle).samples_marched


def test_constant_reflection_volume_and_2d_data():
    flat = np.asfortranarray(np.random.default_rng(1).random((30, 18), dtype=np.float32))
    r = vrt.example1_scene(vrt.OracleVolumeRender, flat, (48, 40), lights=1)
    r.VolumeReflection = api.Volume(0.25)
    ref = r.render()
    img, _ = vrt.emul_render(r, layout=1, skip=1)
    assert np.array_equal(img, ref)


def test_stereo_offset_passes():
    vol = vrt.synth_brain((32, 32, 32))
    r = vrt.example1_scene(vrt.OracleVolumeRender, vol, (64, 48), lights=1)
    r.CameraXOffset = 0.06
    base, delta = r.stereo_delta()
    res = [48, 64 + delta]
    for off in (base, -base):
        ref = r.p_render(off, res)
        img, _ = vrt.emul_render(r, layout=1, skip=1, resolution=res, xoffset=off)
        assert np.array_equal(img, ref)


def test_closed_form_jumps_axis_parallel_and_zero_crossings():
    """The exact fast-forward must survive rays whose coordinates sit at 0 / cross 0 (binade
    borders) and rays parallel to an axis: no rotation, centre pixel looks straight down z."""
    vol = np.zeros((48, 40, 56), dtype=np.float32, order="F")
    vol[20:28, 16:24, 40:52] = vrt.synth_brain((8, 8, 12)) + np.float32(0.2)
    vol[2:5, 30:36, 3:9] = 0.7
    for dist, rot in ((6.0, (0, 0, 0)), (2.5, (90, 0, 0)), (6.0, (45, 45, 45)), (0.3, (10, 200, 5))):
        r = vrt.example1_scene(vrt.OracleVolumeRender, vol, (65, 49), lights=1)
        r.RotationMatrix = np.eye(3)
        r.rotate(*rot)
        r.DistanceToObject = dist
        r.ElementSizeUm = [1.0, 1.3, 0.7]
        ref = r.render()
        o = r._backend.stats(r.objectHandle)
        for skip in (1, 2):
            img, st = vrt.emul_render(r, layout=1, skip=skip)
            assert np.array_equal(img, ref), (dist, rot, skip)
            assert st[0] == o.samples_marched and st[2] == o.rays_hit, (dist, rot, skip, st, o.samples_marched)


@pytest.mark.parametrize("splits", [(2, 1, 1), (1, 1, 3), (2, 2, 2)])
@pytest.mark.parametrize("grad", [0, 1])
def test_volume_space_bricks_match_one_pass(splits, grad):
    """Sort-last mode on virtual shards: per-brick opacity pass, shading pass that starts from the
    opacity of the bricks in front, sum of the shares.  Against the oracle's single pass it may differ
    only by re-association of the opacity product (well inside the 1e-3 bar) - including across the
    early-termination threshold, which is crossed inside some brick."""
    from volume_renderer_b200 import multigpu

    vol = vrt.synth_brain((40, 36, 44))
    r = vrt.example1_scene(vrt.OracleVolumeRender, vol, (72, 64), lights=2)
    r.FactorAbsorption = 40.0  # rays do reach the 0.9 threshold inside the volume
    if grad:
        gx, gy, gz = r.VolumeEmission.grad()
        r.VolumeGradientX, r.VolumeGradientY, r.VolumeGradientZ = gx, gy, gz
    ref = r.render()
    bricks = multigpu.brick_grid(vol.shape, splits)
    img, _ = vrt.emul_render(r, layout=1, skip=2, bricks=bricks)
    m = vrt.compare(img, ref)
    assert m["max_abs_norm"] < 2e-5, m
    one, _ = vrt.emul_render(r, layout=1, skip=2, bricks=multigpu.brick_grid(vol.shape, (1, 1, 1)))
    assert np.array_equal(one, ref)  # a single brick is the plain march


@pytest.mark.parametrize("splits,slices", [((1, 1, 1), 2), ((1, 1, 2), 3), ((2, 2, 2), 4), ((1, 2, 1), 16)])
@pytest.mark.parametrize("grad", [0, 1])
def test_depth_slices_visit_every_sample_once(splits, slices, grad):
    """Depth slices of the volume-space passes with the product's own per-lane code on the host: slice starts by
    the exact fast-forward, a slice's end as the ray's tfar, opacities folded as a product, shares summed.  Every lattice sample
    is visited and shaded by exactly one slice (equal counters), the image differs from whole rays by the rounding of the folds,
    also where rays end at the threshold inside a slice."""
    from volume_renderer_b200 import multigpu

    vol = vrt.synth_brain((40, 36, 44))
    bricks = multigpu.brick_grid(vol.shape, splits)
    for absorption, bar in ((4.0, 2e-6), (40.0, 2e-5)):  # translucent / rays that reach the 0.9 threshold
        r = vrt.example1_scene(vrt.OracleVolumeRender, vol, (72, 64), lights=2)
        r.FactorAbsorption = absorption
        if grad:
            gx, gy, gz = r.VolumeEmission.grad()
            r.VolumeGradientX, r.VolumeGradientY, r.VolumeGradientZ = gx, gy, gz
        whole, st1 = vrt.emul_render(r, layout=1, skip=2, bricks=bricks)
        cut, stS = vrt.emul_render(r, layout=1, skip=2, bricks=bricks, slices=slices)
        m = vrt.compare(cut, whole)
        assert m["max_abs_norm"] < bar, (absorption, m)
        if absorption == 4.0:
            assert stS[0] == st1[0] and stS[1] == st1[1] and st1[1] > 0, (st1, stS)  # samples visited, samples shaded
        assert vrt.compare(cut, r.render())["max_abs_norm"] < 2e-5


def test_brick_grid_and_splits():
    from volume_renderer_b200 import multigpu

    assert multigpu.splits_for(8) == (2, 2, 2) and multigpu.splits_for(2) == (1, 1, 2) and multigpu.splits_for(4) == (1, 2, 2)
    assert multigpu.splits_for(6) == (1, 1, 6) or np.prod(multigpu.splits_for(6)) == 6
    b = multigpu.brick_grid((10, 7, 9), (2, 2, 2))
    assert len(b) == 8 and sum(np.prod(np.array(hi) - np.array(lo)) for lo, hi in b) == 10 * 7 * 9
