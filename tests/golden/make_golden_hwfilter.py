"""Fixture of REAL texture-unit fetches: tools/hwfilter_probe (run on a B200: `gpurun -- tools/hwfilter_probe gpurun_out/hwfilter`)
dumps 1 M linear-filtered tex3D<float> fetches of a 16 x 12 x 20 fp32 array with the reference's texture settings
(volumeRender_kernel.cu:544-548).  This script keeps the array, 16384 of the random coordinates (with everything outside
[0,1] that exercises the clamp) and the hardware's answers as tests/golden/hwfilter_b200.npz.
    python tests/golden/make_golden_hwfilter.py gpurun_out/hwfilter"""
import os
import sys

import numpy as np

prefix = sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/hwfilter"
c = np.fromfile(prefix + "_coords.f32", np.float32).reshape(-1, 3)
hw = np.fromfile(prefix + "_rand.f32", np.float32)
vol = np.fromfile(prefix + "_vol.f32", np.float32).reshape(20, 12, 16)  # z, y, x
m = len(c)
sel = np.arange(m // 2, m // 2 + 16384)  # the random part of the probe
out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "hwfilter_b200.npz")
np.savez_compressed(out, vol=vol, coords=c[sel], hw=hw[sel])
print(out, os.path.getsize(out))
