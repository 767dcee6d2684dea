"""A/B of the arithmetic flavours (VR_OPT_ARITHMETIC) on the GPU: each golden scene against the reference kernel's own image
(tests/golden/refcuda_*.npy) and against the CPU checker with the measured hardware filter.  Prints one line per scene and
flavour; bench timings are taken separately (bench.py --arith)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import vrt  # noqa: E402
from test_golden_refcuda import GOLD, envelope  # noqa: E402
from volume_renderer_b200 import _abi, api  # noqa: E402

for name, make in sorted(vrt.golden_scenes().items()):
    gold = np.load(os.path.join(GOLD, "refcuda_%s.npy" % name))
    orc = make(vrt.OracleVolumeRender)
    orc._backend.lib.orc_set_option(orc.objectHandle, vrt.ORC_OPT_FILTER_FRAC_BITS, 9)
    ref = orc.render()
    for arith in (0, 1, 2):
        r = make(api.VolumeRender)
        r.setOption(_abi.OPT_TEXTURE_UNIT, 1)
        r.setOption(_abi.OPT_ARITHMETIC, arith)
        img = r.render()
        e = envelope(img, gold)
        o = envelope(img, ref)
        d = np.abs(np.asarray(img, np.float64) - gold).max(axis=2) / gold.max()
        print("%-28s arith %d | vs reference kernel: psnr %6.1f max %.2e within1e-3 %.5f n>2e-3 %d | vs checker: psnr %6.1f max %.2e within1e-3 %.5f" % (
            name, arith, e["psnr_db"], e["max"], e["frac_within_1e-3"], int((d > 2e-3).sum()), o["psnr_db"], o["max"], o["frac_within_1e-3"]), flush=True)
