"""Runs the gather probes (csrc/probe_kernels.cu) on their own, for an ncu capture of their L1 / L2 hit rates:
    ncu --metrics <list> -k regex:k_probe --csv --log-file probe.csv python tools/probe_run.py
The hit rates say which probe footprint serves its taps from the level the ray march's own fetches come from."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from volume_renderer_b200 import _abi  # noqa: E402

lib = _abi.lib()
ctx = C.c_void_p()
_abi.check(lib.vr_create(C.byref(ctx)))
kinds = [int(k) for k in (sys.argv[1].split(",") if len(sys.argv) > 1 else ["2", "0", "4"])]  # 4 = TMA-staged neighbourhood, taps from shared memory
for kind in kinds:
    for e in (5, 8, 10):
        f, ms = C.c_double(), C.c_double()
        _abi.check(lib.vr_probe_gather(ctx, kind, e, C.byref(f), C.byref(ms)), ctx)
        print("kind %d footprint (2^%d)^3: %.1f G trilinear fetches/s, %.3f ms" % (kind, e, f.value / 1e9, ms.value))
_abi.check(lib.vr_destroy(ctx))
