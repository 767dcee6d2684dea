"""Generates the golden fixtures in this directory from the reference's OWN CPU sources, compiled
unmodified from /root/reference into oracle/_ref/libref_cpu.so (oracle/Makefile target `ref`).
Run here (the container with /root/reference); the fixtures travel to the GPU box.

    python tests/golden/make_golden.py
"""
import ctypes as C
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import vrt  # noqa: E402

ref = vrt.ref_cpu()
assert ref is not None, "needs /root/reference"

# 1. HenyeyGreenstein(64) (HenyeyGreenstein.cc, default g = 0.8)
lut = np.zeros(64 ** 3, dtype=np.float32)
err = C.create_string_buffer(256)
assert ref.ref_henyey_greenstein(64, 0, 0.0, lut.ctypes.data, err, 256) == 0
np.save(os.path.join(HERE, "hg_lut_64_g0.8.npy"), lut.reshape((64, 64, 64), order="F"))

# 2. vr::initRender (volumeRender.cpp:112-156): boxmax and tstep for a set of extents / element sizes
cases = [((64, 64, 64), (1, 1, 1)), ((1024, 1024, 1024), (1, 1, 1)), ((4096, 4096, 4096), (1, 1, 1)), ((300, 200, 100), (0.5, 0.25, 2.0)),
         ((7, 130, 31), (1.0, 3.0, 0.1)), ((512, 512, 170), (0.32, 0.32, 1.0)), ((40, 24, 1), (1, 1, 1))]
dims, ess, bmaxs, tsteps = [], [], [], []
for d, es in cases:
    d3 = (C.c_uint64 * 3)(*d)
    esx = (C.c_float * 3)(*es)
    ext, mem = (C.c_uint64 * 3)(), C.c_uint64()
    rmin, rmax, rt, fo = (C.c_float * 3)(), (C.c_float * 3)(), C.c_float(), (C.c_float * 3)()
    factors = (C.c_float * 3)(1, 1, 1)
    data = np.zeros(1, dtype=np.float32)
    ref.ref_init_render(data.ctypes.data, d3, 3 if d[2] > 1 else 2, 1, esx, 64, 64, factors, 0.9, ext, C.byref(mem), rmin, rmax, C.byref(rt), fo)
    dims.append(d), ess.append(es), bmaxs.append(list(rmax)), tsteps.append(rt.value)
np.savez(os.path.join(HERE, "init_render.npz"), dims=np.array(dims, dtype=np.uint64), es=np.array(ess, dtype=np.float32),
         boxmax=np.array(bmaxs, dtype=np.float32), tstep=np.array(tsteps, dtype=np.float32))
print("golden fixtures written to", HERE)
