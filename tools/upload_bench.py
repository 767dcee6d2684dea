#!/usr/bin/env python
"""Developer tool: host -> resident-brick upload throughput of vr_sync_volumes.

  python tools/upload_bench.py [--size 1024] [--reps 3]

Times a dirty re-upload of one N^3 fp32 volume (the emission role) from pageable and from pinned host
memory, and a partial update of half the volume (vr_update_box), wall clock around the call plus a
synchronise.  Prints one JSON line.
"""
import argparse
import ctypes as C
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))


def main():
    p = argparse.ArgumentParser()
    p.add_argument("--size", type=int, default=1024)
    p.add_argument("--reps", type=int, default=3)
    a = p.parse_args()
    import torch

    from volume_renderer_b200 import _abi

    lib = _abi.lib()
    ctx = C.c_void_p()
    _abi.check(lib.vr_create(C.byref(ctx)))
    n = a.size
    out = {"size": n, "gib": n ** 3 * 4 / 2 ** 30}
    one = np.ones(1, dtype=np.float32)

    def sync(ptr, stamp, T):
        E = _abi.volume_struct((ptr, (n, n, n), _abi.DTYPE_F32, _abi.MEM_HOST), stamp)
        R = _abi.volume_struct((one.ctypes.data, (1, 1, 1), _abi.DTYPE_F32, _abi.MEM_HOST), 1)
        t0 = time.perf_counter()
        _abi.check(lib.vr_sync_volumes(ctx, C.c_uint64(T), C.byref(E), C.byref(R), C.byref(E), None, None, None), ctx)
        torch.cuda.synchronize()
        return time.perf_counter() - t0

    pageable = np.empty(n ** 3, dtype=np.float32)
    pageable[:] = 0.25
    pinned = torch.empty(n ** 3, dtype=torch.float32).pin_memory()
    pinned.fill_(0.25)
    for name, ptr in (("pageable", pageable.ctypes.data), ("pinned", pinned.data_ptr())):
        ts = []
        stamp, T = 10, 0
        for _ in range(a.reps + 1):
            ts.append(sync(ptr, stamp, T))
            T = stamp  # "last sync" = now; the next stamp is newer -> dirty again
            stamp += 10
        best = min(ts[1:])
        out[name + "_s"] = best
        out[name + "_GBps"] = n ** 3 * 4 / best / 1e9
    if hasattr(lib, "vr_update_box"):
        lo = (C.c_uint64 * 3)(0, 0, 0)
        hi = (C.c_uint64 * 3)(n, n, n // 2)
        E = _abi.volume_struct((pageable.ctypes.data, (n, n, n), _abi.DTYPE_F32, _abi.MEM_HOST), 1000)
        ts = []
        for _ in range(a.reps + 1):
            t0 = time.perf_counter()
            _abi.check(lib.vr_update_box(ctx, 0, C.byref(E), lo, hi), ctx)
            torch.cuda.synchronize()
            ts.append(time.perf_counter() - t0)
        out["half_box_pageable_s"] = min(ts[1:])
        out["half_box_pageable_GBps"] = n ** 3 * 2 / min(ts[1:]) / 1e9
    _abi.check(lib.vr_destroy(ctx))
    print(json.dumps(out))


if __name__ == "__main__":
    main()
