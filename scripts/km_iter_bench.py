#!/usr/bin/env python
"""One KModesIter over C4 (1 048 576 points x 16 384 centroids, 1 093 bins of 960, kmodes.pas:845-909) through the
device-resident handle: a few changed centroid rows before every bin, raw C-ABI calls, host wall clock.  Prints one
JSON line; TM_KM_POLL=0 selects the copy + synchronise completion for a same-box A/B."""
import ctypes as C
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import tiler_b200 as tb
    from tiler_b200._lib import lib
    rows_per_bin = int(sys.argv[1]) if len(sys.argv) > 1 else 4
    tb.set_device(0)
    rng = np.random.default_rng(0)
    npnt, k = 1 << 20, 16384
    X = rng.integers(0, 16, (npnt, 80), dtype=np.uint8)
    X[:, 64:] &= 1
    Cn = X[rng.choice(npnt, k, replace=False)].copy()
    vp = C.c_void_p
    with tb.KModesState(X) as st:
        st.set_centroids(Cn)
        st.assign_range(0, 960)
        cl, di = np.empty(960, np.int32), np.empty(960, np.uint64)
        nbins = (npnt + 959) // 960
        idx = rng.integers(0, k, (nbins, rows_per_bin)).astype(np.int32)
        rows = np.ascontiguousarray(Cn[idx])
        best = None
        sweeps = []
        for rep in range(3):
            t0 = time.perf_counter()
            nb = 0
            for lo in range(0, npnt, 960):
                lib.tm_kmodes_update_centroids(st._h, vp(idx[nb].ctypes.data), vp(rows[nb].ctypes.data), rows_per_bin)
                lib.tm_kmodes_assign_range(st._h, lo, min(npnt, lo + 960), vp(cl.ctypes.data), vp(di.ctypes.data))
                nb += 1
            dt = time.perf_counter() - t0
            sweeps.append(round(1e6 * dt / nb, 1))
            best = dt if best is None else min(best, dt)
    print(json.dumps({"case": "kmodes_iter_c4_handle", "poll": os.environ.get("TM_KM_POLL", "1"), "rows_per_bin": rows_per_bin,
                      "bins": nb, "ms": 1e3 * best, "us_per_bin": 1e6 * best / nb,
                      "us_per_bin_by_sweep": sweeps, "note": "sweep 1 has no labels from a previous pass to bound the search"}))


if __name__ == "__main__":
    main()
