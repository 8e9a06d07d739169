"""Run by tests/test_gpu_tc.py::test_tc_chained_plan_equals_exact_engine in a process of its own: TM_TC_CHAIN is
read once per process, and =2 (2 MB of operand per split) makes a 131 072-row dataset take the chained, split-major
plan that a 1 048 576-row tileset takes by default.  Compares the tensor-core engine with the exact engine."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import tiler_b200 as tb
    k = int(sys.argv[1])
    rng = np.random.default_rng(100 + k)
    n, nq = 131072, 76800  # 512 tiles -> 8 splits of 64; 300 query blocks >= 4 x 74 pairs
    ds = rng.standard_normal((n, 192)).astype(np.float32)
    ds[1::7] = ds[0:-1:7][: len(ds[1::7])] + (1e-3 * rng.standard_normal((len(ds[1::7]), 192))).astype(np.float32)  # near-duplicate pairs
    q = ds[rng.integers(0, n, nq)] + (0.05 * rng.standard_normal((nq, 192))).astype(np.float32)
    q[:4096] = rng.standard_normal((4096, 192)).astype(np.float32)  # no close row at all: loose thresholds, full lists
    with tb.KDTree.from_rows(ds) as t:
        tb.set_engine(tb.ENGINE_SIMT)
        si, se = t.search_batch(q, k=k)
        tb.set_engine(tb.ENGINE_TC)
        tb.reset_stats()
        ti, te = t.search_batch(q, k=k)
        st = tb.stats()
    print(json.dumps({"idx_equal": bool(np.array_equal(si, ti)),
                      "err_equal": bool(np.array_equal(se.view(np.uint32), te.view(np.uint32))),
                      "coarse_launches": int(st["coarse_launches"]), "fallback_queries": int(st["fallback_queries"]), "nq": nq}))


if __name__ == "__main__":
    main()
