#!/usr/bin/env python
"""Latency of the per-frame drop-in call: tm_frametiling_search with HOST buffers, one
frame per call, from 1 and from 8 caller threads (the reference runs DoFrameTiling one frame per worker thread)."""
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import numpy as np
    import torch
    import tiler_b200 as tb
    from bench import synth_tileset, synth_queries
    dev = torch.device("cuda", 0)
    tb.set_device(0)
    gen = torch.Generator(device=dev).manual_seed(0x42381337 + 9)
    for name, pos, ntiles in (("720p frame, 16k tiles", 160 * 90, 16384), ("1080p frame, 64k tiles", 240 * 135, 65536)):
        tiles, pal = synth_tileset(torch, ntiles, 16, gen)
        tree = tb.KDTree.from_tileset(tiles, pal.view(1, -1))
        nfr = 64
        rgb = synth_queries(torch, pos * nfr, tiles, pal, gen).cpu().numpy().reshape(nfr, pos, 64)
        out_i = np.empty((nfr, pos), np.int32)
        out_e = np.empty((nfr, pos), np.float32)
        for f in range(4):
            tree.frametiling_search(rgb[f], out_idx=out_i[f], out_err=out_e[f])
        t0 = time.perf_counter()
        for f in range(nfr):
            tree.frametiling_search(rgb[f], out_idx=out_i[f], out_err=out_e[f])
        t1 = time.perf_counter()
        serial = (t1 - t0) / nfr
        tb.reset_stats()
        tb.set_profiling(True)
        for f in range(nfr):
            tree.frametiling_search(rgb[f], out_idx=out_i[f], out_err=out_e[f])
        st = tb.stats()
        tb.set_profiling(False)
        dev_ms = {k: st[k] / nfr for k in ("desc_ms", "search_ms", "coarse_ms", "rerank_ms")}

        bar = threading.Barrier(9)

        def worker(w):
            tree.frametiling_search(rgb[w], out_idx=out_i[w], out_err=out_e[w])  # per-thread staging exists after this
            bar.wait()
            for f in range(w, nfr, 8):
                tree.frametiling_search(rgb[f], out_idx=out_i[f], out_err=out_e[f])
            bar.wait()
        th = [threading.Thread(target=worker, args=(w,)) for w in range(8)]
        for t in th:
            t.start()
        bar.wait()
        t0 = time.perf_counter()
        bar.wait()
        t1 = time.perf_counter()
        for t in th:
            t.join()
        print(json.dumps({"case": name, "queries_per_call": pos, "rows": tree.n, "ms_per_call_1_thread": serial * 1e3,
                          "queries_per_s_1_thread": pos / serial, "device_ms_per_call": dev_ms,
                          "ms_per_call_8_threads": (t1 - t0) / nfr * 1e3, "queries_per_s_8_threads": pos * nfr / (t1 - t0)}))
        tree.destroy()


def kmodes_bin_case():
    """KModesIter's unit of work (kmodes.pas:866-878): one bin of cBinSize = 960 points against all centroids,
    host buffers, centroids re-uploaded every call because the Pascal side changes them between bins."""
    import numpy as np
    import tiler_b200 as tb
    rng = np.random.default_rng(3)
    X = rng.integers(0, 16, (960, 80), dtype=np.uint8)
    X[:, 64:] = rng.integers(0, 2, (960, 16), dtype=np.uint8)
    for k in (4096, 16384):
        C = rng.integers(0, 16, (k, 80), dtype=np.uint8)
        C[:, 64:] = rng.integers(0, 2, (k, 16), dtype=np.uint8)
        for _ in range(3):
            tb.kmodes_assign(X, C)
        t0 = time.perf_counter()
        for _ in range(50):
            tb.kmodes_assign(X, C)
        ms = (time.perf_counter() - t0) / 50 * 1e3
        print(json.dumps({"case": "kmodes bin (960 points)", "centroids": k, "ms_per_call": ms,
                          "pairs_per_s": 960 * k / ms * 1e3}))


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "kmodes":
        kmodes_bin_case()
        sys.exit(0)
    main()
