#!/usr/bin/env python
"""Times the HBM-bound helper kernels and the KModes assignment pass on their BASELINE configs
(CUDA events, inputs resident in HBM) and prints one JSON line per kernel:
  pack     psyv_tileset_kernel   C3 tileset: 65 536 tiles x 4 orientations  (784 B algorithmic per row)
  queries  psyv_queries_kernel   2 097 152 query tiles                       (1 024 B per query)
  kmodes   kmodes_assign_kernel  C4: 1 048 576 points x 16 384 centroids      (92 B per point; pairs/s)
Also the command ncu is pointed at for the per-kernel captures under profiles/.
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--only", default="pack,queries,kmodes")
    ap.add_argument("--points", type=int, default=1 << 20)
    ap.add_argument("--centroids", type=int, default=1 << 14)
    args = ap.parse_args()
    import torch
    import tiler_b200 as tb
    from bench import synth_tileset, synth_queries
    dev = torch.device("cuda", 0)
    tb.set_device(0)
    stream = torch.cuda.current_stream()
    tb.set_stream(stream.cuda_stream)
    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
    hbm = peaks.get("hbm_gbs", 6500.0)
    gen = torch.Generator(device=dev).manual_seed(0x42381337)
    only = set(args.only.split(","))

    def timed(fn):
        fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        best = 1e30
        for _ in range(args.reps):
            e0.record(stream)
            fn()
            e1.record(stream)
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        return best

    tiles, pal = synth_tileset(torch, 65536, 16, gen)
    if "pack" in only:
        from tiler_b200._lib import lib
        from tiler_b200.ann import _dp
        out = torch.empty((65536 * 4, 192), dtype=torch.float32, device=dev)
        # the pack kernel alone (dataset creation also builds the fp16 operand): time the whole create, report both
        def create():
            t = tb.KDTree.from_tileset(tiles, pal.view(1, -1))
            t.destroy()
        ms = timed(create)
        nrows = 65536 * 4
        print(json.dumps({"kernel": "dataset create (pack/mirror + fp16 operand + norm tiles)", "rows": nrows, "ms": ms}))
    if "queries" in only:
        nq = 1 << 21
        rgb = synth_queries(torch, nq, tiles[:4096], pal, gen)
        out = torch.empty((nq, 192), dtype=torch.float32, device=dev)
        ms = timed(lambda: tb.psyv_queries(rgb, out=out))
        b = nq * 1024
        print(json.dumps({"kernel": "psyv_queries_kernel", "queries": nq, "ms": ms, "algorithmic_bytes": b,
                          "achieved_gbs": b / ms / 1e6, "hbm_peak_gbs": hbm, "frac": b / ms / 1e6 / hbm}))
    if "kmodes" in only:
        npnt, k = args.points, args.centroids
        t2, _ = synth_tileset(torch, npnt, 16, gen)
        # KModes rows: 64 palette indices + 16 zone flags (main.pas:4142-4183), built with torch for speed
        zone = (t2.long() * 16) // 16
        flags = torch.stack([(zone == z).sum(1) > 1 for z in range(16)], 1).to(torch.uint8)
        X = torch.cat([t2, flags], 1).contiguous()
        cent = X[:: max(1, npnt // k)][:k].contiguous()
        clust = torch.empty(npnt, dtype=torch.int32, device=dev)
        dis = torch.empty(npnt, dtype=torch.int64, device=dev)
        ms = timed(lambda: tb.kmodes_assign(X, cent, tb.KM_ASM, out_clust=clust, out_dis=dis))
        b = npnt * 92 + k * 80
        pairs = npnt * k
        print(json.dumps({"kernel": "kmodes_assign_kernel", "points": npnt, "centroids": k, "ms": ms,
                          "pairs_per_s": pairs / ms * 1e3, "byte_compares_per_s": pairs * 80 / ms * 1e3,
                          "algorithmic_bytes": b, "achieved_gbs": b / ms / 1e6, "hbm_peak_gbs": hbm}))


if __name__ == "__main__":
    main()
