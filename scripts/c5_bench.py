#!/usr/bin/env python
"""Times the BASELINE C5-shaped search on one GPU's shard (k = 8 neighbours + FP32 re-rank against a 256k-tile
set in four orientations = 1 048 576 rows) and the temporal-smoothing pass, inputs resident in HBM, CUDA events.
Prints one JSON line per case.  Helper for profiles/, not the driver's bench."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tiles", type=int, default=262144)
    ap.add_argument("--queries", type=int, default=1 << 20)
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--only", default="k8,k1,smooth,keyframe")
    args = ap.parse_args()
    import torch
    import tiler_b200 as tb
    from bench import synth_tileset, synth_queries
    dev = torch.device("cuda", 0)
    tb.set_device(0)
    stream = torch.cuda.current_stream()
    tb.set_stream(stream.cuda_stream)
    gen = torch.Generator(device=dev).manual_seed(0x42381337 + 5)
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

    tiles, pal = synth_tileset(torch, args.tiles, 16, gen)
    if only & {"k8", "k1"}:
        tree = tb.KDTree.from_tileset(tiles, pal.view(1, -1))
        rgb = synth_queries(torch, args.queries, tiles[:65536], pal, gen)
        q = torch.empty((args.queries, 192), dtype=torch.float32, device=dev)
        tb.psyv_queries(rgb, out=q)
        for k in (8, 1):
            if f"k{k}" not in only:
                continue
            oi = torch.empty((args.queries, k), dtype=torch.int32, device=dev)
            oe = torch.empty((args.queries, k), dtype=torch.float32, device=dev)
            ms = timed(lambda: tree.search_batch(q, k, out_idx=oi, out_err=oe))
            tb.set_profiling(True)
            tb.reset_stats()
            tree.search_batch(q, k, out_idx=oi, out_err=oe)  # one more pass with the phase events on
            st = tb.stats()
            tb.set_profiling(False)
            flop = 2.0 * args.queries * tree.n * 192
            print(json.dumps({"case": f"search_batch k={k}", "queries": args.queries, "rows": tree.n, "ms": ms,
                              "queries_per_s": args.queries / ms * 1e3, "tflops": flop / ms / 1e9,
                              "fallback_queries": st.get("fallback_queries"),
                              "phases_ms": {"search": st["search_ms"], "coarse": st["coarse_ms"], "rerank": st["rerank_ms"]}}))
        tree.destroy()
    if "keyframe" in only:
        # PrepareGlobalFT + PrepareFrameTiling for one keyframe: 65 536 stored tiles, 16 palettes, the tilemaps of
        # 30 frames at 1080p (972 000 items), ftMedium quality; host buffers, two keyframes (the second reuses the
        # cached 8-NN of the stored tiles)
        import time
        import numpy as np
        g = np.random.default_rng(11)
        nt, npal = 65536, 16
        t_h = tiles[:nt].cpu().numpy()
        pals = g.integers(0, 1 << 24, (npal, 16), dtype=np.int64).astype(np.int32)
        t0 = time.perf_counter()
        gtree, _, _ = tb.KDTree.global_ft(t_h)
        t_global = (time.perf_counter() - t0) * 1e3
        cen = g.standard_normal((npal, 192))
        res = []
        for kf in range(2):
            item_tile = g.integers(0, nt, 972000).astype(np.int32)
            item_pal = g.integers(0, npal, 972000).astype(np.int32)
            t0 = time.perf_counter()
            st2 = gtree.keyframe_dataset(t_h, pals, item_pal, item_tile, quality=1, pal_tol=0.05, pal_centroids=cen)
            res.append(((time.perf_counter() - t0) * 1e3, st2.n))
            st2.destroy()
        gtree.destroy()
        print(json.dumps({"case": "keyframe preparation (host buffers)", "stored_tiles": nt, "palettes": npal,
                          "items": 972000, "global_ft_ms": t_global, "first_keyframe_ms": res[0][0],
                          "next_keyframe_ms": res[1][0], "stage2_rows": res[1][1]}))
    if "smooth" in only:
        # 1080p tilemap, 64 frames, 8 keyframes, 16 palettes: items persist with probability 0.7
        nframes, npos, nkf, npal = 64, 240 * 135, 8, 16
        g = torch.Generator(device="cpu").manual_seed(7)
        pals = torch.randint(0, 1 << 24, (nkf, npal, 16), dtype=torch.int32, generator=g)
        frame_kf = (torch.arange(nframes) // (nframes // nkf)).to(torch.int32)
        tile = torch.empty((nframes, npos), dtype=torch.int32)
        tile[0] = torch.randint(0, args.tiles, (npos,), generator=g, dtype=torch.int32)
        for f in range(1, nframes):
            keep = torch.rand(npos, generator=g) < 0.7
            tile[f] = torch.where(keep, tile[f - 1], torch.randint(0, args.tiles, (npos,), generator=g, dtype=torch.int32))
        palidx = torch.randint(0, npal, (nframes, npos), generator=g, dtype=torch.int32)
        attrs = torch.randint(0, 4, (nframes, npos), generator=g, dtype=torch.uint8)
        import time
        tl = tiles.cpu().numpy()
        args_np = (tl, pals.numpy(), frame_kf.numpy(), tile.numpy(), palidx.numpy(), attrs.numpy(), 0.02)
        tb.temporal_smoothing(*args_np)
        t0 = time.perf_counter()
        out = tb.temporal_smoothing(*args_np)
        ms = (time.perf_counter() - t0) * 1e3
        print(json.dumps({"case": "temporal_smoothing (host buffers, copies included)", "frames": nframes, "positions": npos,
                          "ms": ms, "items_per_s": nframes * npos / ms * 1e3,
                          "smoothed": int(((out[3] & 4) != 0).sum())}))


if __name__ == "__main__":
    main()
