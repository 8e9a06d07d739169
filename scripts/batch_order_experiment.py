"""search_batch (float descriptor rows, k = 8 and k = 1, C5-shaped shard) with the rows as generated vs sorted by a
Morton code of the three plane DC coefficients: would the coherence order pay on the generic batch path too?"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import tiler_b200 as tb
    from bench import synth_tileset, synth_queries
    dev = torch.device("cuda", 0)
    tb.set_device(0)
    stream = torch.cuda.current_stream()
    tb.set_stream(stream.cuda_stream)
    gen = torch.Generator(device=dev).manual_seed(0x42381337 + 5)
    ntiles, nq = 262144, 1 << 20
    tiles, pal = synth_tileset(torch, ntiles, 16, gen)
    tree = tb.KDTree.from_tileset(tiles, pal.view(1, -1))
    rgb = synth_queries(torch, nq, tiles[:65536], pal, gen)
    q = torch.empty((nq, 192), dtype=torch.float32, device=dev)
    tb.psyv_queries(rgb, out=q)

    def spread(x):
        o = torch.zeros_like(x)
        for i in range(7):
            o |= ((x >> i) & 1) << (3 * i)
        return o
    dc = q[:, [0, 64, 128]]
    lo, hi = dc.min(0).values, dc.max(0).values
    qq = ((dc - lo) / (hi - lo) * 127.99).long().clamp_(0, 127)
    key = (spread(qq[:, 0]) << 2) | (spread(qq[:, 1]) << 1) | spread(qq[:, 2])
    qs = q[torch.argsort(key, stable=True)].contiguous()
    out = {}
    for k in (8, 1):
        oi = torch.empty((nq, k), dtype=torch.int32, device=dev)
        oe = torch.empty((nq, k), dtype=torch.float32, device=dev)
        for name, x in (("as_generated", q), ("dc_morton_sorted", qs), ("as_generated_again", q)):
            tree.search_batch(x, k, out_idx=oi, out_err=oe)
            torch.cuda.synchronize()
            tb.set_profiling(True)
            tb.reset_stats()
            for _ in range(2):
                tree.search_batch(x, k, out_idx=oi, out_err=oe)
            torch.cuda.synchronize()
            st = tb.stats()
            tb.set_profiling(False)
            out[f"k{k}_{name}"] = {"coarse_ms": st["coarse_ms"] / 2, "rerank_ms": st["rerank_ms"] / 2,
                                   "search_ms": st["search_ms"] / 2, "fallback": st["fallback_queries"] / 2,
                                   "tflops": 2.0 * nq * tree.n * 192 / (st["coarse_ms"] / 2) / 1e9}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
