"""How much of the coarse pass's candidate-tracking cost depends on the ORDER of the queries?
A warp of the search epilogue serves 32 consecutive queries and pays one out-of-line call whenever ANY of
them admits a row, so queries with similar score profiles in one warp share their calls.
Runs the C2 workload with the queries (a) as generated, (b) sorted by the planted (tile, mirror) - the upper
bound of what clustering can give, (c) sorted by a key computed from the query pixels alone (mean colour,
Morton-interleaved), (d) a random permutation.  Same process, same box: a fair A/B.  One JSON line."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import tiler_b200 as tb
    import bench

    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    tb.set_device(0)
    stream = torch.cuda.current_stream()
    tb.set_stream(stream.cuda_stream)
    frames, pos, ntiles = bench.WORKLOADS["c2_720p_300f_16k"]
    nq = frames * pos
    gen = torch.Generator(device=dev).manual_seed(0x42381337)
    tiles, pal = bench.synth_tileset(torch, ntiles, 16, gen)
    gen.manual_seed(0x42381337 + 1)
    rgb = bench.synth_queries(torch, nq, tiles, pal, gen)
    tree = tb.KDTree.from_tileset(tiles, pal.view(1, -1).contiguous())
    d_idx = torch.empty(nq, dtype=torch.int32, device=dev)
    d_err = torch.empty(nq, dtype=torch.float32, device=dev)

    def run(q, steps=3):
        for _ in range(2):
            tree.frametiling_search(q, out_idx=d_idx, out_err=d_err)
        torch.cuda.synchronize()
        tb.reset_stats()
        tb.set_profiling(True)
        for _ in range(steps):
            tree.frametiling_search(q, out_idx=d_idx, out_err=d_err)
        torch.cuda.synchronize()
        st = tb.stats()
        tb.set_profiling(False)
        return {"coarse_ms": st["coarse_ms"] / steps, "rerank_ms": st["rerank_ms"] / steps,
                "fallback": st["fallback_queries"] / steps}

    out = {}
    out["as_generated"] = run(rgb)
    base_idx = d_idx.clone()
    # (b) by the answer itself (row index = tile * 4 + attrs): best possible clustering
    order = torch.argsort(base_idx.long(), stable=True)
    out["by_answer"] = run(rgb[order].contiguous())
    assert torch.equal(d_idx, base_idx[order])
    # (c) by mean colour of the query tile, 3 x 7 bits Morton
    r = (rgb & 0xff).float().mean(1)
    g = ((rgb >> 8) & 0xff).float().mean(1)
    b = ((rgb >> 16) & 0xff).float().mean(1)
    qy = ((0.2126 * r + 0.7152 * g + 0.0722 * b) / 2).long().clamp_(0, 127)
    qu = (((b - (0.2126 * r + 0.7152 * g + 0.0722 * b)) * 0.5 / 0.9278 + 128) / 2).long().clamp_(0, 127)
    qv = (((r - (0.2126 * r + 0.7152 * g + 0.0722 * b)) * 0.5 / 0.7874 + 128) / 2).long().clamp_(0, 127)

    def spread(x):
        o = torch.zeros_like(x)
        for i in range(7):
            o |= ((x >> i) & 1) << (3 * i)
        return o
    key = (spread(qy) << 2) | (spread(qu) << 1) | spread(qv)
    order = torch.argsort(key, stable=True)
    out["by_mean_colour_morton"] = run(rgb[order].contiguous())
    assert torch.equal(d_idx, base_idx[order])
    # (c2) luma-major: Y at 8 bits, then U, V at 4 bits
    key2 = ((0.2126 * r + 0.7152 * g + 0.0722 * b).long().clamp_(0, 255) << 8) | ((qu >> 3) << 4) | (qv >> 3)
    order = torch.argsort(key2, stable=True)
    out["by_luma_then_chroma"] = run(rgb[order].contiguous())
    # (d) random permutation
    order = torch.randperm(nq, device=dev, generator=gen)
    out["random_permutation"] = run(rgb[order].contiguous())
    print(json.dumps(out))


if __name__ == "__main__":
    main()
