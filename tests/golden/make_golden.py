#!/usr/bin/env python
"""Regenerates tests/golden/ft_small.npz from the CPU oracle (oracle/).

The reference (b0nefish/tiler) cannot run here (FreePascal + Win32 + ANN.dll), so these
vectors are produced by the restatement, not by the reference: PARITY UNPINNED.  They freeze
the oracle's answers on fixed seeded inputs (the reference's own LCG, kmodes.pas:82-86) so
(a) an oracle regression is caught on CPU, (b) the CUDA path is checked against committed
numbers on the GPU box, (c) a future FPC build of the reference can be diffed on the same inputs.

    python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import tmoracle as o  # noqa: E402


def main():
    pal, s = o.gen_palette(o.BASE_SEED + 7)
    pal2, s = o.gen_palette(s)
    tiles, s = o.gen_tileset(s, 48)
    rgb, s = o.gen_queries(s, 96, tiles, pal)
    out = dict(pal=pal, pal2=pal2, tiles=tiles, rgb=rgb)
    # descriptors (float32 = what crosses the FFI, main.pas:3913-3914, 4024-4025)
    out["q_haar"] = o.psyv_queries_f32(rgb, True, -1)
    out["q_dct"] = o.psyv_queries_f32(rgb[:16], False, -1)
    out["q_haar_gamma0"] = o.psyv_queries_f32(rgb[:16], True, 0)
    out["ds_haar"] = o.psyv_tileset_f32(tiles, pal, True, -1)
    out["ds_dct"] = o.psyv_tileset_f32(tiles[:8], pal2, False, -1)
    # stage-2 exact 1-NN (DoFrameTiling main.pas:4027) and the gap to the runner-up
    i1, d1, nxt = o.knn_f64(out["ds_haar"], out["q_haar"], 1)
    f1i, f1d = o.knn_f32(out["ds_haar"], out["q_haar"], 1)
    out.update(nn_idx=i1[:, 0], nn_dist_f64=d1[:, 0], nn_next_f64=nxt, nn_idx_f32=f1i[:, 0], nn_dist_f32=f1d[:, 0])
    # stage-1 rows and k = 8 pre-filter (PrepareGlobalFT main.pas:3736-3780, UseOne 3830)
    g, tr_tile, tr_attrs = o.global_rows_f32(tiles)
    k8i, k8d = o.knn_f32(g, g[::7], 8)
    out.update(g_rows=g, tr_tile=tr_tile, tr_attrs=tr_attrs, k8_dist=k8d, k8_idx=k8i)
    # KModes rows, centroids, assignment in both semantics, one UpdateMinDistance pass
    X = o.kmodes_rows(tiles)
    cent = X[::5].copy()
    cent[1] = cent[0]  # duplicate centroid: "last index wins" is exercised
    ca, da = o.kmodes_assign(X, cent, o.ASM)
    cg, dg = o.kmodes_assign(X, cent, o.GENERIC)
    md = o.update_min_distance(X, X[3], np.full(len(X), np.iinfo(np.uint64).max, np.uint64), o.ASM)
    ffc, ffi = o.farthest_first(X, 6, 2, o.ASM)
    out.update(km_X=X, km_cent=cent, km_clust_asm=ca, km_dis_asm=da, km_clust_gen=cg, km_dis_gen=dg,
               km_mindist=md, km_ff_chosen=ffi)
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "ft_small.npz")
    np.savez_compressed(path, **out)
    print(path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
