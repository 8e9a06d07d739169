"""Committed golden vectors (tests/golden/ft_small.npz, made by tests/golden/make_golden.py).

CPU tests: the oracle still reproduces them (guards the checker itself).
GPU tests: the CUDA path, called through the C ABI, reproduces them — without touching the oracle.
PARITY UNPINNED: the vectors come from the restatement, the reference cannot run here.
"""
import os

import numpy as np
import pytest

G = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ft_small.npz"))


# ------------------------------------------------------------------ CPU: oracle vs golden
def test_oracle_reproduces_golden_descriptors(oracle):
    o = oracle
    assert np.array_equal(o.psyv_queries_f32(G["rgb"], True, -1), G["q_haar"])
    assert np.array_equal(o.psyv_queries_f32(G["rgb"][:16], False, -1), G["q_dct"])
    assert np.array_equal(o.psyv_queries_f32(G["rgb"][:16], True, 0), G["q_haar_gamma0"])
    assert np.array_equal(o.psyv_tileset_f32(G["tiles"], G["pal"], True, -1), G["ds_haar"])
    assert np.array_equal(o.psyv_tileset_f32(G["tiles"][:8], G["pal2"], False, -1), G["ds_dct"])


def test_oracle_reproduces_golden_search_and_kmodes(oracle):
    o = oracle
    i1, d1, nxt = o.knn_f64(G["ds_haar"], G["q_haar"], 1)
    assert np.array_equal(i1[:, 0], G["nn_idx"]) and np.array_equal(d1[:, 0], G["nn_dist_f64"])
    kd = o.KDTreeRef(G["ds_haar"])
    ki, kdist = kd.search(G["q_haar"], 1)
    assert np.array_equal(kdist[:, 0], G["nn_dist_f32"])
    g, tt, ta = o.global_rows_f32(G["tiles"])
    assert np.array_equal(g, G["g_rows"]) and np.array_equal(ta, G["tr_attrs"]) and np.array_equal(tt, G["tr_tile"])
    ca, da = o.kmodes_assign(G["km_X"], G["km_cent"], o.ASM)
    cg, dg = o.kmodes_assign(G["km_X"], G["km_cent"], o.GENERIC)
    assert np.array_equal(ca, G["km_clust_asm"]) and np.array_equal(da, G["km_dis_asm"])
    assert np.array_equal(cg, G["km_clust_gen"]) and np.array_equal(dg, G["km_dis_gen"])
    assert not np.any(ca == 0)  # centroid 1 duplicates centroid 0: the last index must win


# ------------------------------------------------------------------ GPU: CUDA path vs golden
@pytest.fixture(scope="module")
def tb():
    import tiler_b200 as t
    t.set_device(0)
    return t


@pytest.mark.gpu
def test_cuda_descriptors_match_golden(tb):
    assert np.array_equal(tb.psyv_queries(G["rgb"], True, -1), G["q_haar"])
    assert np.array_equal(tb.psyv_queries(G["rgb"][:16], False, -1), G["q_dct"])
    assert np.array_equal(tb.psyv_queries(G["rgb"][:16], True, 0), G["q_haar_gamma0"])
    with tb.KDTree.from_tileset(G["tiles"], G["pal"]) as t:
        assert np.array_equal(t.rows(), G["ds_haar"])
    with tb.KDTree.from_tileset(G["tiles"][:8], G["pal2"], wavelets=False) as t:
        assert np.array_equal(t.rows(), G["ds_dct"])


@pytest.mark.gpu
@pytest.mark.parametrize("engine", ["simt", "tc"])
def test_cuda_search_matches_golden(tb, engine):
    tb.set_engine(tb.ENGINE_SIMT if engine == "simt" else tb.ENGINE_TC)
    try:
        with tb.KDTree.from_tileset(G["tiles"], G["pal"]) as t:
            idx, err = t.frametiling_search(G["rgb"])
        d1, d2 = G["nn_dist_f64"], G["nn_next_f64"]
        solid = (d2 - d1) > 1e-5 * np.maximum(d1, 1e-12)
        assert np.array_equal(idx[solid], G["nn_idx"][solid])
        assert np.allclose(err, d1, rtol=1e-5, atol=1e-7)
        assert np.array_equal(err, G["nn_dist_f32"])  # same float32 add chain as ANN's leaf scan
        tree, tr_tile, tr_attrs = tb.KDTree.global_ft(G["tiles"])
        with tree:
            assert np.array_equal(tree.rows(), G["g_rows"])
            assert np.array_equal(tr_attrs, G["tr_attrs"]) and np.array_equal(tr_tile, G["tr_tile"])
            i8, d8 = tree.search_batch(G["g_rows"][::7], 8)
            assert np.array_equal(d8, G["k8_dist"])
            for r in range(len(i8)):  # integer-valued rows: ties are common, any tied row is legal
                for j in range(8):
                    if i8[r, j] != G["k8_idx"][r, j]:
                        assert d8[r, j] in G["k8_dist"][r]
    finally:
        tb.set_engine(tb.ENGINE_AUTO)


@pytest.mark.gpu
def test_cuda_kmodes_matches_golden(tb):
    ca, da = tb.kmodes_assign(G["km_X"], G["km_cent"], tb.KM_ASM)
    cg, dg = tb.kmodes_assign(G["km_X"], G["km_cent"], tb.KM_GENERIC)
    assert np.array_equal(ca, G["km_clust_asm"]) and np.array_equal(da, G["km_dis_asm"])
    assert np.array_equal(cg, G["km_clust_gen"]) and np.array_equal(dg, G["km_dis_gen"])
    md = tb.kmodes_update_min_distance(G["km_X"], G["km_X"][3],
                                       np.full(len(G["km_X"]), np.iinfo(np.uint64).max, np.uint64))
    assert np.array_equal(md, G["km_mindist"])
