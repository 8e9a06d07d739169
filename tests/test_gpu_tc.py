"""Tensor-core engine vs the exact SIMT engine: results must be IDENTICAL (index and float32
distance bit-for-bit), because uncertified queries are re-run exactly; and vs the oracle."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def tb():
    import tiler_b200
    return tiler_b200


def _both(tb, tree, q, k):
    tb.set_engine(tb.ENGINE_SIMT)
    si, se = tree.search_batch(q, k=k)
    tb.set_engine(tb.ENGINE_TC)
    tb.reset_stats()
    ti, te = tree.search_batch(q, k=k)
    st = tb.stats()
    tb.set_engine(tb.ENGINE_AUTO)
    return si, se, ti, te, st


def _both_auto(tb, tree, q, k=1):
    tb.set_engine(tb.ENGINE_SIMT)
    si, se = tree.search_batch(q, k=k)
    tb.set_engine(tb.ENGINE_AUTO)
    tb.reset_stats()
    tb.set_profiling(True)
    ti, te = tree.search_batch(q, k=k)
    st = tb.stats()
    tb.set_profiling(False)
    return si, se, ti, te, st


@pytest.mark.parametrize("ntiles,nq,k", [(512, 1500, 1), (2048, 5000, 1), (300, 257, 1), (700, 3000, 4), (512, 900, 8)])
def test_tc_equals_simt_psyv(tb, oracle, ntiles, nq, k):
    o = oracle
    pal, s = o.gen_palette(o.BASE_SEED + ntiles)
    tiles, s = o.gen_tileset(s, ntiles)
    rgb, s = o.gen_queries(s, nq, tiles, pal)
    ds = o.psyv_tileset_f32(tiles, pal)
    q = o.psyv_queries_f32(rgb)
    with tb.KDTree.from_rows(ds) as t:
        si, se, ti, te, st = _both(tb, t, q, k)
    assert np.array_equal(si, ti)
    assert np.array_equal(se.view(np.uint32), te.view(np.uint32))
    assert st["coarse_launches"] > 0, st  # counted unconditionally: the tcgen05 kernel ran, not the exact engine
    assert st["fallback_queries"] < 0.2 * nq, st  # the coarse pass must certify the bulk
    oi, od = o.knn_f32(ds, q, k)
    assert np.array_equal(te.view(np.uint32), od.view(np.uint32))


@pytest.mark.parametrize("dd,k", [(96, 1), (128, 1), (128, 8), (64, 1), (80, 4), (192, 8)])
def test_tc_every_operand_width(tb, oracle, dd, k):
    """NCH = 1 (dd <= 64), 2 (64 < dd <= 128) and 3 (128 < dd <= 192) instantiations of the kernel, padded operand
    rows included (dd = 80, 96: kp = 128); the forced tensor-core engine must have launched and must equal the exact
    engine bit for bit."""
    rng = np.random.default_rng(1000 + dd + k)
    ds = rng.standard_normal((9000, dd)).astype(np.float32)
    q = (ds[rng.integers(0, 9000, 2500)] + 0.1 * rng.standard_normal((2500, dd))).astype(np.float32)
    with tb.KDTree.from_rows(ds) as t:
        si, se, ti, te, st = _both(tb, t, q, k)
    assert st["coarse_launches"] > 0, st
    assert np.array_equal(si, ti) and np.array_equal(se.view(np.uint32), te.view(np.uint32))
    oi, od = oracle.knn_f32(ds, q[:64], k)
    assert np.array_equal(te[:64].view(np.uint32), od.view(np.uint32))


@pytest.mark.parametrize("nblocks,extra,k", [(75, 0, 1), (80, 5, 1), (100, -3, 1), (149, 1, 8), (81, 0, 4), (223, 17, 1)])
def test_tc_tail_plan_equals_exact_engine(tb, oracle, nblocks, extra, k):
    """Query-block counts that leave a few blocks for the last round of the 74 CTA pairs: those blocks are cut into
    finer row splits (TcPlan, search_tc.h) and carry more candidate lists than the others.  Bit-identical to the exact
    engine, whichever part of the plan a query falls in."""
    rng = np.random.default_rng(nblocks * 7 + k)
    n, nq = 9000, nblocks * 256 + extra
    ds = rng.standard_normal((n, 192)).astype(np.float32)
    q = (ds[rng.integers(0, n, nq)] + 0.2 * rng.standard_normal((nq, 192))).astype(np.float32)
    with tb.KDTree.from_rows(ds) as t:
        si, se, ti, te, st = _both(tb, t, q, k)
    assert st["coarse_launches"] > 0, st
    assert np.array_equal(si, ti) and np.array_equal(se.view(np.uint32), te.view(np.uint32))
    oi, od = oracle.knn_f32(ds, q[-64:], k)           # the tail blocks against the oracle
    assert np.array_equal(te[-64:].view(np.uint32), od.view(np.uint32))


def test_tc_envelope_is_visible_in_the_counters(tb):
    """k > 8 and datasets wider than 192 are outside the tensor-core path: the exact engine answers, and the
    launch counter says so (no silent substitution inside a 'tensor-core' test)."""
    rng = np.random.default_rng(3)
    ds = rng.standard_normal((2000, 192)).astype(np.float32)
    q = rng.standard_normal((300, 192)).astype(np.float32)
    with tb.KDTree.from_rows(ds) as t:
        tb.reset_stats()
        tb.set_engine(tb.ENGINE_TC)
        t.search_batch(q, k=12)
        assert tb.stats()["coarse_launches"] == 0
        t.search_batch(q, k=8)
        assert tb.stats()["coarse_launches"] > 0
        tb.set_engine(tb.ENGINE_AUTO)


def test_tc_random_gaussian_hard_case(tb, oracle):
    # no planted answers: many near-ties, exercises the certificate + fallback
    rng = np.random.default_rng(11)
    ds = rng.standard_normal((3000, 192)).astype(np.float32)
    q = rng.standard_normal((700, 192)).astype(np.float32)
    with tb.KDTree.from_rows(ds) as t:
        si, se, ti, te, st = _both(tb, t, q, 1)
    assert np.array_equal(si, ti) and np.array_equal(se.view(np.uint32), te.view(np.uint32))


def test_tc_dd64_palette_rows(tb, oracle, small_case):
    rows, _, _ = oracle.global_rows_f32(small_case["tiles"])
    q = rows[::5][:400]
    with tb.KDTree.from_rows(rows) as t:
        si, se, ti, te, st = _both(tb, t, q, 8)
    assert np.array_equal(se, te)
    assert np.array_equal(si, ti)


def test_tc_duplicates_and_ragged(tb, oracle):
    rng = np.random.default_rng(12)
    base = rng.standard_normal((130, 192)).astype(np.float32)
    ds = np.concatenate([base, base[:40], base[:7]])  # exact duplicates: lowest index must win
    q = base[:100] + 1e-3 * rng.standard_normal((100, 192)).astype(np.float32)
    with tb.KDTree.from_rows(ds) as t:
        si, se, ti, te, st = _both(tb, t, q, 1)
    assert np.array_equal(si, ti) and np.array_equal(se.view(np.uint32), te.view(np.uint32))
    assert np.array_equal(ti[:, 0], np.arange(100))


def test_tc_out_of_fp16_range_falls_back_to_exact(tb, oracle):
    rng = np.random.default_rng(13)
    ds = (rng.standard_normal((500, 192)) * 1e5).astype(np.float32)
    q = (rng.standard_normal((100, 192)) * 1e5).astype(np.float32)
    with tb.KDTree.from_rows(ds) as t:
        tb.set_engine(tb.ENGINE_AUTO)
        ti, te = t.search_batch(q, k=1)
    oi, od = oracle.knn_f32(ds, q, 1)
    assert np.array_equal(ti, oi) and np.array_equal(te, od)
    # dataset in range, one query far outside it
    ds = rng.standard_normal((500, 192)).astype(np.float32)
    q = rng.standard_normal((100, 192)).astype(np.float32)
    q[3] *= 1e6
    with tb.KDTree.from_rows(ds) as t:
        ti, te = t.search_batch(q, k=1)
    oi, od = oracle.knn_f32(ds, q, 1)
    assert np.array_equal(ti, oi) and np.array_equal(te, od)


@pytest.mark.parametrize("layout", ["interleaved", "contiguous"])
def test_tc_second_chance_for_rejected_queries(tb, oracle, layout):
    """Clusters of 40 near-identical rows: with k = 8 more rows sit inside the certificate's band than a
    16-entry list holds, so the first pass rejects the queries.  Interleaved clusters spread over the 16
    lists of the finest row split (second tensor-core pass certifies them); contiguous clusters overflow
    there too and end on the exact engine.  Either way the result is the exact engine's, bit for bit."""
    rng = np.random.default_rng(17)
    nbase, rep = 3200, 40                      # 128 000 rows
    base = rng.standard_normal((nbase, 192)).astype(np.float32)
    jit = (1e-4 * rng.standard_normal((rep, nbase, 192))).astype(np.float32)
    ds = base[None] + jit                      # [rep][nbase][192]
    ds = ds.reshape(-1, 192) if layout == "interleaved" else ds.transpose(1, 0, 2).reshape(-1, 192)
    ds = np.ascontiguousarray(ds)
    q = base[:2304] + (1e-3 * rng.standard_normal((2304, 192))).astype(np.float32)
    with tb.KDTree.from_rows(ds) as t:
        si, se, ti, te, st = _both(tb, t, q, 8)
    assert np.array_equal(se.view(np.uint32), te.view(np.uint32))
    assert np.array_equal(si, ti)
    if layout == "interleaved":
        assert st["fallback_queries"] < 100, st   # certified on the second pass
    else:
        assert st["fallback_queries"] > 2000, st  # 40 rows of one cluster in one list: exact engine
    oi, od = oracle.knn_f32(ds, q[:64], 8)
    assert np.array_equal(te[:64].view(np.uint32), od.view(np.uint32))


@pytest.mark.parametrize("seed", range(6))
def test_tc_certificate_soundness_sweep(tb, seed):
    """The certificate's error model (search_tc.h) against data it was not tuned on: offsets, heavy-tailed norms,
    near-duplicate clusters at several jitter scales, integer rows.  Tensor-core results must equal the exact
    engine's bit for bit whatever the coarse pass thinks."""
    rng = np.random.default_rng(100 + seed)
    n, nq = 6000, 1800
    base = rng.standard_normal((n // 6, 192))
    jitter = 10.0 ** rng.uniform(-5, -1, size=(6, 1, 1))
    ds = (base[None] + jitter * rng.standard_normal((6, n // 6, 192))).reshape(n, 192)  # clusters, mixed scales
    scale = np.where(rng.random((n, 1)) < 0.02, 40.0, 1.0)                              # a few huge rows
    offset = rng.standard_normal((1, 192)) * 10.0 ** rng.uniform(-1, 2)                 # common offset
    ds = (ds * scale + offset)
    if seed == 5:
        ds = np.round(ds * 4)  # integer rows (exactly representable after centring only by luck)
    ds = ds.astype(np.float32)
    pick = rng.integers(0, n, nq)
    q = (ds[pick] + (10.0 ** rng.uniform(-4, 0, size=(nq, 1))) * rng.standard_normal((nq, 192))).astype(np.float32)
    with tb.KDTree.from_rows(ds) as t:
        for k in (1, 4, 8):
            si, se, ti, te, st = _both(tb, t, q, k)
            assert np.array_equal(se.view(np.uint32), te.view(np.uint32)), (seed, k)
            assert np.array_equal(si, ti), (seed, k)


def test_tc_error_model_audit(tb, oracle, small_case):
    """Profiling level 2 reports the largest observed coarse-score error over all re-ranked candidates as a
    fraction of the bound the certificate assumes (search_tc.h): it must be positive and well below 1."""
    o = oracle
    ds = o.psyv_tileset_f32(small_case["tiles"], small_case["pal"])
    q = o.psyv_queries_f32(small_case["rgb"])
    rng = np.random.default_rng(5)
    cases = [(ds, q), (rng.standard_normal((4000, 192)).astype(np.float32) * 7 + 3,
                       rng.standard_normal((1500, 192)).astype(np.float32) * 7 + 3)]
    for rows, qq in cases:
        with tb.KDTree.from_rows(rows) as t:
            tb.set_engine(tb.ENGINE_TC)
            tb.reset_stats()
            tb.set_profiling(2)
            for k in (1, 8):
                t.search_batch(qq, k=k)
            st = tb.stats()
            tb.set_profiling(0)
            tb.set_engine(tb.ENGINE_AUTO)
        assert st["audited_candidates"] >= 2 * len(qq), st
        assert 0.0 < st["max_error_ratio"] < 0.5, st


def test_tc_nan_and_inf_rows_take_the_exact_engine(tb, oracle):
    """A NaN or Inf anywhere in the dataset must be seen when the fp16 operand is built (the tensor-core path is
    then not used for that dataset); results equal the exact engine's."""
    rng = np.random.default_rng(21)
    q = rng.standard_normal((300, 192)).astype(np.float32)
    for bad in (np.nan, np.inf):
        ds = rng.standard_normal((5000, 192)).astype(np.float32)
        ds[4321, 17] = bad
        with tb.KDTree.from_rows(ds) as t:
            si, se, ti, te, st = _both_auto(tb, t, q)
        assert np.array_equal(si, ti) and np.array_equal(se.view(np.uint32), te.view(np.uint32))
        assert st["coarse_launches"] == 0


@pytest.mark.parametrize("k", [1, 2, 8])
def test_tc_long_items_equal_exact_engine(tb, oracle, k):
    """k = 1: 1 024 tiles per work item select the norm-free kernel over the operand in scan order (rows sorted by
    norm, candidate positions translated back through perm[]), here with the queries in the caller's order.
    k > 1 against 262 144 rows with enough query blocks for one work item per block (1 024 tiles per item): the
    first 16 tiles of every item are scanned for the k smallest group minima only, seed the admission threshold and
    are scanned again at the end (search_tc2_kernel.inc).  Results must still be the exact engine's, bit for bit;
    clusters of near-identical rows inside the seed tiles make the seeded threshold tight from the start."""
    rng = np.random.default_rng(29 + k)
    n, nq = 262144, 20000
    ds = rng.standard_normal((n, 192)).astype(np.float32)
    ds[:4096:4] = ds[1:4096:4] + (1e-3 * rng.standard_normal((1024, 192))).astype(np.float32)  # near-duplicates up front
    q = ds[rng.integers(0, n, nq)] + (0.05 * rng.standard_normal((nq, 192))).astype(np.float32)
    q[:512] = ds[:2048:4][:512] + (1e-3 * rng.standard_normal((512, 192))).astype(np.float32)  # answers inside the seed tiles
    with tb.KDTree.from_rows(ds) as t:
        si, se, ti, te, st = _both(tb, t, q, k)
    assert st["coarse_launches"] > 0, st
    assert np.array_equal(se.view(np.uint32), te.view(np.uint32))
    assert np.array_equal(si, ti)
    assert st["fallback_queries"] < nq // 20, st
    oi, od = oracle.knn_f32(ds, q[:32], k)
    assert np.array_equal(te[:32].view(np.uint32), od.view(np.uint32))


@pytest.mark.parametrize("k,nf_min", [(1, "32"), (1, "1000000"), (2, None), (8, None)])
def test_tc_chained_plan_equals_exact_engine(k, nf_min):
    """Tilesets beyond the L2: the rows are cut into splits, the work items run split-major and every item continues
    the candidate lists and the threshold its query block reached in the previous split (search_tc.h: TcPlan.chain;
    k = 1: the candidate ring, with the norm-free kernel over the scan-order operand - TM_TC_NF_MIN=32 - and with the
    norm MMA).  Forced here on 131 072 rows with TM_TC_CHAIN=2 in a process of its own (the switches are read once);
    TM_TC_CHAIN=0 in a second process must give the same, exact, answer."""
    import json
    import os
    import subprocess
    import sys
    here = os.path.dirname(os.path.abspath(__file__))
    for chain in ("2", "0"):
        env = dict(os.environ, TM_TC_CHAIN=chain)
        if nf_min:
            env["TM_TC_NF_MIN"] = nf_min
        out = subprocess.run([sys.executable, os.path.join(here, "chain_check.py"), str(k)], env=env, capture_output=True,
                             text=True, timeout=600)
        assert out.returncode == 0, out.stderr[-2000:]
        res = json.loads(out.stdout.strip().splitlines()[-1])
        assert res["idx_equal"] and res["err_equal"], (chain, res)
        assert res["coarse_launches"] > 0, res
        assert res["fallback_queries"] < res["nq"] // 10, (chain, res)
