"""GPU parity tests proper: the CUDA path, called through the C ABI, against the CPU
oracle on the same seeded inputs.

Bars (north_star): descriptor rows bit-exact (FP64 arithmetic in the reference's
operation order); search index bit-exact wherever the best/second-best gap exceeds
EPS_GAP relative, distances within RTOL of the oracle's float64 value and bit-equal
to its float32 sequential accumulation; KModes assignments and dissimilarities bit-exact.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

EPS_GAP = 1e-5   # relative best/second-best gap below which a query is in the "gap set"
RTOL = 1e-5      # float32 accumulation over 192 terms: 192 * 2^-24 ~ 1.1e-5


@pytest.fixture(scope="module")
def tb():
    import tiler_b200
    return tiler_b200


def _check_search(oracle, ds, q, idx, err, k=1):
    oi, od, nxt = oracle.knn_f64(ds, q, k)
    o32i, o32d = oracle.knn_f32(ds, q, k)
    idx = np.asarray(idx).reshape(len(q), k)
    err = np.asarray(err).reshape(len(q), k)
    # distances: every returned distance is the true distance of the returned row
    true = np.array([[np.sum((q[i].astype(np.float64) - ds[idx[i, j]].astype(np.float64)) ** 2)
                      for j in range(k)] for i in range(len(q))])
    assert np.allclose(err, true, rtol=RTOL, atol=1e-7)
    assert np.allclose(err, od, rtol=RTOL, atol=1e-7)
    assert np.all(np.diff(err, axis=1) >= 0)
    # indices: exact outside the gap set (gap judged on consecutive oracle distances)
    allod = np.concatenate([od, nxt[:, None]], axis=1)
    gap = np.diff(allod, axis=1) > EPS_GAP * np.maximum(allod[:, :-1], 1e-12)
    prev_ok = np.concatenate([np.ones((len(q), 1), bool), gap[:, :-1]], axis=1)
    solid = gap & prev_ok
    assert np.array_equal(idx[solid], oi[solid])
    return int((~solid).sum())


def test_psyv_queries_bit_exact(tb, oracle, small_case):
    for wavelets in (True, False):
        for gamma in (-1, 0, 1):
            got = tb.psyv_queries(small_case["rgb"][:300], wavelets=wavelets, gamma=gamma)
            want = oracle.psyv_queries_f32(small_case["rgb"][:300], use_wavelets=wavelets, gamma=gamma)
            assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), (wavelets, gamma)


def test_psyv_edge_tiles(tb, oracle):
    rgb = np.stack([np.zeros(64, np.int32), np.full(64, 0x00FFFFFF, np.int32),
                    np.arange(64, dtype=np.int32) * 0x00040404, np.full(64, 0x7F0000FF, np.int32)])
    got = tb.psyv_queries(rgb)
    want = oracle.psyv_queries_f32(rgb)  # the top byte of $00BBGGRR is ignored (FromRGB main.pas:571)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    assert tb.psyv_queries(np.zeros((0, 64), np.int32)).shape == (0, 192)


def test_tileset_dataset_bit_exact(tb, oracle, small_case):
    for wavelets in (True, False):
        with tb.KDTree.from_tileset(small_case["tiles"], small_case["pal"], wavelets=wavelets) as t:
            assert (t.n, t.dd) == (512 * 4, 192)
            want = oracle.psyv_tileset_f32(small_case["tiles"], small_case["pal"], use_wavelets=wavelets)
            assert np.array_equal(t.rows().view(np.uint32), want.view(np.uint32))


def test_tileset_dataset_multi_palette_and_tile_flags(tb, oracle, small_case):
    o = oracle
    pals = np.stack([o.gen_palette(o.BASE_SEED + i)[0] for i in range(3)])
    tiles = small_case["tiles"][:100]
    flags = (np.arange(100) % 4).astype(np.uint8)
    with tb.KDTree.from_tileset(tiles, pals, tile_flags=flags) as t:
        got = t.rows().reshape(3, 100, 2, 2, 192)
    for p in range(3):
        for ti in (0, 1, 2, 3, 57):
            for vm in (0, 1):
                for hm in (0, 1):
                    # hmir xor T^.HMirror, vmir xor T^.VMirror (main.pas:3912)
                    want = o.psyv(palpix=tiles[ti], pal=pals[p], hmirror=hm ^ (flags[ti] & 1),
                                  vmirror=vm ^ (flags[ti] >> 1)).astype(np.float32)
                    assert np.array_equal(got[p, ti, vm, hm], want)


def test_global_ft_rows(tb, oracle, small_case):
    tree, tr_tile, tr_attrs = tb.KDTree.global_ft(small_case["tiles"])
    want, wt, wa = oracle.global_rows_f32(small_case["tiles"])
    assert np.array_equal(tree.rows(), want) and np.array_equal(tr_tile, wt) and np.array_equal(tr_attrs, wa)
    # stage-1 search (UseOne main.pas:3830): k = 8 on palette-index rows, exact small integers
    q = want[::37][:40]
    idx, err = tree.search_batch(q, k=8)
    oi, od, _ = oracle.knn_f64(want, q, 8)
    assert np.array_equal(err, od.astype(np.float32))
    assert np.array_equal(idx[:, 0] // 4, oi[:, 0] // 4) and np.all(err[:, 0] == 0)
    tree.destroy()


@pytest.mark.parametrize("engine", ["simt", "auto"])
def test_search_batch_vs_oracle(tb, oracle, small_case, engine):
    tb.set_engine(tb.ENGINE_SIMT if engine == "simt" else tb.ENGINE_AUTO)
    try:
        ds = oracle.psyv_tileset_f32(small_case["tiles"], small_case["pal"])
        q = oracle.psyv_queries_f32(small_case["rgb"])
        with tb.KDTree.from_rows(ds) as t:
            idx, err = t.search_batch(q, k=1)
            _check_search(oracle, ds, q, idx, err, 1)
            o32i, o32d = oracle.knn_f32(ds, q, 1)
            assert np.array_equal(err.view(np.uint32), o32d.view(np.uint32))  # same fp32 accumulation order
            idx8, err8 = t.search_batch(q[:200], k=8)
            _check_search(oracle, ds, q[:200], idx8, err8, 8)
    finally:
        tb.set_engine(tb.ENGINE_AUTO)


def test_legacy_entry_points(tb, oracle, small_case):
    ds = oracle.psyv_tileset_f32(small_case["tiles"][:64], small_case["pal"])
    q = oracle.psyv_queries_f32(small_case["rgb"][:8])
    t = tb.KDTree.create(ds)  # array-of-row-pointers path (extern.pas:63)
    oi, od = oracle.knn_f32(ds, q, 8)
    for i in range(8):
        idx, err = t.search(q[i])
        assert idx == oi[i, 0] and np.float32(err) == od[i, 0]
        assert t.pri_search(q[i]) == (idx, err)
        idxs, errs = t.search_multi(q[i], 8)
        assert np.array_equal(idxs, oi[i]) and np.array_equal(errs, od[i])
    t.destroy()
    t.destroy()  # idempotent / NULL-safe


def test_concurrent_callers_share_handles(tb, oracle, small_case):
    """The reference calls ANN from a pool of worker threads over one shared tree (DoFrameTiling, one frame per
    thread, main.pas:972 -> 4033): concurrent calls of every flavour must return what serial calls return."""
    import threading
    ds = oracle.psyv_tileset_f32(small_case["tiles"][:256], small_case["pal"])
    t = tb.KDTree.create(ds)
    ft = tb.KDTree.from_tileset(small_case["tiles"], small_case["pal"])
    rgb = small_case["rgb"]
    q = oracle.psyv_queries_f32(rgb[:64])
    X = oracle.kmodes_rows(small_case["tiles"][:600])
    Cn = X[::7].copy()
    want_single = [t.search(q[i]) for i in range(64)]
    want_batch = t.search_batch(q, 4)
    want_ft = [ft.frametiling_search(rgb[o:o + 300 + 17 * j]) for j, o in enumerate(range(0, 1200, 300))]
    want_km = [tb.kmodes_assign(X[o:o + 150], Cn[: 20 + 11 * j]) for j, o in enumerate(range(0, 600, 150))]
    errors = []

    def worker(w):
        try:
            for rep in range(6):
                kind = (w + rep) % 4
                if kind == 0:
                    for i in range(w % 8, 64, 8):
                        assert t.search(q[i]) == want_single[i]
                elif kind == 1:
                    gi, ge = t.search_batch(q, 4)
                    assert np.array_equal(gi, want_batch[0]) and np.array_equal(ge, want_batch[1])
                elif kind == 2:
                    j = (w + rep) % 4
                    gi, ge = ft.frametiling_search(rgb[300 * j:300 * j + 300 + 17 * j])
                    assert np.array_equal(gi, want_ft[j][0]) and np.array_equal(ge, want_ft[j][1])
                else:
                    j = (w + rep) % 4
                    gc, gd = tb.kmodes_assign(X[150 * j:150 * j + 150], Cn[: 20 + 11 * j])
                    assert np.array_equal(gc, want_km[j][0]) and np.array_equal(gd, want_km[j][1])
        except BaseException as e:  # noqa: BLE001 - reported in the main thread
            errors.append((w, repr(e)))

    threads = [threading.Thread(target=worker, args=(w,)) for w in range(16)]
    for th in threads:
        th.start()
    for th in threads:
        th.join()
    assert not errors, errors[:3]
    t.destroy()
    ft.destroy()


def test_frame_calls_with_rejected_queries(tb, oracle, small_case):
    """Frame-sized host calls are pipelined by the device's worker (api_combine.inc): the tensor-core engine's rejected
    queries are handled after the batch.  100 copies of one tile put 400 rows at the same distance from their queries -
    more than a candidate list holds - so those queries are rejected and recomputed on the exact engine; 70 000
    queries per call also take the coherence order (reject positions translated back).  Result: the exact engine's."""
    o = oracle
    tiles = np.concatenate([small_case["tiles"], np.repeat(small_case["tiles"][7:8], 100, 0)])
    pal = small_case["pal"]
    rgb_a, s = o.gen_queries(o.BASE_SEED + 5, 35000, tiles[7:8], pal)      # all near the repeated tile (or pure random)
    rgb_b, s = o.gen_queries(s, 35000, tiles, pal)
    rgb = np.concatenate([rgb_a, rgb_b])[np.random.default_rng(1).permutation(70000)]
    with tb.KDTree.from_tileset(tiles, pal) as t:
        tb.set_engine(tb.ENGINE_SIMT)
        want = t.frametiling_search(rgb)
        tb.set_engine(tb.ENGINE_AUTO)
        for n in (3000, 70000):
            tb.reset_stats()
            gi, ge = t.frametiling_search(rgb[:n])
            st = tb.stats()
            assert st["coarse_launches"] > 0 and st["fallback_queries"] > n // 8, st
            assert np.array_equal(gi, want[0][:n]) and np.array_equal(ge.view(np.uint32), want[1][:n].view(np.uint32))


def test_search_edge_cases(tb, oracle):
    rng = np.random.default_rng(0)
    # ragged sizes: n not a multiple of the tile, nq not a multiple of the block, dd not a multiple of 4
    for n, nq, dd in [(1, 1, 192), (3, 5, 64), (65, 130, 192), (1000, 77, 50), (129, 64, 7)]:
        ds = rng.standard_normal((n, dd)).astype(np.float32)
        q = rng.standard_normal((nq, dd)).astype(np.float32)
        k = min(n, 4)
        with tb.KDTree.from_rows(ds) as t:
            idx, err = t.search_batch(q, k=k)
        oi, od = oracle.knn_f32(ds, q, k)
        assert np.array_equal(idx, oi) and np.array_equal(err, od), (n, nq, dd)
    # k > n: missing neighbours are index -1 / +inf
    ds = rng.standard_normal((2, 16)).astype(np.float32)
    with tb.KDTree.from_rows(ds) as t:
        idx, err = t.search_batch(ds[:1], k=4)
        assert list(idx[0, 2:]) == [-1, -1] and np.all(np.isinf(err[0, 2:]))
        assert t.search_batch(np.zeros((0, 16), np.float32))[0].shape == (0, 1)
    # exact duplicates: lowest index wins
    ds = np.repeat(rng.standard_normal((1, 192)).astype(np.float32), 5, axis=0)
    with tb.KDTree.from_rows(ds) as t:
        idx, err = t.search_batch(ds[:1], k=3)
        assert list(idx[0]) == [0, 1, 2] and np.all(err == 0)
    # empty dataset
    with tb.KDTree.from_rows(np.zeros((0, 192), np.float32)) as t:
        idx, err = t.search_batch(np.zeros((3, 192), np.float32))
        assert np.all(idx == -1)


def test_frametiling_fused(tb, oracle, small_case):
    ds = oracle.psyv_tileset_f32(small_case["tiles"], small_case["pal"])
    q = oracle.psyv_queries_f32(small_case["rgb"])
    with tb.KDTree.from_tileset(small_case["tiles"], small_case["pal"]) as t:
        idx, err = t.frametiling_search(small_case["rgb"])
    ingap = _check_search(oracle, ds, q, idx, err, 1)
    assert ingap < 0.01 * len(q)
    # row -> (tile, attrs) mapping of DoPsyV (main.pas:3908-3910): planted queries come back
    # as their tile with err << 1 (noise only)
    assert np.median(err) < 0.1


def test_frametiling_ragged_sizes_all_paths(tb, oracle, small_case):
    """Every query count around the engine / padding / staging boundaries, through the host call (pageable
    and pinned memory) and the device-pointer call, against the exact engine on the same descriptors."""
    import torch
    rgb = small_case["rgb"]
    ds = oracle.psyv_tileset_f32(small_case["tiles"], small_case["pal"])
    q = oracle.psyv_queries_f32(rgb)
    oi, od = oracle.knn_f32(ds, q, 1)
    with tb.KDTree.from_tileset(small_case["tiles"], small_case["pal"]) as t:
        for n in (1, 2, 63, 64, 65, 127, 128, 255, 256, 257, 511, 513, 1000, 1499, 1500):
            for off in (0, 3):
                if off + n > len(rgb):
                    continue
                sl = slice(off, off + n)
                gi, ge = t.frametiling_search(rgb[sl])                      # pageable host memory
                assert np.array_equal(ge.view(np.uint32), od[sl, 0].view(np.uint32)), (n, off)
                same = gi == oi[sl, 0]
                assert np.all(same | (ge == 0) | np.isclose(ge, od[sl, 0])), (n, off)
                pin = torch.from_numpy(np.ascontiguousarray(rgb[sl])).pin_memory()
                pi, pe = t.frametiling_search(pin.numpy())                  # pinned host memory
                assert np.array_equal(pi, gi) and np.array_equal(pe, ge), (n, off)
                d = torch.from_numpy(np.ascontiguousarray(rgb[sl])).cuda()
                di = torch.empty(n, dtype=torch.int32, device="cuda")
                de = torch.empty(n, dtype=torch.float32, device="cuda")
                t.frametiling_search(d, out_idx=di, out_err=de)             # device pointers
                torch.cuda.synchronize()
                assert np.array_equal(di.cpu().numpy(), gi) and np.array_equal(de.cpu().numpy(), ge), (n, off)


def test_kmodes_assign_bit_exact(tb, oracle, small_case):
    o = oracle
    tiles, _ = o.gen_tileset(o.BASE_SEED + 9, 6000)
    X = o.kmodes_rows(tiles)
    cent = X[::23][:200].copy()
    for sem in (tb.KM_ASM, tb.KM_GENERIC):
        clust, dis = tb.kmodes_assign(X, cent, sem)
        wc, wd = o.kmodes_assign(X, cent, sem)
        assert np.array_equal(clust, wc) and np.array_equal(dis, wd), sem
    # arbitrary bytes (wrap-around of pabsb/psubb), duplicate centroids (last index wins), k not a tile multiple
    rng = np.random.default_rng(4)
    X = rng.integers(0, 256, (777, 80)).astype(np.uint8)
    cent = rng.integers(0, 256, (513, 80)).astype(np.uint8)
    cent[500] = cent[3]
    X[5] = cent[3]
    for sem in (tb.KM_ASM, tb.KM_GENERIC):
        clust, dis = tb.kmodes_assign(X, cent, sem)
        wc, wd = o.kmodes_assign(X, cent, sem)
        assert np.array_equal(clust, wc) and np.array_equal(dis, wd), sem
        assert clust[5] == 500 and dis[5] == 0
    # nibble fast path and byte path mixed inside one call: narrow rows, one wide centroid in the third
    # 256-centroid tile, a few wide points in one block; ties between narrow duplicates
    Xn = rng.integers(0, 16, (3000, 80)).astype(np.uint8)
    cn = rng.integers(0, 16, (700, 80)).astype(np.uint8)
    cn[600, 7] = 200
    cn[650] = cn[10]
    Xn[40] = cn[10]
    Xn[1500, 79] = 16
    for sem in (tb.KM_ASM, tb.KM_GENERIC):
        clust, dis = tb.kmodes_assign(Xn, cn, sem)
        wc, wd = o.kmodes_assign(Xn, cn, sem)
        assert np.array_equal(clust, wc) and np.array_equal(dis, wd), sem
        assert clust[40] == 650 and dis[40] == 0
    # single centroid, empty inputs
    clust, dis = tb.kmodes_assign(X[:10], cent[:1])
    assert np.all(clust == 0)
    assert tb.kmodes_assign(X[:0], cent)[0].shape == (0,)
    clust, dis = tb.kmodes_assign(X[:4], cent[:0])
    assert np.all(clust == -1) and np.all(dis == np.uint64(2 ** 64 - 1))


def test_kmodes_bin_sized_calls_split_centroids(tb, oracle):
    """KModesIter works in bins of 960 points (kmodes.pas:847): few points, many centroids.  The launch splits
    the centroid range over the grid; duplicates of a centroid in different ranges must still resolve to the
    LAST index (kmodes.pas:414-415)."""
    o = oracle
    tiles, _ = o.gen_tileset(o.BASE_SEED + 77, 960 + 5000)
    rows = o.kmodes_rows(tiles)
    X, cent = rows[:960], rows[960:].copy()
    cent[4900] = cent[10]     # same centroid far apart: the later one must win every tie
    cent[2600] = cent[10]
    cent[3000] = X[5]         # exact hits (dis 0) duplicated across ranges
    cent[17] = X[5]
    for sem in (tb.KM_ASM, tb.KM_GENERIC):
        for n, k in ((960, 5000), (960, 513), (37, 5000), (1, 4097)):
            clust, dis = tb.kmodes_assign(X[:n], cent[:k], sem)
            wc, wd = o.kmodes_assign(X[:n], cent[:k], sem)
            assert np.array_equal(clust, wc) and np.array_equal(dis, wd), (sem, n, k)
    clust, _ = tb.kmodes_assign(X, cent)
    assert clust[5] == 3000 and not np.any(clust == 10) and not np.any(clust == 2600)


def test_kmodes_update_min_distance(tb, oracle):
    rng = np.random.default_rng(6)
    X = rng.integers(0, 16, (5000, 80)).astype(np.uint8)
    md = rng.integers(0, 60000, 5000).astype(np.uint64)
    md[::3] = np.uint64(2 ** 64 - 1)
    for sem in (tb.KM_ASM, tb.KM_GENERIC):
        got = tb.kmodes_update_min_distance(X, X[17], md, sem)
        assert np.array_equal(got, oracle.update_min_distance(X, X[17], md, sem))


def test_kmodes_farthest_first_and_representatives(tb, oracle):
    o = oracle
    tiles, _ = o.gen_tileset(o.BASE_SEED + 31, 3000)
    X = o.kmodes_rows(tiles)
    X[100] = X[7]      # duplicate rows: ties in the arg-max (last index wins, kmodes.pas:765-770)
    X[2999] = X[7]
    for sem in (tb.KM_ASM, tb.KM_GENERIC):
        cent, chosen = tb.kmodes_farthest_first(X, 40, 17, sem)
        wc, wch = o.farthest_first(X, 40, 17, sem)
        assert np.array_equal(chosen, wch) and np.array_equal(cent, wc), sem
        labels, _ = tb.kmodes_assign(X, cent, sem)
        rep, cnt = tb.kmodes_representatives(X, labels, cent, sem)
        wr, wn = o.kmodes_representatives(X, labels, cent, sem)
        assert np.array_equal(rep, wr) and np.array_equal(cnt, wn), sem
    # arbitrary bytes, k = npoints (every point chosen exactly once), empty cluster
    rng = np.random.default_rng(8)
    Xr = rng.integers(0, 256, (50, 80)).astype(np.uint8)
    cent, chosen = tb.kmodes_farthest_first(Xr, 50, 0)
    wc, wch = o.farthest_first(Xr, 50, 0)
    assert np.array_equal(chosen, wch) and sorted(chosen.tolist()) == list(range(50))
    lab = np.zeros(50, np.int32)
    rep, cnt = tb.kmodes_representatives(Xr, lab, Xr[:3])
    wr, wn = o.kmodes_representatives(Xr, lab, Xr[:3])
    assert np.array_equal(rep, wr) and rep[1] == -1 and cnt.tolist() == [50, 0, 0]
    from tiler_b200._lib import TMError
    with pytest.raises(TMError):
        tb.kmodes_farthest_first(Xr, 51, 0)
    with pytest.raises(TMError):
        tb.kmodes_representatives(Xr, lab + 5, Xr[:3])


def test_reindex_matches_oracle(tb, oracle):
    import torch
    rng = np.random.default_rng(12)
    ntiles = 5000
    # zipf-like use counts with many ties and many unused tiles
    tm = np.minimum(rng.zipf(1.3, 200000) - 1, ntiles - 1).astype(np.int32)
    tm[rng.integers(0, len(tm), 3000)] = rng.integers(0, ntiles, 3000)
    new, cnt, imap, order = tb.reindex(tm, ntiles)
    wn, wc, wi, wo = oracle.reindex(tm, ntiles)
    assert np.array_equal(new, wn) and np.array_equal(cnt, wc) and np.array_equal(imap, wi) and np.array_equal(order, wo)
    # properties at a larger size: counts are non-increasing along the new order, the remap preserves use counts
    assert np.all(np.diff(cnt[order]) <= 0)
    assert np.array_equal(np.bincount(new, minlength=len(order)), cnt[order])
    # device-pointer variant rewrites in place
    d = torch.from_numpy(tm).cuda()
    dc, di, do, nact = tb.reindex(d, ntiles)
    torch.cuda.synchronize()
    assert nact == len(wo) and np.array_equal(d.cpu().numpy(), wn) and np.array_equal(di.cpu().numpy(), wi)
    # empty tilemap, out-of-range index
    e = tb.reindex(np.zeros(0, np.int32), 4)
    assert len(e[0]) == 0 and e[1].tolist() == [0, 0, 0, 0] and len(e[3]) == 0
    from tiler_b200._lib import TMError
    with pytest.raises(TMError):
        tb.reindex(np.array([0, 9], np.int32), 4)


def test_kmodes_rows_and_binned_assign(tb, oracle):
    """K1 rows on the device and ReloadPreviousTiling's binned arg-min (main.pas:4377-4409)."""
    o = oracle
    for ps in (16, 8, 64):
        tiles, _ = o.gen_tileset(o.BASE_SEED + 40 + ps, 1500, ps)
        rows, signi = tb.kmodes_rows(tiles, ps)
        for t in range(0, 1500, 7):
            wr, ws = o.kmodes_row(tiles[t], ps)
            assert np.array_equal(rows[t], wr) and signi[t] == ws, (ps, t)
    tiles, _ = o.gen_tileset(o.BASE_SEED + 41, 4000)
    saved, _ = o.gen_tileset(o.BASE_SEED + 42, 900)
    X, xs = tb.kmodes_rows(tiles)
    Cn, cs = tb.kmodes_rows(saved)
    Cn[5] = Cn[300]; cs[5] = cs[300]  # duplicate saved tiles in one bin: the later one wins
    clust, dis = tb.kmodes_assign_binned(X, xs, Cn, cs, 65)
    # restated with the oracle's arg-min bin by bin
    for i in range(0, 4000, 13):
        members = np.where(cs == xs[i])[0]
        if len(members):
            j, d = o.min_dissim(Cn[members], X[i])
            j = members[j]
        else:
            j, d = o.min_dissim(Cn, X[i])
        assert clust[i] == j and dis[i] == d, i
    assert np.any(~np.isin(xs, cs))  # some points fall into bins without saved tiles (the all-tiles branch ran)


def _smoothing_case(o, seed, nframes, npos, ntiles, nkf, npal, change=0.35):
    """Tilemaps that mostly persist from frame to frame, with near-duplicate tiles so that the
    threshold test has hits, misses and both replacement directions."""
    rng = np.random.default_rng(seed)
    tiles, _ = o.gen_tileset(o.BASE_SEED + seed, ntiles)
    near = tiles[: ntiles // 2].copy()  # second half: one-pixel variants of the first half
    for t in range(near.shape[0]):
        near[t, rng.integers(64)] = rng.integers(16)
    tiles = np.concatenate([tiles[: ntiles - near.shape[0]], near])
    pals = np.stack([np.stack([o.gen_palette(o.BASE_SEED + seed + 7 * k + p)[0] for p in range(npal)]) for k in range(nkf)])
    frame_kf = np.sort(rng.integers(0, nkf, nframes)).astype(np.int32)
    tile = np.empty((nframes, npos), np.int32)
    pal = np.empty((nframes, npos), np.int32)
    attrs = np.empty((nframes, npos), np.uint8)
    tile[0] = rng.integers(0, ntiles, npos); pal[0] = rng.integers(0, npal, npos); attrs[0] = rng.integers(0, 8, npos)
    half = ntiles - near.shape[0]
    for f in range(1, nframes):
        ch = rng.random(npos) < change
        flip = rng.random(npos) < 0.5  # change to the near-duplicate (small distance) or to a random tile
        twin = np.where(tile[f - 1] < near.shape[0], tile[f - 1] + half, tile[f - 1] - half) % ntiles
        tile[f] = np.where(ch, np.where(flip, twin, rng.integers(0, ntiles, npos)), tile[f - 1])
        pal[f] = np.where(ch & ~flip, rng.integers(0, npal, npos), pal[f - 1])
        attrs[f] = np.where(ch & ~flip, rng.integers(0, 8, npos), attrs[f - 1]).astype(np.uint8)
    return tiles, pals, frame_kf, tile, pal, attrs


def test_temporal_smoothing_matches_oracle(tb, oracle):
    """btnSmoothClick + DoTemporalSmoothing (main.pas:1336-1370, 4071-4119): bit-exact decisions."""
    o = oracle
    tiles, pals, frame_kf, tile, pal, attrs = _smoothing_case(o, 61, 24, 700, 400, 3, 4)
    hits = []
    for strength in (0.0, 0.01, 0.03, 0.1, 1.0):
        want = o.temporal_smoothing(tiles, pals, frame_kf, tile, pal, attrs, strength)
        got = tb.temporal_smoothing(tiles, pals, frame_kf, tile, pal, attrs, strength)
        for w, g, name in zip(want, got, ("src_frame", "tile", "pal", "attrs")):
            assert np.array_equal(w, g), (strength, name, int((w != g).sum()))
        hits.append(int(((want[3] & 4) != 0).sum()))
    assert hits[0] < hits[2] < hits[4]  # the thresholds separate the cases
    # single frame / single keyframe boundary everywhere: nothing changes
    src, t2, p2, a2 = tb.temporal_smoothing(tiles, pals, np.arange(3, dtype=np.int32), tile[:3], pal[:3], attrs[:3], 1.0)
    assert np.array_equal(t2, tile[:3]) and np.array_equal(a2, attrs[:3]) and np.array_equal(src, np.repeat(np.arange(3), 700).reshape(3, 700))
    with pytest.raises(Exception):
        tb.temporal_smoothing(tiles, pals, frame_kf, tile + 400, pal, attrs, 0.1)


def test_device_pointer_entry_points(tb, oracle, small_case):
    import torch
    ds = oracle.psyv_tileset_f32(small_case["tiles"], small_case["pal"])
    q = oracle.psyv_queries_f32(small_case["rgb"])
    d_rgb = torch.from_numpy(small_case["rgb"]).cuda()
    d_idx = torch.empty(len(q), dtype=torch.int32, device="cuda")
    d_err = torch.empty(len(q), dtype=torch.float32, device="cuda")
    t = tb.KDTree.from_tileset(torch.from_numpy(small_case["tiles"]).cuda(),
                               torch.from_numpy(small_case["pal"]).cuda().reshape(1, -1))
    t.frametiling_search(d_rgb, out_idx=d_idx, out_err=d_err)
    torch.cuda.synchronize()
    _check_search(oracle, ds, q, d_idx.cpu().numpy(), d_err.cpu().numpy(), 1)
    d_desc = torch.empty((len(q), 192), dtype=torch.float32, device="cuda")
    tb.psyv_queries(d_rgb, out=d_desc)
    torch.cuda.synchronize()
    assert np.array_equal(d_desc.cpu().numpy(), q)
    st = tb.stats()
    assert st["kernel_launches"] > 0 and st["queries"] >= len(q)
    t.destroy()
