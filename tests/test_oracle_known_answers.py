"""Known-answer tests that pin the CPU oracle to the reference's arithmetic.

The reference has no tests or golden vectors for this path (SURVEY.md §4, §8c:
"parity unpinned"), so the anchors are identities derived by hand from the cited
Pascal code: Haar orthonormality, DCT scaling, constant-tile and colour constants,
mirror consistency, the asm-vs-generic dissimilarity values and tie-breaks.
"""
import numpy as np
import pytest


def test_colour_constants(oracle):
    # main.pas:2656-2679 with cRedMul/cGreenMul/cBlueMul = 2126/7152/722 (main.pas:25-27)
    assert oracle.rgb_to_yuv(255, 255, 255) == pytest.approx((1.0, 0.0, 0.0), abs=1e-15)
    y, u, v = oracle.rgb_to_yuv(255, 0, 0)
    assert y == pytest.approx(0.2126) and v == pytest.approx(0.5)
    assert u == pytest.approx(-0.2126 * 0.5 / 0.9278)
    y, u, v = oracle.rgb_to_yuv(0, 0, 255)
    assert y == pytest.approx(0.0722) and u == pytest.approx(0.5)


def test_gamma_lut(oracle):
    # gGamma = (2.0, 0.6) main.pas:586
    y, _, _ = oracle.rgb_to_yuv(128, 128, 128, gamma=0)
    assert y == pytest.approx((128 / 255.0) ** 2.0)
    y, _, _ = oracle.rgb_to_yuv(128, 128, 128, gamma=1)
    assert y == pytest.approx((128 / 255.0) ** 0.6)


def test_constant_tile(oracle):
    col = (30 << 16) | (200 << 8) | 90  # $00BBGGRR
    y, u, v = oracle.rgb_to_yuv(90, 200, 30)
    f = oracle.psyv(rgb=np.full(64, col))
    assert f[0] == pytest.approx(8 * y) and f[64] == pytest.approx(8 * u) and f[128] == pytest.approx(8 * v)
    mask = np.ones(192, bool)
    mask[[0, 64, 128]] = False
    assert np.abs(f[mask]).max() < 1e-14
    f = oracle.psyv(rgb=np.full(64, col), use_wavelets=False)
    # DC term only: the reference's LUT uses cos((x+0.5)*u*PI/16) (main.pas:620), i.e. HALF the DCT-II
    # frequency, so its "DCT" basis is not orthogonal and a constant tile has non-zero AC terms.
    assert f[0] == pytest.approx(32 * y) and f[64] == pytest.approx(32 * u) and f[128] == pytest.approx(32 * v)


def _planes(oracle, rgb):
    out = np.empty((3, 64))
    for i, c in enumerate(rgb):
        out[:, i] = oracle.rgb_to_yuv(c & 0xff, (c >> 8) & 0xff, (c >> 16) & 0xff)
    return out.reshape(-1)


def _dct_numpy(planes):
    # literal restatement of main.pas:613-623 + 3075-3174 in numpy (independent of oracle/psyv.c)
    x = np.arange(8)
    c = np.cos((x[None, :] + 0.5) * x[:, None] * np.pi / 16.0)  # c[u, x]
    ratio = np.where(x == 0, np.sqrt(0.5), 1.0)
    out = np.empty((3, 8, 8))
    for p in range(3):
        cp = planes[p].reshape(8, 8)  # [y, x]
        out[p] = np.einsum("yx,ux,vy->vu", cp, c, c) * ratio[:, None] * ratio[None, :]
    return out.reshape(-1)


def test_haar_isometry_and_dct_formula(oracle):
    rng = np.random.default_rng(1)
    a = rng.integers(0, 1 << 24, 64).astype(np.int32)
    b = rng.integers(0, 1 << 24, 64).astype(np.int32)
    pix = np.sum((_planes(oracle, a) - _planes(oracle, b)) ** 2)
    haar = np.sum((oracle.psyv(rgb=a) - oracle.psyv(rgb=b)) ** 2)
    assert haar == pytest.approx(pix, rel=1e-12)        # orthonormal per plane
    dct = oracle.psyv(rgb=a, use_wavelets=False)
    assert np.allclose(dct, _dct_numpy(_planes(oracle, a).reshape(3, 64)), rtol=1e-12, atol=1e-12)


def test_haar_layout(oracle):
    # a single bright pixel at (y=0, x=0): level-1 HL/LH/HH of the top-left 2x2 block are +0.5*Y
    rgb = np.zeros(64, np.int32)
    rgb[0] = 0x00FFFFFF
    f = oracle.psyv(rgb=rgb)[:64].reshape(8, 8)
    assert f[0, 4] == pytest.approx(0.5) and f[4, 0] == pytest.approx(0.5) and f[4, 4] == pytest.approx(0.5)
    assert f[0, 0] == pytest.approx(1 / 8.0)  # DC after three levels: 1 * (1/2)^3


def test_mirror_consistency(oracle):
    rng = np.random.default_rng(2)
    pal = rng.integers(0, 1 << 24, 16).astype(np.int32)
    t = rng.integers(0, 16, (8, 8)).astype(np.uint8)
    for hm in (0, 1):
        for vm in (0, 1):
            phys = t.copy()
            if hm:
                phys = oracle.hmirror(phys)
            if vm:
                phys = oracle.vmirror(phys)
            a = oracle.psyv(palpix=phys, pal=pal)
            b = oracle.psyv(palpix=t, pal=pal, hmirror=hm, vmirror=vm)
            assert np.array_equal(a, b)
    assert np.array_equal(oracle.hmirror(t), t[:, ::-1])
    assert np.array_equal(oracle.vmirror(t), t[::-1, :])


def test_global_rows_order(oracle):
    rng = np.random.default_rng(3)
    t = rng.integers(0, 16, (5, 64)).astype(np.uint8)
    rows, tr_tile, tr_attrs = oracle.global_rows_f32(t)
    assert list(tr_attrs[:4]) == [0, 1, 3, 2]  # main.pas:3753 with order none,H,HV,V
    assert list(tr_tile[:8]) == [0, 0, 0, 0, 1, 1, 1, 1]
    t0 = t[0].reshape(8, 8)
    assert np.array_equal(rows[0], t0.reshape(-1))
    assert np.array_equal(rows[1], t0[:, ::-1].reshape(-1))
    assert np.array_equal(rows[2], t0[::-1, ::-1].reshape(-1))
    assert np.array_equal(rows[3], t0[::-1, :].reshape(-1))


def test_prepare_tile_mirrors(oracle):
    t = np.zeros((8, 8), np.uint8)
    t[6, 6] = 9  # brightest quadrant bottom-right -> H and V
    out, f = oracle.prepare_tile_mirrors(t)
    assert f == 3 and out[1, 1] == 9
    out, f = oracle.prepare_tile_mirrors(np.zeros((8, 8), np.uint8))
    assert f == 0  # first wins ties (strict > main.pas:4059)


def test_dissim_known_values(oracle):
    a = np.zeros(80, np.uint8)
    for byte, want_asm, want_gen in [(5, 2048, 2051), (1, 2048 + 768, 2051), (20, 2051, 2051),
                                     (0, 2051, 2051), (9, 2048 + 768, 2051), (8, 2051, 2051), (79, 2051, 2051)]:
        b = a.copy()
        b[byte] = 3
        assert oracle.dissim(a, b, oracle.ASM) == want_asm, byte
        assert oracle.dissim(b, a, oracle.ASM) == want_asm, byte
        assert oracle.dissim(a, b, oracle.GENERIC) == want_gen, byte


def test_dissim_wrapping_bytes(oracle):
    # psubb/pabsb wrap for byte differences > 127 (kmodes.pas:363-365)
    a = np.zeros(80, np.uint8)
    b = a.copy()
    b[0] = 200  # 200 - 0 = 0xC8 = -56 -> 56
    assert oracle.dissim(a, b, oracle.ASM) == 2048 + 56
    b[0] = 128
    assert oracle.dissim(a, b, oracle.ASM) == 2048 + 128
    assert oracle.dissim(a, b, oracle.GENERIC) == 2048 + 128


def test_argmin_last_index_wins(oracle):
    item = np.zeros(80, np.uint8)
    rows = np.zeros((4, 80), np.uint8)
    rows[0, 30] = 2
    rows[2, 40] = 2   # same dissimilarity as row 0? no: rows 1 and 3 are exact (0)
    idx, best = oracle.min_dissim(rows, item)
    assert (idx, best) == (3, 0)
    rows[1, 3] = 1
    rows[3, 3] = 1
    idx, best = oracle.min_dissim(rows, item)
    assert best == 2048 and idx == 3
    assert oracle.min_dissim(rows[:0], item) == (-1, 2 ** 64 - 1)


def test_update_min_distance_strict(oracle):
    X = np.zeros((3, 80), np.uint8)
    X[1, 20] = 4
    md = np.array([5, 5, 2 ** 64 - 1], np.uint64)
    out = oracle.update_min_distance(X, np.zeros(80, np.uint8), md)
    assert list(out) == [0, 5, 0]


def test_farthest_first(oracle):
    rng = np.random.default_rng(5)
    X = rng.integers(0, 16, (200, 80)).astype(np.uint8)
    X[:, 64:] = rng.integers(0, 2, (200, 16))
    cent, chosen = oracle.farthest_first(X, 5, 7)
    assert chosen[0] == 7 and len(set(chosen)) == 5
    # second centre = LAST index among the farthest points from X[7] (>= test kmodes.pas:767)
    d = np.array([oracle.dissim(X[7], X[i]) for i in range(200)])
    d[7] = 0
    assert chosen[1] == np.flatnonzero(d == d.max())[-1]
    assert np.array_equal(cent[1], X[chosen[1]])


def test_kmodes_row(oracle):
    t = np.zeros(64, np.uint8)
    t[:10] = 15
    row, res = oracle.kmodes_row(t, 16)
    assert np.array_equal(row[:64], t)
    assert row[64] == 1 and row[64 + 15] == 1 and row[65:79].sum() == 0  # counts 54 and 10 exceed S div 16 = 1
    assert res == 10


def test_randint_matches_reference_lcg(oracle):
    # Seed := Seed * $08088405 + 1; Result := (Seed * Range) shr 32  (kmodes.pas:82-86)
    seed = 0x42381337
    for rng in (10, 1 << 24, 7):
        want_seed = (seed * 0x08088405 + 1) & 0xffffffff
        want = (want_seed * rng) >> 32
        got, seed = oracle.randint(rng, seed)
        assert (got, seed) == (want, want_seed)


def test_knn_paths_agree(oracle, small_case):
    ds = oracle.psyv_tileset_f32(small_case["tiles"], small_case["pal"])
    q = oracle.psyv_queries_f32(small_case["rgb"])
    i64_, d64, nxt = oracle.knn_f64(ds, q, 1)
    i32_, d32 = oracle.knn_f32(ds, q, 1)
    ib, db = oracle.nn1_blocked_f32(ds, q)
    gap_ok = (nxt - d64[:, 0]) > 1e-5 * np.maximum(d64[:, 0], 1e-12)
    assert np.array_equal(i64_[gap_ok, 0], i32_[gap_ok, 0])
    assert np.array_equal(i64_[gap_ok, 0], ib[gap_ok])
    assert np.allclose(d32[:, 0], d64[:, 0], rtol=1e-5, atol=1e-9)
    # k = 8 ascending
    i8, d8, _ = oracle.knn_f64(ds, q[:50], 8)
    assert np.all(np.diff(d8, axis=1) >= 0) and np.array_equal(i8[:, 0], i64_[:50, 0])


def test_planted_queries_recovered(oracle):
    pal, s = oracle.gen_palette(oracle.BASE_SEED + 1)
    tiles, s = oracle.gen_tileset(s, 300)
    tiles = np.unique(tiles, axis=0)
    ds = oracle.psyv_tileset_f32(tiles, pal)
    # query = exact render of dataset row r (tile 17, vmir=1, hmir=0 -> row 17*4+2) => err 0
    t = tiles[17].reshape(8, 8)[::-1, :]
    rgb = pal[t.reshape(-1)]
    q = oracle.psyv_queries_f32(rgb[None])
    idx, d, _ = oracle.knn_f64(ds, q, 1)
    assert d[0, 0] == 0.0
    assert np.array_equal(ds[idx[0, 0]], ds[17 * 4 + 2])


# ---- ANN-style kd-tree restatement vs the brute-force restatement (two independently written paths)
def test_kdtree_matches_bruteforce_k1_and_k8(oracle, small_case):
    o = oracle
    ds = o.psyv_tileset_f32(small_case["tiles"], small_case["pal"])
    q = o.psyv_queries_f32(small_case["rgb"])
    kd = o.KDTreeRef(ds)
    i1, d1, st = kd.search(q, 1, with_stats=True)
    fi, fd = o.knn_f32(ds, q, 1)
    assert np.array_equal(d1, fd)  # same float32 sequential sum => bit-identical distances
    same = i1[:, 0] == fi[:, 0]
    # a differing index is only legal among exact float32 ties (ANN's tie order is traversal order)
    assert np.array_equal(d1[~same], fd[~same])
    assert st[0] < len(q) * len(ds)  # pruning happened at all
    i8, d8 = kd.search(q[:200], 8)
    f8i, f8d = o.knn_f32(ds, q[:200], 8)
    assert np.array_equal(d8, f8d)
    kd.close()


def test_kdtree_edge_cases(oracle):
    o = oracle
    rng = np.random.default_rng(3)
    # duplicates, n < k, n = 1, stage-1 shaped integer rows (D = 64)
    rows = rng.integers(0, 16, (37, 64)).astype(np.float32)
    rows[5] = rows[20]
    kd = o.KDTreeRef(rows)
    i, d = kd.search(rows[20], 2)
    assert set(i[0]) == {5, 20} and d[0, 0] == 0 and d[0, 1] == 0
    i, d = kd.search(rows[:4], 8)
    fi, fd = o.knn_f32(rows, rows[:4], 8)
    assert np.array_equal(d, fd)
    kd.close()
    one = o.KDTreeRef(rows[:1])
    i, d = one.search(rows[3], 3)
    assert i[0, 0] == 0 and i[0, 1] == -1 and i[0, 2] == -1
    one.close()


# ---- PrepareFrameTiling restatement (main.pas:3791-3967): hand-checkable cases
def test_prepare_frame_tiling_known_cases(oracle):
    o = oracle
    pal, s = o.gen_palette(o.BASE_SEED + 3)
    pals = np.stack([pal, pal[::-1].copy()])
    tiles = np.zeros((3, 64), np.uint8)
    tiles[0] = 5                                  # flat tile: its 4 stage-1 rows are identical
    tiles[1] = np.arange(64) % 16                 # no symmetry
    tiles[2] = 9                                  # another flat tile, distance 64 * 16 from tile 0
    # one tilemap item (pal 1, tile 0), ftFast: hits = the 4 identical rows of tile 0 (err 0, only the first kept:
    # row 0 -> attrs 0), then the 4 identical rows of tile 2 (err 1024, first kept), 8 hits in total
    rows, tt, tp, ta, used = o.prepare_frame_tiling(tiles, pals, [1], [0], quality=0)
    assert used[0].sum() == 0                      # ftFast: own palette only (main.pas:3842)
    assert used[1].tolist() == [[1, 0, 0, 0], [0, 0, 0, 0], [1, 0, 0, 0]]
    assert tt.tolist() == [0, 2] and tp.tolist() == [1, 1] and ta.tolist() == [0, 0]
    # rows are the palette-rendered descriptors in DoPsyV order
    want = o.psyv(palpix=tiles[0], pal=pals[1]).astype(np.float32)
    assert np.array_equal(rows[0], want)
    # ftSlow marks every palette (main.pas:3848-3850); duplicates of the item are ignored (3814-3821)
    rows2, tt2, tp2, ta2, used2 = o.prepare_frame_tiling(tiles, pals, [1, 1, 1], [0, 0, 0], quality=2)
    assert np.array_equal(used2[0], used2[1]) and used2[0].sum() == 2 and tp2.tolist() == [0, 0, 1, 1]
    # ftMedium: palette 0 is selected for an item of palette 1 only if their centroid distance < tol * max
    cen = np.zeros((2, 192))
    cen[1, 0] = 2.0                                # squared distance 4 = the maximum
    _, _, tpm, _, um = o.prepare_frame_tiling(tiles, pals, [1], [0], quality=1, pal_tol=0.05, pal_centroids=cen)
    assert um[0].sum() == 0 and um[1].sum() == 2   # 4 < 0.05 * 4 is false; own palette: 0 < 0.2 is true
    _, _, _, _, um2 = o.prepare_frame_tiling(tiles, pals, [1], [0], quality=1, pal_tol=1.5, pal_centroids=cen)
    assert um2[0].sum() == 2
    # tile flags flip the rendered orientation, not the attrs (main.pas:3910-3912)
    r3, t3, _, a3, _ = o.prepare_frame_tiling(tiles, pals, [0], [1], quality=0, tile_flags=[0, 1, 0])
    i = int(np.where((a3 == 0) & (t3 == 1))[0][0])
    assert np.array_equal(r3[i], o.psyv(palpix=tiles[1], pal=pals[0], hmirror=True).astype(np.float32))


def test_reindex_known_case(oracle):
    # tiles 0..5; uses: tile 4 x3, tile 1 x3, tile 2 x1, tile 5 x2; tiles 0 and 3 unused
    tm = np.array([4, 1, 4, 5, 2, 1, 4, 1, 5], np.int32)
    new, cnt, imap, order = oracle.reindex(tm, 6)
    assert cnt.tolist() == [0, 3, 1, 0, 3, 2]
    assert order.tolist() == [1, 4, 5, 2]          # count desc, then old index asc (main.pas:4472-4481)
    assert imap.tolist() == [-1, 0, 3, -1, 1, 2]
    assert new.tolist() == [1, 0, 1, 2, 3, 0, 1, 0, 2]


def test_temporal_smoothing_known_case(oracle):
    """DoTemporalSmoothing (main.pas:4071-4119) on a hand-built chain of one position."""
    o = oracle
    flat = np.zeros((4, 64), np.uint8)
    flat[1] = 1                     # tile 1: other flat colour
    flat[2] = 0; flat[2, 0] = 1     # tile 2: tile 0 with one pixel changed
    flat[3] = 0
    pal = np.zeros((2, 1, 16), np.int32)
    pal[:, 0, 0] = 0x000000
    pal[:, 0, 1] = 0xFFFFFF
    # flat white tile: luma DC term 64 * 1.0 * cQ/sqrt(16) * 0.5 = 32, chroma planes zero.  (gDCTLut's angle is
    # (x + 0.5) * u * PI / 16, main.pas:621 - half the DCT-II frequency - so a flat tile has AC terms too.)
    d = o.psyv(palpix=flat[1], pal=pal[0, 0], use_wavelets=False, qweighting=True)
    assert abs(d[0] - 32.0) < 1e-12 and np.abs(d[64:]).max() < 1e-12
    want1 = sum(np.cos((x + 0.5) * np.pi / 16.0) for x in range(8)) * 8 * (4.0 / np.sqrt(11.0)) * np.sqrt(0.5)
    assert abs(d[1] - want1) < 1e-9
    # frames: tile 2, tile 0, tile 0 | keyframe change | tile 1, tile 3 (different tile, same pixels as tile 0)
    kf = np.array([0, 0, 0, 1, 1], np.int32)
    tile = np.array([[2], [0], [0], [1], [3]], np.int32)
    z = np.zeros((5, 1), np.int32)
    at = np.zeros((5, 1), np.uint8)
    d2 = o.psyv(palpix=flat[2], pal=pal[0, 0], use_wavelets=False, qweighting=True)
    d0 = o.psyv(palpix=flat[0], pal=pal[0, 0], use_wavelets=False, qweighting=True)
    cmp20 = np.sqrt(((d0 - d2) ** 2).sum() / 192.0)
    # below the threshold of the 2 -> 0 pair: only identical items are merged
    src, t, p, a = o.temporal_smoothing(flat, pal, kf, tile, z, at, cmp20 * 0.5)
    assert t[:, 0].tolist() == [2, 0, 0, 1, 3] and a[:, 0].tolist() == [0, 0, 4, 0, 0]
    assert src[:, 0].tolist() == [0, 1, 1, 3, 4]      # frame 2's item is a copy of frame 1's
    # above it: frame 1 (tile 0 < tile 2) overwrites frame 0; frames 3, 4 differ by a full white tile
    src, t, p, a = o.temporal_smoothing(flat, pal, kf, tile, z, at, cmp20 * 1.5)
    assert t[:, 0].tolist() == [0, 0, 0, 1, 3] and a[:, 0].tolist() == [0, 4, 4, 0, 0]
    assert src[:, 0].tolist() == [1, 1, 1, 3, 4]
    # large threshold: inside keyframe 1 tile 3 >= tile 1, so frame 4 takes frame 3's item; no merge across keyframes
    src, t, p, a = o.temporal_smoothing(flat, pal, kf, tile, z, at, 10.0)
    assert t[:, 0].tolist() == [0, 0, 0, 1, 1] and a[:, 0].tolist() == [0, 4, 4, 0, 4]
    assert src[:, 0].tolist() == [1, 1, 1, 3, 3]
