"""PrepareFrameTiling (main.pas:3791-3967) in one library call vs the CPU oracle, and the whole
GlobalFT -> PrepareFrameTiling -> DoFrameTiling chain (F1 -> F2 -> F3 -> F8 of SURVEY.md §8a)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def tb():
    import tiler_b200 as t
    t.set_device(0)
    return t


def _case(o, ntiles=300, npal=5, nitems=4000, seed=21):
    s = o.BASE_SEED + seed
    pals = []
    for _ in range(npal):
        p, s = o.gen_palette(s)
        pals.append(p)
    pals = np.stack(pals)
    tiles, s = o.gen_tileset(s, ntiles)
    rng = np.random.default_rng(seed)
    # symmetric tiles: identical stage-1 rows -> equal errs -> the skip rule of main.pas:3835-3837 matters
    tiles[3] = 7
    tiles[4] = np.tile(np.array([1, 2, 3, 4, 4, 3, 2, 1], np.uint8), 8)
    tiles[5] = tiles[4]  # duplicate tile
    flags = rng.integers(0, 4, ntiles).astype(np.uint8)
    item_pal = rng.integers(0, npal, nitems).astype(np.int32)
    item_tile = rng.integers(0, ntiles // 2, nitems).astype(np.int32)  # only half the tiles appear in tilemaps
    item_tile[:6] = [3, 4, 5, 3, 4, 5]
    cen = rng.normal(size=(npal, 192))
    cen[1] = cen[0] + 0.01  # two close palettes: ftMedium selects both
    return pals, tiles, flags, item_pal, item_tile, cen


@pytest.mark.parametrize("quality", [0, 1, 2])
def test_keyframe_dataset_matches_oracle(tb, oracle, quality):
    o = oracle
    pals, tiles, flags, item_pal, item_tile, cen = _case(o)
    rows, tt, tp, ta, used = o.prepare_frame_tiling(tiles, pals, item_pal, item_tile, quality=quality, pal_tol=0.05,
                                                    pal_centroids=cen, tile_flags=flags)
    g, _, _ = tb.KDTree.global_ft(tiles)
    with g:
        t = g.keyframe_dataset(tiles, pals, item_pal, item_tile, quality=quality, pal_tol=0.05, pal_centroids=cen,
                               tile_flags=flags)
        with t:
            assert t.n == len(rows), (t.n, len(rows))
            assert np.array_equal(t.attrs["tr_tile"], tt)
            assert np.array_equal(t.attrs["tr_pal"], tp)
            assert np.array_equal(t.attrs["tr_attrs"], ta)
            assert np.array_equal(t.rows(), rows)
        # second keyframe on the same stage-1 handle reuses the cached 8-NN lists
        t2 = g.keyframe_dataset(tiles, pals, item_pal[:100], item_tile[:100], quality=0, tile_flags=flags)
        r2 = o.prepare_frame_tiling(tiles, pals, item_pal[:100], item_tile[:100], quality=0, tile_flags=flags)
        with t2:
            assert np.array_equal(t2.rows(), r2[0]) and np.array_equal(t2.attrs["tr_tile"], r2[1])
    if quality == 1:
        assert used[1].sum() > 0 and used[0].sum() > 0


def test_keyframe_chain_frametiling(tb, oracle):
    """F1 -> F2 -> F3 -> F8: tilemap items (tile, pal, HMirror, VMirror) equal the oracle's wherever the gap is solid."""
    o = oracle
    pals, tiles, flags, item_pal, item_tile, cen = _case(o, ntiles=400, npal=3, nitems=3000, seed=5)
    rgb, _ = o.gen_queries(o.BASE_SEED + 77, 2500, tiles, pals[0])
    rows, tt, tp, ta, _ = o.prepare_frame_tiling(tiles, pals, item_pal, item_tile, quality=2, tile_flags=flags)
    q = o.psyv_queries_f32(rgb)
    oi, od, nxt = o.knn_f64(rows, q, 1)
    g, _, _ = tb.KDTree.global_ft(tiles)
    with g, g.keyframe_dataset(tiles, pals, item_pal, item_tile, quality=2, tile_flags=flags) as t:
        idx, err = t.frametiling_search(rgb)
        solid = (nxt - od[:, 0]) > 1e-5 * np.maximum(od[:, 0], 1e-12)
        assert solid.sum() > 0.9 * len(rgb)
        a = t.attrs
        for name, ref in (("tr_tile", tt), ("tr_pal", tp), ("tr_attrs", ta)):
            assert np.array_equal(a[name][idx[solid]], ref[oi[solid, 0]]), name  # main.pas:4029-4034
        assert np.allclose(err, od[:, 0], rtol=1e-5, atol=1e-7)


def test_keyframe_bad_arguments(tb, oracle):
    from tiler_b200._lib import TMError
    o = oracle
    pals, tiles, flags, item_pal, item_tile, cen = _case(o, ntiles=64, npal=2, nitems=50)
    g, _, _ = tb.KDTree.global_ft(tiles)
    with g:
        bad = item_tile.copy()
        bad[7] = 64  # outside the tileset
        with pytest.raises(TMError):
            g.keyframe_dataset(tiles, pals, item_pal, bad, quality=0)
        with pytest.raises(TMError):
            g.keyframe_dataset(tiles[:32], pals, item_pal, item_tile % 32, quality=0)  # not this handle's tileset
        with pytest.raises(TMError):
            g.keyframe_dataset(tiles, pals, item_pal, item_tile, quality=1)  # ftMedium needs centroids
        t = g.keyframe_dataset(tiles, pals, item_pal[:0], item_tile[:0], quality=0)  # empty keyframe
        assert t.n == 0
        t.destroy()


@pytest.mark.parametrize("quality", [0, 1])
def test_keyframe_dataset_with_inactive_tiles(tb, oracle, quality):
    """FTiles with inactive entries (after MergeTiles, before Reindex): PrepareGlobalFT emits stage-1 rows for Active
    tiles only and maps hits back through TRToTileIdx (main.pas:3763-3777, 3843-3850); tilemap items and the used[]
    table keep indexing the whole tile array."""
    o = oracle
    pals, tiles, flags, item_pal, item_tile, cen = _case(o, ntiles=400, seed=33)
    rng = np.random.default_rng(5)
    active = (rng.random(400) < 0.6).astype(np.uint8)
    active[[3, 4]] = 1
    active[5] = 0                                  # the duplicate of tile 4 is inactive: it may be queried, never hit
    item_tile = item_tile.copy()
    rows, tt, tp, ta, used = o.prepare_frame_tiling(tiles, pals, item_pal, item_tile, quality=quality, pal_tol=0.05,
                                                    pal_centroids=cen, tile_flags=flags, active=active)
    assert not used[:, active == 0].any()          # inactive tiles never enter the stage-2 dataset
    g, tr_tile, tr_attrs = tb.KDTree.global_ft(tiles, active=active)
    assert g.n == 4 * int(active.sum())
    assert np.array_equal(tr_tile, np.repeat(np.flatnonzero(active), 4))
    assert np.array_equal(tr_attrs, np.tile(np.array([0, 1, 3, 2], np.uint8), int(active.sum())))
    want_rows, _, _ = o.global_rows_f32(tiles[active != 0])
    with g:
        assert np.array_equal(g.rows(), want_rows)
        t = g.keyframe_dataset(tiles, pals, item_pal, item_tile, quality=quality, pal_tol=0.05, pal_centroids=cen,
                               tile_flags=flags)
        with t:
            assert t.n == len(rows), (t.n, len(rows))
            assert np.array_equal(t.attrs["tr_tile"], tt)
            assert np.array_equal(t.attrs["tr_pal"], tp)
            assert np.array_equal(t.attrs["tr_attrs"], ta)
            assert np.array_equal(t.rows(), rows)
    # all-active mask == no mask
    g2, tr2, _ = tb.KDTree.global_ft(tiles, active=np.ones(400, np.uint8))
    with g2:
        assert g2.n == 1600 and np.array_equal(tr2, np.repeat(np.arange(400), 4))
