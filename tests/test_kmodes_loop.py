"""The whole KModes loop (kmodes.pas:845-909, 917-1060): the C restatement in oracle/kmodes_ref.c against a second
restatement in numpy (tests/kmodes_loop.py) on the CPU, and - on a GPU - the same loop driven through the library's
device-resident handle, one tm_kmodes_assign_range call per 960-point bin with only the changed centroid rows sent."""
import numpy as np
import pytest

from tests import kmodes_loop as kl


def _rows(o, n, seed, pal_size=16):
    tiles, _ = o.gen_tileset(o.BASE_SEED + seed, n, pal_size)
    return o.kmodes_rows(tiles, pal_size)


@pytest.mark.parametrize("n,k,start,sem", [(2500, 24, 7, 0), (1500, 60, 0, 1), (961, 5, 960, 0)])
def test_c_and_numpy_restatements_agree(oracle, n, k, start, sem):
    o = oracle
    X = _rows(o, n, n + k)
    want = o.compute_kmodes(X, k, start, 16, sem)
    got = kl.compute_kmodes(
        X, k, start, 16,
        farthest_first=lambda s: o.farthest_first(X, k, s, sem)[0],
        assign_all=lambda c: o.kmodes_assign(X, c, sem)[0],
        assign_bin=lambda lo, hi, L: o.kmodes_assign(X[lo:hi], L.cent, sem))
    assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1])
    assert got[2:] == want[2:], (got[2:], want[2:])
    assert want[2] >= 2 and want[4] > 0  # the loop did iterate and move points


def test_empty_cluster_paths_are_exercised(oracle):
    """Many duplicate points + more clusters than distinct rows: farthest-first picks duplicates, clusters come out of
    the first assignment empty (re-seeded attribute by attribute from random points, kmodes.pas:1008-1009) and clusters
    are emptied by moves (re-seeded from the largest cluster through RandInt, 890-905)."""
    o = oracle
    base = _rows(o, 40, 99)
    X = np.concatenate([base] * 30)[np.random.default_rng(2).permutation(1200)]
    want = o.compute_kmodes(X, 64, 3, 16, 0)
    got = kl.compute_kmodes(
        X, 64, 3, 16,
        farthest_first=lambda s: o.farthest_first(X, 64, s, 0)[0],
        assign_all=lambda c: o.kmodes_assign(X, c, 0)[0],
        assign_bin=lambda lo, hi, L: o.kmodes_assign(X[lo:hi], L.cent, 0))
    assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1]) and got[2:] == want[2:]


@pytest.mark.gpu
@pytest.mark.parametrize("n,k,sem", [(6000, 48, 0), (3000, 700, 1)])
def test_gpu_drives_the_whole_loop_bit_exact(oracle, n, k, sem):
    import tiler_b200 as tb
    o = oracle
    X = _rows(o, n, 5 * n + k)
    want = o.compute_kmodes(X, k, 11, 16, sem)
    calls = {"bins": 0, "rows_sent": 0}
    with tb.KModesState(X, sem) as st:
        def assign_bin(lo, hi, L):
            if L.touched:  # MovePointCat's edits since the previous bin
                idx = np.array(sorted(L.touched), np.int32)
                st.update_centroids(idx, L.cent[idx])
                calls["rows_sent"] += len(idx)
            calls["bins"] += 1
            return st.assign_range(lo, hi)
        got = kl.compute_kmodes(
            X, k, 11, 16,
            farthest_first=lambda s: tb.kmodes_farthest_first(X, k, s, sem)[0],
            assign_all=lambda c: tb.kmodes_assign(X, c, sem)[0],
            assign_bin=assign_bin, on_centroids=st.set_centroids)
        dev_cent = st.centroids()
    assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1])
    assert got[2:] == want[2:], (got[2:], want[2:])
    assert calls["bins"] == want[2] * ((n + 959) // 960)
    # the device copy tracked every MovePointCat edit but the ones after the last bin
    assert dev_cent.shape == (k, 80)


@pytest.mark.gpu
def test_kmodes_handle_matches_one_shot_calls(oracle):
    import tiler_b200 as tb
    o = oracle
    X = _rows(o, 5000, 77)
    C = X[::50].copy()
    rng = np.random.default_rng(0)
    with tb.KModesState(X) as st:
        st.set_centroids(C)
        for lo, hi in [(0, 960), (960, 1920), (4800, 5000), (0, 5000), (17, 18)]:
            gc, gd = st.assign_range(lo, hi)
            wc, wd = o.kmodes_assign(X[lo:hi], C)
            assert np.array_equal(gc, wc) and np.array_equal(gd, wd)
        idx = rng.choice(len(C), 17, replace=False).astype(np.int32)
        rows = X[rng.integers(0, 5000, 17)]
        rows[3, 5] = 200                           # a byte >= 16: the byte-wise path must take over
        C2 = C.copy()
        C2[idx] = rows
        st.update_centroids(idx, rows)
        gc, gd = st.assign_range(100, 2100)
        wc, wd = o.kmodes_assign(X[100:2100], C2)
        assert np.array_equal(gc, wc) and np.array_equal(gd, wd)
        assert np.array_equal(st.centroids(), C2)
        with pytest.raises(Exception):
            st.update_centroids(np.array([len(C)], np.int32), rows[:1])
        with pytest.raises(Exception):
            st.assign_range(0, 5001)


@pytest.mark.gpu
@pytest.mark.parametrize("m", [5, 24, 25, 300])
def test_kmodes_handle_updates_with_duplicates_last_row_wins(oracle, m):
    """tm_kmodes_update_centroids with repeated centroid indices behaves like sequential stores, through the rows-in-
    kernel-parameters path (m <= 24) and the staged path; bin-sized ranges (results polled from mapped host memory)
    and a whole-set range (copied) see the same centroids."""
    import tiler_b200 as tb
    o = oracle
    X = _rows(o, 4000, 123 + m)
    C = X[::10].copy()                             # 400 centroids: bin-sized calls split them over the grid
    rng = np.random.default_rng(m)
    idx = rng.integers(0, 40, m).astype(np.int32)  # few distinct targets: many duplicates
    rows = X[rng.integers(0, 4000, m)].copy()
    C2 = C.copy()
    for i in range(m):
        C2[idx[i]] = rows[i]
    with tb.KModesState(X) as st:
        st.set_centroids(C)
        st.update_centroids(idx, rows)
        assert np.array_equal(st.centroids(), C2)
        for lo, hi in [(0, 960), (960, 1920), (0, 4000), (3999, 4000)]:
            gc, gd = st.assign_range(lo, hi)
            wc, wd = o.kmodes_assign(X[lo:hi], C2)
            assert np.array_equal(gc, wc) and np.array_equal(gd, wd)
        gc, _ = st.assign_range(0, 960, with_dis=False)
        assert np.array_equal(gc, o.kmodes_assign(X[:960], C2)[0])
        # two updates and two bins back to back: the staging of the first is still in flight when the second lands
        st.update_centroids(idx[::-1].copy(), rows[::-1].copy())
        st.update_centroids(idx[:3], rows[:3])
        C3 = C2.copy()
        for i in range(m - 1, -1, -1):
            C3[idx[i]] = rows[i]
        for i in range(3):
            C3[idx[i]] = rows[i]
        gc, gd = st.assign_range(1000, 1960)
        wc, wd = o.kmodes_assign(X[1000:1960], C3)
        assert np.array_equal(gc, wc) and np.array_equal(gd, wd)
