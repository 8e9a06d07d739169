"""Python view of the C ABI (include/tmann.h), named after the reference's bindings
(extern.pas:63-67) so parity tests read like calls the encoder makes.

numpy arrays are HOST buffers (the h_* entry points, copies inside the call);
objects with a ``data_ptr()`` (torch CUDA tensors) are DEVICE buffers (the d_*
entry points).  torch is never imported here: it is plumbing of the callers.
"""
import ctypes as C

import numpy as np

from ._lib import TMStats, check, check_handle, lib

PSYV_WAVELETS, PSYV_QWEIGHT = 1, 2
KM_ASM, KM_GENERIC = 0, 1
ENGINE_AUTO, ENGINE_SIMT, ENGINE_TC = 0, 1, 2


def _hp(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def _is_dev(x):
    return hasattr(x, "data_ptr")


def _dp(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


def set_device(dev):
    check(lib.tm_set_device(int(dev)))


def set_devices(devs):
    """Drive several GPUs from this process (tm_set_devices): datasets are built on devs[0] and broadcast over NCCL."""
    devs = [int(d) for d in devs]
    arr = (C.c_int * len(devs))(*devs)
    check(lib.tm_set_devices(arr, len(devs)))


def get_devices():
    arr = (C.c_int * 16)()
    n = lib.tm_get_devices(arr, 16)
    if n < 0:
        check(n)
    return [int(arr[i]) for i in range(n)]


def comm_init_ms():
    return float(lib.tm_comm_init_ms())


def set_engine(engine):
    check(lib.tm_set_engine(int(engine)))


def set_stream(cuda_stream_ptr):
    check(lib.tm_set_stream(C.c_void_p(int(cuda_stream_ptr))))


def stats():
    s = TMStats()
    check(lib.tm_get_stats(C.byref(s)))
    return {f: getattr(s, f) for f, _ in TMStats._fields_}


def set_profiling(on):
    """False/0 off, True/1 phase timing, 2 phase timing + error-model audit of the tensor-core engine."""
    check(lib.tm_set_profiling(int(on)))


def reset_stats():
    check(lib.tm_reset_stats())


class KDTree:
    """Dataset handle (PANNkdtree, extern.pas:36-39)."""

    def __init__(self, handle, attrs=None):
        self._h = None
        self._h = check_handle(handle)
        self.attrs = attrs or {}

    # -- constructors ---------------------------------------------------------
    @classmethod
    def create(cls, rows):
        """ann_kdtree_create through an array of row pointers (extern.pas:63)."""
        rows = np.ascontiguousarray(rows, np.float32)
        n, dd = rows.shape
        ptrs = (C.c_void_p * max(n, 1))(*[rows[i].ctypes.data for i in range(n)])
        return cls(lib.ann_kdtree_create(ptrs, n, dd, 1, 0))

    @classmethod
    def from_rows(cls, rows):
        if _is_dev(rows):
            n, dd = rows.shape
            return cls(lib.tm_dataset_create(_dp(rows), n, dd, 1))
        rows = np.ascontiguousarray(rows, np.float32)
        return cls(lib.tm_dataset_create(_hp(rows), rows.shape[0], rows.shape[1], 0))

    @classmethod
    def from_tileset(cls, palpix, pals, tile_flags=None, wavelets=True, gamma=-1, qweight=False):
        """Stage-2 dataset of DoPsyV (main.pas:3883-3919), all (pal, tile, vmir, hmir) used."""
        flags = (PSYV_WAVELETS if wavelets else 0) | (PSYV_QWEIGHT if qweight else 0)
        if _is_dev(palpix):
            npal, ps = pals.shape
            h = lib.tm_tileset_dataset_create_dev(_dp(palpix), palpix.shape[0], _dp(pals), npal, ps,
                                                  _dp(tile_flags), flags, gamma)
            return cls(h)
        palpix = np.ascontiguousarray(palpix, np.uint8).reshape(-1, 64)
        pals = np.ascontiguousarray(pals, np.int32)
        pals = pals.reshape(1, -1) if pals.ndim == 1 else pals
        tf = None if tile_flags is None else np.ascontiguousarray(tile_flags, np.uint8)
        h = lib.tm_tileset_dataset_create(_hp(palpix), palpix.shape[0], _hp(pals), pals.shape[0],
                                          pals.shape[1], _hp(tf), flags, gamma)
        return cls(h)

    @classmethod
    def global_ft(cls, palpix, active=None):
        """PrepareGlobalFT (main.pas:3736-3780): returns (tree, TRToTileIdx, TRToAttrs).
        active: optional bool/uint8 [ntiles] = T^.Active (rows only for active tiles, main.pas:3763-3777)."""
        palpix = np.ascontiguousarray(palpix, np.uint8).reshape(-1, 64)
        act = None if active is None else np.ascontiguousarray(active, np.uint8)
        n = (palpix.shape[0] if act is None else int((act != 0).sum())) * 4
        tr_tile = np.empty(n, np.int32)
        tr_attrs = np.empty(n, np.uint8)
        h = lib.tm_global_dataset_create_active(_hp(palpix), palpix.shape[0], _hp(act), _hp(tr_tile), _hp(tr_attrs))
        return cls(h), tr_tile, tr_attrs

    def keyframe_dataset(self, palpix, pals, item_pal, item_tile, quality=1, pal_tol=0.05, pal_centroids=None,
                         tile_flags=None, wavelets=True, gamma=-1):
        """PrepareFrameTiling (main.pas:3791-3967) in one call; self is the stage-1 tree of global_ft().
        Returns the stage-2 tree; its .attrs holds TRToTileIdx / TRToPalIdx / TRToAttrs."""
        palpix = np.ascontiguousarray(palpix, np.uint8).reshape(-1, 64)
        pals = np.ascontiguousarray(pals, np.int32)
        pals = pals.reshape(1, -1) if pals.ndim == 1 else pals
        item_pal = np.ascontiguousarray(item_pal, np.int32)
        item_tile = np.ascontiguousarray(item_tile, np.int32)
        cen = None if pal_centroids is None else np.ascontiguousarray(pal_centroids, np.float64)
        tf = None if tile_flags is None else np.ascontiguousarray(tile_flags, np.uint8)
        flags = PSYV_WAVELETS if wavelets else 0
        h = lib.tm_keyframe_dataset_create(self._h, _hp(palpix), palpix.shape[0], _hp(tf), _hp(pals), pals.shape[0],
                                           pals.shape[1], _hp(item_pal), _hp(item_tile), len(item_pal), int(quality),
                                           float(pal_tol), _hp(cen), flags, gamma)
        t = KDTree(h)
        n = t.n
        tr_tile, tr_pal, tr_attrs = np.empty(n, np.int32), np.empty(n, np.int32), np.empty(n, np.uint8)
        check(lib.tm_keyframe_tables(t._h, _hp(tr_tile), _hp(tr_pal), _hp(tr_attrs)))
        t.attrs = dict(tr_tile=tr_tile, tr_pal=tr_pal, tr_attrs=tr_attrs)
        return t

    # -- lifetime ---------------------------------------------------------------
    def destroy(self):
        if self._h:
            lib.ann_kdtree_destroy(self._h)
            self._h = None

    __del__ = destroy

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.destroy()

    @property
    def n(self):
        return int(lib.tm_dataset_rows(self._h))

    @property
    def dd(self):
        return int(lib.tm_dataset_dim(self._h))

    @property
    def replicas(self):
        return int(lib.tm_dataset_replicas(self._h))

    def rows(self, replica=0):
        out = np.empty((self.n, self.dd), np.float32)
        check(lib.tm_dataset_read_replica(self._h, int(replica), _hp(out)))
        return out

    # -- legacy single-query entry points (extern.pas:65-67) ----------------------
    def search(self, q, eps=0.0):
        q = np.ascontiguousarray(q, np.float32)
        err = C.c_float()
        idx = lib.ann_kdtree_search(self._h, _hp(q), eps, C.byref(err))
        return idx, err.value

    def pri_search(self, q, eps=0.0):
        q = np.ascontiguousarray(q, np.float32)
        err = C.c_float()
        idx = lib.ann_kdtree_pri_search(self._h, _hp(q), eps, C.byref(err))
        return idx, err.value

    def search_multi(self, q, cnt, eps=0.0):
        q = np.ascontiguousarray(q, np.float32)
        idxs = np.full(cnt, -1, np.int32)
        errs = np.zeros(cnt, np.float32)
        lib.ann_kdtree_search_multi(self._h, _hp(idxs), _hp(errs), cnt, _hp(q), eps)
        return idxs, errs

    # -- batch ------------------------------------------------------------------
    def search_batch(self, q, k=1, out_idx=None, out_err=None):
        if _is_dev(q):
            check(lib.tm_search_batch_dev(self._h, _dp(q), q.shape[0], k, _dp(out_idx), _dp(out_err)))
            return out_idx, out_err
        q = np.ascontiguousarray(q, np.float32).reshape(-1, self.dd)
        idx = np.empty((q.shape[0], k), np.int32)
        err = np.empty((q.shape[0], k), np.float32)
        check(lib.ann_kdtree_search_batch(self._h, _hp(q), q.shape[0], k, _hp(idx), _hp(err)))
        return idx, err

    def search_batch_into(self, q, k, out_idx, out_err):
        """ann_kdtree_search_batch with caller-owned (e.g. pinned) host buffers: no allocation, no copy in Python"""
        check(lib.ann_kdtree_search_batch(self._h, _hp(q), q.shape[0], int(k), _hp(out_idx), _hp(out_err)))
        return out_idx, out_err

    def frametiling_search(self, rgb, wavelets=True, gamma=-1, out_idx=None, out_err=None):
        """DoFrameTiling hot loop (main.pas:4020-4038) for a batch of RGB tiles."""
        flags = PSYV_WAVELETS if wavelets else 0
        if _is_dev(rgb):
            check(lib.tm_frametiling_search_dev(self._h, _dp(rgb), rgb.shape[0], flags, gamma,
                                                _dp(out_idx), _dp(out_err)))
            return out_idx, out_err
        rgb = np.ascontiguousarray(rgb, np.int32).reshape(-1, 64)
        idx = out_idx if out_idx is not None else np.empty(rgb.shape[0], np.int32)
        err = out_err if out_err is not None else np.empty(rgb.shape[0], np.float32)
        check(lib.tm_frametiling_search(self._h, _hp(rgb), rgb.shape[0], flags, gamma, _hp(idx), _hp(err)))
        return idx, err


def psyv_queries(rgb, wavelets=True, gamma=-1, qweight=False, out=None):
    flags = (PSYV_WAVELETS if wavelets else 0) | (PSYV_QWEIGHT if qweight else 0)
    if _is_dev(rgb):
        check(lib.tm_psyv_queries_dev(_dp(rgb), rgb.shape[0], flags, gamma, _dp(out)))
        return out
    rgb = np.ascontiguousarray(rgb, np.int32).reshape(-1, 64)
    out = np.empty((rgb.shape[0], 192), np.float32)
    check(lib.tm_psyv_queries(_hp(rgb), rgb.shape[0], flags, gamma, _hp(out)))
    return out


def kmodes_assign(X, centroids, semantics=KM_ASM, out_clust=None, out_dis=None):
    """TKModes.DoGMMD over all points (kmodes.pas:825-835, 984-997)."""
    if _is_dev(X):
        check(lib.tm_kmodes_assign_dev(_dp(X), X.shape[0], _dp(centroids), centroids.shape[0], semantics,
                                       _dp(out_clust), _dp(out_dis)))
        return out_clust, out_dis
    X = np.ascontiguousarray(X, np.uint8).reshape(-1, 80)
    centroids = np.ascontiguousarray(centroids, np.uint8).reshape(-1, 80)
    clust = np.empty(X.shape[0], np.int32)
    dis = np.empty(X.shape[0], np.uint64)
    check(lib.tm_kmodes_assign(_hp(X), X.shape[0], _hp(centroids), centroids.shape[0], semantics,
                               _hp(clust), _hp(dis)))
    return clust, dis


class KModesState:
    """Device-resident points + centroids for the per-bin loop of KModesIter (kmodes.pas:845-909)."""

    def __init__(self, X, semantics=KM_ASM):
        self._h = None
        X = np.ascontiguousarray(X, np.uint8).reshape(-1, 80)
        self.n = X.shape[0]
        self.k = 0
        self._h = check_handle(lib.tm_kmodes_open(_hp(X), self.n, int(semantics)))

    def set_centroids(self, centroids):
        centroids = np.ascontiguousarray(centroids, np.uint8).reshape(-1, 80)
        check(lib.tm_kmodes_set_centroids(self._h, _hp(centroids), centroids.shape[0]))
        self.k = centroids.shape[0]

    def update_centroids(self, idx, rows):
        idx = np.ascontiguousarray(idx, np.int32)
        rows = np.ascontiguousarray(rows, np.uint8).reshape(-1, 80)
        check(lib.tm_kmodes_update_centroids(self._h, _hp(idx), _hp(rows), len(idx)))

    def assign_range(self, lo, hi, with_dis=True):
        clust = np.empty(hi - lo, np.int32)
        dis = np.empty(hi - lo, np.uint64) if with_dis else None
        check(lib.tm_kmodes_assign_range(self._h, int(lo), int(hi), _hp(clust), _hp(dis)))
        return clust, dis

    def centroids(self):
        out = np.empty((self.k, 80), np.uint8)
        check(lib.tm_kmodes_read_centroids(self._h, _hp(out)))
        return out

    def close(self):
        if self._h:
            lib.tm_kmodes_close(self._h)
            self._h = None

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()


def kmodes_update_min_distance(X, center, mindist, semantics=KM_ASM):
    X = np.ascontiguousarray(X, np.uint8).reshape(-1, 80)
    center = np.ascontiguousarray(center, np.uint8)
    mindist = np.ascontiguousarray(mindist, np.uint64).copy()
    check(lib.tm_kmodes_update_min_distance(_hp(X), X.shape[0], _hp(center), semantics, _hp(mindist)))
    return mindist


def kmodes_farthest_first(X, k, init_point, semantics=KM_ASM):
    """TKModes.InitFarthestFirst (kmodes.pas:698-776): returns (centroids[k][80], chosen[k])."""
    X = np.ascontiguousarray(X, np.uint8).reshape(-1, 80)
    cent = np.empty((k, 80), np.uint8)
    chosen = np.empty(k, np.int64)
    check(lib.tm_kmodes_farthest_first(_hp(X), X.shape[0], int(k), int(init_point), semantics, _hp(cent), _hp(chosen)))
    return cent, chosen


def kmodes_representatives(X, labels, centroids, semantics=KM_ASM):
    """Per-cluster representative pick of DoKModes (main.pas:4231-4253): returns (rep[k], count[k])."""
    X = np.ascontiguousarray(X, np.uint8).reshape(-1, 80)
    labels = np.ascontiguousarray(labels, np.int32)
    centroids = np.ascontiguousarray(centroids, np.uint8).reshape(-1, 80)
    k = centroids.shape[0]
    rep = np.empty(k, np.int32)
    cnt = np.empty(k, np.int32)
    check(lib.tm_kmodes_representatives(_hp(X), X.shape[0], _hp(labels), _hp(centroids), k, semantics, _hp(rep), _hp(cnt)))
    return rep, cnt


def reindex(tilemap, ntiles):
    """btnReindexClick + ReindexTiles (main.pas:1199-1230, 4483-4527).  numpy in: returns
    (new_tilemap, use_count, idx_map, order[:n_active]); torch CUDA int32 tensor in: rewritten in place,
    returns (use_count, idx_map, order, n_active) as tensors allocated like the input."""
    act = C.c_int64()
    if _is_dev(tilemap):
        import torch
        cnt = torch.empty(ntiles, dtype=torch.int32, device=tilemap.device)
        imap = torch.empty_like(cnt)
        order = torch.empty_like(cnt)
        check(lib.tm_reindex_dev(_dp(tilemap), tilemap.numel(), int(ntiles), _dp(cnt), _dp(imap), _dp(order), C.byref(act)))
        return cnt, imap, order, int(act.value)
    tm = np.ascontiguousarray(tilemap, np.int32).copy()
    cnt = np.empty(ntiles, np.int32)
    imap = np.empty(ntiles, np.int32)
    order = np.empty(ntiles, np.int32)
    check(lib.tm_reindex(_hp(tm), tm.size, int(ntiles), _hp(cnt), _hp(imap), _hp(order), C.byref(act)))
    return tm, cnt, imap, order[: act.value]


def kmodes_rows(palpix, pal_size=16):
    """WriteTileDatasetLine + GetTilePalZoneThres (main.pas:4142-4183): returns (rows[n][80], signi[n])."""
    palpix = np.ascontiguousarray(palpix, np.uint8).reshape(-1, 64)
    rows = np.empty((palpix.shape[0], 80), np.uint8)
    signi = np.empty(palpix.shape[0], np.int32)
    check(lib.tm_kmodes_rows(_hp(palpix), palpix.shape[0], int(pal_size), _hp(rows), _hp(signi)))
    return rows, signi


def kmodes_assign_binned(X, xbin, centroids, cbin, nbins=65, semantics=KM_ASM):
    """DoFindBest of ReloadPreviousTiling (main.pas:4377-4409)."""
    X = np.ascontiguousarray(X, np.uint8).reshape(-1, 80)
    centroids = np.ascontiguousarray(centroids, np.uint8).reshape(-1, 80)
    xbin = np.ascontiguousarray(xbin, np.int32)
    cbin = np.ascontiguousarray(cbin, np.int32)
    clust = np.empty(X.shape[0], np.int32)
    dis = np.empty(X.shape[0], np.uint64)
    check(lib.tm_kmodes_assign_binned(_hp(X), X.shape[0], _hp(xbin), _hp(centroids), centroids.shape[0], _hp(cbin),
                                      int(nbins), semantics, _hp(clust), _hp(dis)))
    return clust, dis


def temporal_smoothing(palpix, pals, frame_kf, tile, pal, attrs, strength):
    """btnSmoothClick + DoTemporalSmoothing (main.pas:1336-1370, 4071-4119) for a whole clip.
    pals [nkf][npal][pal_size]; tile / pal / attrs [nframes][npos].  Returns (src_frame, tile, pal, attrs)."""
    palpix = np.ascontiguousarray(palpix, np.uint8).reshape(-1, 64)
    pals = np.ascontiguousarray(pals, np.int32)
    if pals.ndim != 3:
        raise ValueError("pals must be [nkf][npal][pal_size]")
    frame_kf = np.ascontiguousarray(frame_kf, np.int32)
    tile = np.ascontiguousarray(tile, np.int32)
    pal = np.ascontiguousarray(pal, np.int32)
    attrs = np.ascontiguousarray(attrs, np.uint8)
    nframes, npos = tile.shape
    src = np.empty((nframes, npos), np.int32)
    ot, op, oa = np.empty_like(tile), np.empty_like(pal), np.empty_like(attrs)
    check(lib.tm_temporal_smoothing(_hp(palpix), palpix.shape[0], _hp(pals), pals.shape[0], pals.shape[1], pals.shape[2],
                                    _hp(frame_kf), nframes, npos, _hp(tile), _hp(pal), _hp(attrs), float(strength),
                                    _hp(src), _hp(ot), _hp(op), _hp(oa)))
    return src, ot, op, oa
