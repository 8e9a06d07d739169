"""ctypes binding of the CPU oracle (oracle/libtmoracle.so).

TEST INFRASTRUCTURE ONLY: importable from tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs.  The product path
(tiler_b200/) never imports this module.  PARITY UNPINNED (see tmoracle.h).
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))


def build(march="x86-64-v3", out="libtmoracle.so"):
    subprocess.run(["make", "-s", "-C", _HERE, f"MARCH={march}", f"OUT={out}"], check=True)
    return os.path.join(_HERE, out)


def _load(name="libtmoracle.so"):
    path = os.path.join(_HERE, name)
    if not os.path.exists(path):
        build(out=name)
    return C.CDLL(path)


_lib = _load()


def use_native():
    """Rebuild with -march=native on this machine (timed CPU baseline only)."""
    global _lib
    _lib = C.CDLL(build(march="native", out="libtmoracle_native.so"))
    _proto(_lib)


def _p(a, t):
    return a.ctypes.data_as(C.POINTER(t)) if a is not None else None


def _proto(lib):
    lib.tmo_randint.restype = C.c_uint32
    lib.tmo_dissim.restype = C.c_uint64
    lib.tmo_min_dissim.restype = C.c_int32
    lib.tmo_max_threads.restype = C.c_int
    lib.tmo_prepare_tile_mirrors.restype = C.c_int
    lib.tmo_kmodes_row.restype = C.c_int
    lib.tmo_kdtree_create.restype = C.c_void_p
    lib.tmo_kdtree_destroy.argtypes = [C.c_void_p]


_proto(_lib)
i64 = C.c_int64


def max_threads():
    return int(_lib.tmo_max_threads())


def set_threads(n):
    """OpenMP team size of the oracle (torchrun forces OMP_NUM_THREADS=1 on its workers)."""
    _lib.tmo_set_threads(int(n))


def rgb_to_yuv(r, g, b, gamma=-1):
    y, u, v = C.c_double(), C.c_double(), C.c_double()
    _lib.tmo_rgb_to_yuv(int(r), int(g), int(b), int(gamma), C.byref(y), C.byref(u), C.byref(v))
    return y.value, u.value, v.value


def psyv(rgb=None, palpix=None, pal=None, use_wavelets=True, qweighting=False,
         hmirror=False, vmirror=False, gamma=-1):
    out = np.zeros(192, np.float64)
    from_pal = rgb is None
    if from_pal:
        palpix = np.ascontiguousarray(palpix, np.uint8).reshape(64)
        pal = np.ascontiguousarray(pal, np.int32)
    else:
        rgb = np.ascontiguousarray(rgb, np.int32).reshape(64)
    _lib.tmo_psyv(_p(rgb, C.c_int32), _p(palpix, C.c_uint8), _p(pal, C.c_int32),
                  int(from_pal), int(use_wavelets), int(qweighting), int(hmirror),
                  int(vmirror), int(gamma), _p(out, C.c_double))
    return out


def psyv_queries_f32(rgb, use_wavelets=True, gamma=-1):
    rgb = np.ascontiguousarray(rgb, np.int32).reshape(-1, 64)
    out = np.empty((rgb.shape[0], 192), np.float32)
    _lib.tmo_psyv_queries_f32(_p(rgb, C.c_int32), i64(rgb.shape[0]), int(use_wavelets),
                              int(gamma), _p(out, C.c_float))
    return out


def psyv_tileset_f32(palpix, pal, use_wavelets=True, gamma=-1):
    palpix = np.ascontiguousarray(palpix, np.uint8).reshape(-1, 64)
    pal = np.ascontiguousarray(pal, np.int32)
    out = np.empty((palpix.shape[0] * 4, 192), np.float32)
    _lib.tmo_psyv_tileset_f32(_p(palpix, C.c_uint8), i64(palpix.shape[0]), _p(pal, C.c_int32),
                              int(use_wavelets), int(gamma), _p(out, C.c_float))
    return out


def hmirror(palpix):
    a = np.array(palpix, np.uint8).reshape(64).copy()
    _lib.tmo_hmirror(_p(a, C.c_uint8))
    return a.reshape(8, 8)


def vmirror(palpix):
    a = np.array(palpix, np.uint8).reshape(64).copy()
    _lib.tmo_vmirror(_p(a, C.c_uint8))
    return a.reshape(8, 8)


def prepare_tile_mirrors(palpix):
    a = np.array(palpix, np.uint8).reshape(64).copy()
    f = _lib.tmo_prepare_tile_mirrors(_p(a, C.c_uint8))
    return a.reshape(8, 8), f


def global_rows_f32(palpix):
    palpix = np.ascontiguousarray(palpix, np.uint8).reshape(-1, 64)
    n = palpix.shape[0]
    out = np.empty((n * 4, 64), np.float32)
    tr_tile = np.empty(n * 4, np.int32)
    tr_attrs = np.empty(n * 4, np.uint8)
    _lib.tmo_global_rows_f32(_p(palpix, C.c_uint8), i64(n), _p(out, C.c_float),
                             _p(tr_tile, C.c_int32), _p(tr_attrs, C.c_uint8))
    return out, tr_tile, tr_attrs


def knn_f64(data, q, k=1):
    data = np.ascontiguousarray(data, np.float32)
    q = np.ascontiguousarray(q, np.float32).reshape(-1, data.shape[1])
    nq = q.shape[0]
    idx = np.empty((nq, k), np.int32)
    dist = np.empty((nq, k), np.float64)
    nxt = np.empty(nq, np.float64)
    _lib.tmo_knn_f64(_p(data, C.c_float), i64(data.shape[0]), int(data.shape[1]),
                     _p(q, C.c_float), i64(nq), int(k), _p(idx, C.c_int32),
                     _p(dist, C.c_double), _p(nxt, C.c_double))
    return idx, dist, nxt


def knn_f32(data, q, k=1):
    data = np.ascontiguousarray(data, np.float32)
    q = np.ascontiguousarray(q, np.float32).reshape(-1, data.shape[1])
    nq = q.shape[0]
    idx = np.empty((nq, k), np.int32)
    dist = np.empty((nq, k), np.float32)
    _lib.tmo_knn_f32(_p(data, C.c_float), i64(data.shape[0]), int(data.shape[1]),
                     _p(q, C.c_float), i64(nq), int(k), _p(idx, C.c_int32), _p(dist, C.c_float))
    return idx, dist


def nn1_blocked_f32(data, q, nthreads=0):
    data = np.ascontiguousarray(data, np.float32)
    q = np.ascontiguousarray(q, np.float32).reshape(-1, data.shape[1])
    nq = q.shape[0]
    idx = np.empty(nq, np.int32)
    dist = np.empty(nq, np.float32)
    _lib.tmo_nn1_blocked_f32(_p(data, C.c_float), i64(data.shape[0]), int(data.shape[1]),
                             _p(q, C.c_float), i64(nq), _p(idx, C.c_int32), _p(dist, C.c_float),
                             int(nthreads))
    return idx, dist


def prepare_frame_tiling(palpix, pals, item_pal, item_tile, quality=1, pal_tol=0.05, pal_centroids=None,
                         tile_flags=None, use_wavelets=True, gamma=-1, active=None):
    """PrepareFrameTiling (main.pas:3791-3967): returns (rows[KNNSize][192] f32, tr_tile, tr_pal, tr_attrs, used)."""
    palpix = np.ascontiguousarray(palpix, np.uint8).reshape(-1, 64)
    pals = np.ascontiguousarray(pals, np.int32)
    pals = pals.reshape(1, -1) if pals.ndim == 1 else pals
    npal, ps = pals.shape
    nt = palpix.shape[0]
    item_pal = np.ascontiguousarray(item_pal, np.int32)
    item_tile = np.ascontiguousarray(item_tile, np.int32)
    cen = None if pal_centroids is None else np.ascontiguousarray(pal_centroids, np.float64).reshape(npal, 192)
    tf = None if tile_flags is None else np.ascontiguousarray(tile_flags, np.uint8)
    used = np.zeros((npal, nt, 4), np.uint8)
    act = None if active is None else np.ascontiguousarray(active, np.uint8)
    _lib.tmo_prepare_frame_tiling_used_active.restype = C.c_int64
    n = _lib.tmo_prepare_frame_tiling_used_active(_p(palpix, C.c_uint8), i64(nt), _p(act, C.c_uint8), int(npal),
                                                  _p(item_pal, C.c_int32), _p(item_tile, C.c_int32), i64(len(item_pal)),
                                                  int(quality), C.c_double(pal_tol), _p(cen, C.c_double),
                                                  _p(used, C.c_uint8))
    rows = np.empty((n, 192), np.float32)
    tr_tile = np.empty(n, np.int32)
    tr_pal = np.empty(n, np.int32)
    tr_attrs = np.empty(n, np.uint8)
    _lib.tmo_frame_tiling_rows(_p(used, C.c_uint8), _p(palpix, C.c_uint8), i64(nt), _p(tf, C.c_uint8),
                               _p(pals, C.c_int32), int(npal), int(ps), int(use_wavelets), int(gamma),
                               _p(rows, C.c_float), _p(tr_tile, C.c_int32), _p(tr_pal, C.c_int32),
                               _p(tr_attrs, C.c_uint8))
    return rows, tr_tile, tr_pal, tr_attrs, used


def reindex(tilemap, ntiles):
    """btnReindexClick + ReindexTiles: returns (new_tilemap, use_count, idx_map, order[:n_active])."""
    tm = np.ascontiguousarray(tilemap, np.int32).copy()
    cnt = np.empty(ntiles, np.int32)
    imap = np.empty(ntiles, np.int32)
    order = np.empty(ntiles, np.int32)
    _lib.tmo_reindex.restype = C.c_int64
    n = _lib.tmo_reindex(_p(tm, C.c_int32), i64(tm.size), i64(ntiles), _p(cnt, C.c_int32), _p(imap, C.c_int32),
                         _p(order, C.c_int32))
    return tm, cnt, imap, order[:n]


class KDTreeRef:
    """ANN-style kd-tree (kdtree_ref.c): ann_kdtree_create(rows, n, dd, 1, ANN_KD_STD) + eps = 0 search."""

    def __init__(self, data, bs=1):
        self.data = np.ascontiguousarray(data, np.float32)  # kept alive: the tree does not copy rows
        self._lib = _lib
        self._h = C.c_void_p(_lib.tmo_kdtree_create(_p(self.data, C.c_float), i64(self.data.shape[0]),
                                                    int(self.data.shape[1]), int(bs)))

    def search(self, q, k=1, nthreads=0, with_stats=False):
        q = np.ascontiguousarray(q, np.float32).reshape(-1, self.data.shape[1])
        nq = q.shape[0]
        idx = np.empty((nq, k), np.int32)
        dist = np.empty((nq, k), np.float32)
        st = np.zeros(2, np.int64)
        self._lib.tmo_kdtree_search(self._h, _p(q, C.c_float), i64(nq), int(k), _p(idx, C.c_int32),
                                    _p(dist, C.c_float), int(nthreads), _p(st, C.c_int64))
        return (idx, dist, st) if with_stats else (idx, dist)

    def close(self):
        if self._h:
            self._lib.tmo_kdtree_destroy(self._h)
            self._h = None

    __del__ = close


# ---- kmodes -----------------------------------------------------------------
ASM, GENERIC = 0, 1


def dissim(a, b, semantics=ASM):
    a = np.ascontiguousarray(a, np.uint8)
    b = np.ascontiguousarray(b, np.uint8)
    assert a.size == 80 and b.size == 80
    return int(_lib.tmo_dissim(_p(a, C.c_uint8), _p(b, C.c_uint8), int(semantics)))


def min_dissim(rows, item, semantics=ASM):
    rows = np.ascontiguousarray(rows, np.uint8).reshape(-1, 80)
    item = np.ascontiguousarray(item, np.uint8)
    best = C.c_uint64()
    r = _lib.tmo_min_dissim(_p(rows, C.c_uint8), i64(rows.shape[0]), _p(item, C.c_uint8),
                            int(semantics), C.byref(best))
    return int(r), int(best.value)


def kmodes_assign(X, centroids, semantics=ASM, nthreads=0):
    X = np.ascontiguousarray(X, np.uint8).reshape(-1, 80)
    centroids = np.ascontiguousarray(centroids, np.uint8).reshape(-1, 80)
    clust = np.empty(X.shape[0], np.int32)
    dis = np.empty(X.shape[0], np.uint64)
    _lib.tmo_kmodes_assign(_p(X, C.c_uint8), i64(X.shape[0]), _p(centroids, C.c_uint8),
                           i64(centroids.shape[0]), int(semantics), _p(clust, C.c_int32),
                           _p(dis, C.c_uint64), int(nthreads))
    return clust, dis


def update_min_distance(X, center, mindist, semantics=ASM):
    X = np.ascontiguousarray(X, np.uint8).reshape(-1, 80)
    center = np.ascontiguousarray(center, np.uint8)
    mindist = np.ascontiguousarray(mindist, np.uint64).copy()
    _lib.tmo_update_min_distance(_p(X, C.c_uint8), i64(X.shape[0]), _p(center, C.c_uint8),
                                 int(semantics), _p(mindist, C.c_uint64))
    return mindist


def farthest_first(X, k, init_point, semantics=ASM):
    X = np.ascontiguousarray(X, np.uint8).reshape(-1, 80)
    cent = np.empty((k, 80), np.uint8)
    chosen = np.empty(k, np.int64)
    _lib.tmo_farthest_first(_p(X, C.c_uint8), i64(X.shape[0]), i64(k), i64(init_point),
                            int(semantics), _p(cent, C.c_uint8), _p(chosen, C.c_int64))
    return cent, chosen


def kmodes_representatives(X, labels, centroids, semantics=ASM):
    X = np.ascontiguousarray(X, np.uint8).reshape(-1, 80)
    labels = np.ascontiguousarray(labels, np.int32)
    centroids = np.ascontiguousarray(centroids, np.uint8).reshape(-1, 80)
    k = centroids.shape[0]
    rep = np.empty(k, np.int32)
    cnt = np.empty(k, np.int32)
    _lib.tmo_kmodes_representatives(_p(X, C.c_uint8), i64(X.shape[0]), _p(labels, C.c_int32), _p(centroids, C.c_uint8),
                                    i64(k), int(semantics), _p(rep, C.c_int32), _p(cnt, C.c_int32))
    return rep, cnt


def compute_kmodes(X, k, start_point, nmod=16, semantics=ASM, max_iter=0):
    """TKModes.ComputeKModes (kmodes.pas:917-1060, one run from start_point).  Returns (labels, centroids, n_iter, cost, moves)."""
    X = np.ascontiguousarray(X, np.uint8).reshape(-1, 80)
    labels = np.empty(X.shape[0], np.int32)
    cent = np.empty((k, 80), np.uint8)
    cost = C.c_uint64()
    moves = C.c_int64()
    it = _lib.tmo_compute_kmodes(_p(X, C.c_uint8), i64(X.shape[0]), i64(k), i64(start_point), int(nmod), int(semantics),
                                 int(max_iter), _p(labels, C.c_int32), _p(cent, C.c_uint8), C.byref(cost), C.byref(moves))
    return labels, cent, int(it), int(cost.value), int(moves.value)


def kmodes_row(palpix, pal_size=16):
    palpix = np.ascontiguousarray(palpix, np.uint8).reshape(64)
    row = np.empty(80, np.uint8)
    r = _lib.tmo_kmodes_row(_p(palpix, C.c_uint8), int(pal_size), _p(row, C.c_uint8))
    return row, int(r)


def kmodes_rows(palpix, pal_size=16):
    palpix = np.ascontiguousarray(palpix, np.uint8).reshape(-1, 64)
    return np.stack([kmodes_row(t, pal_size)[0] for t in palpix])


def temporal_smoothing(palpix, pals, frame_kf, tile, pal, attrs, strength):
    """btnSmoothClick + DoTemporalSmoothing (main.pas:1336-1370, 4071-4119).
    pals [nkf][npal][pal_size]; tile/pal/attrs [nframes][npos].  Returns (src_frame, tile, pal, attrs)."""
    palpix = np.ascontiguousarray(palpix, np.uint8).reshape(-1, 64)
    pals = np.ascontiguousarray(pals, np.int32)
    assert pals.ndim == 3
    frame_kf = np.ascontiguousarray(frame_kf, np.int32)
    tile = np.ascontiguousarray(tile, np.int32)
    pal = np.ascontiguousarray(pal, np.int32)
    attrs = np.ascontiguousarray(attrs, np.uint8)
    nframes, npos = tile.shape
    src = np.empty((nframes, npos), np.int32)
    ot, op, oa = np.empty_like(tile), np.empty_like(pal), np.empty_like(attrs)
    r = _lib.tmo_temporal_smoothing(_p(palpix, C.c_uint8), i64(palpix.shape[0]), _p(pals, C.c_int32), pals.shape[0],
                                    pals.shape[1], pals.shape[2], _p(frame_kf, C.c_int32), i64(nframes), i64(npos),
                                    _p(tile, C.c_int32), _p(pal, C.c_int32), _p(attrs, C.c_uint8), C.c_double(strength),
                                    _p(src, C.c_int32), _p(ot, C.c_int32), _p(op, C.c_int32), _p(oa, C.c_uint8))
    if r:
        raise ValueError("temporal_smoothing: index out of range")
    return src, ot, op, oa


# ---- synthetic generators ---------------------------------------------------
BASE_SEED = 0x42381337  # kmodes.pas:930


def randint(rng, seed):
    s = C.c_uint32(seed)
    r = _lib.tmo_randint(C.c_uint32(rng), C.byref(s))
    return int(r), int(s.value)


def gen_palette(seed, pal_size=16):
    s = C.c_uint32(seed)
    pal = np.empty(pal_size, np.int32)
    _lib.tmo_gen_palette(C.byref(s), int(pal_size), _p(pal, C.c_int32))
    return pal, int(s.value)


def gen_tileset(seed, ntiles, pal_size=16):
    s = C.c_uint32(seed)
    px = np.empty((ntiles, 64), np.uint8)
    _lib.tmo_gen_tileset(C.byref(s), i64(ntiles), int(pal_size), _p(px, C.c_uint8))
    return px, int(s.value)


def gen_queries(seed, nq, palpix, pal, noise=6):
    s = C.c_uint32(seed)
    palpix = np.ascontiguousarray(palpix, np.uint8).reshape(-1, 64)
    pal = np.ascontiguousarray(pal, np.int32)
    rgb = np.empty((nq, 64), np.int32)
    _lib.tmo_gen_queries(C.byref(s), i64(nq), _p(palpix, C.c_uint8), i64(palpix.shape[0]),
                         _p(pal, C.c_int32), int(noise), _p(rgb, C.c_int32))
    return rgb, int(s.value)
