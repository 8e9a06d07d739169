"""TKModes.ComputeKModes (kmodes.pas:917-1060) written a second time, in numpy, from the Pascal text - NOT from
oracle/kmodes_ref.c - with the assignment pass injected: the CPU tests plug the oracle's arg-min in and compare the
whole run with tmo_compute_kmodes (two independently written loops must agree on labels, centroids, iteration count,
cost and moves); the GPU test plugs the library's device-resident KModes handle in, one call per 960-point bin."""
import numpy as np

BIN = 960            # cBinSize kmodes.pas:847
SEED0 = 0x42381337   # kmodes.pas:930


def _randint(rng, seed):  # kmodes.pas:82-86
    seed = (seed * 0x08088405 + 1) & 0xFFFFFFFF
    return (seed * rng) >> 32, seed


class Loop:
    def __init__(self, X, k, nmod):
        self.X = np.ascontiguousarray(X, np.uint8)
        self.n, self.k, self.nmod = len(X), k, nmod
        self.memb = np.full(self.n, -1, np.int32)
        self.freq = np.zeros((k, 80, nmod), np.int32)
        self.cent = None
        self.ar = np.arange(80)
        self.touched = set()  # centroids edited since the last assignment call

    def move_point(self, i, to, frm):  # MovePointCat kmodes.pas:778-806, all 80 attributes at once (they are independent)
        p = self.X[i].astype(np.int64)
        self.memb[i] = to
        tc, fc = self.freq[to], self.freq[frm]
        tc[self.ar, p] += 1
        upd = tc[self.ar, self.cent[to]] < tc[self.ar, p]
        self.cent[to][upd] = p[upd]
        fc[self.ar, p] -= 1
        m = self.cent[frm] == p
        if m.any():
            self.cent[frm][m] = np.argmax(fc[m], axis=1)  # GetMaxValueIndex: first maximum
        self.touched.update((int(to), int(frm)))


def compute_kmodes(X, k, start_point, nmod, farthest_first, assign_all, assign_bin, max_iter=0, on_centroids=None):
    """farthest_first(start) -> centroids[k][80]; assign_all(centroids) -> labels; assign_bin(lo, hi, loop) ->
    (clust, dis) for points [lo, hi) against loop.cent (loop.touched = rows changed since the previous call).
    on_centroids(centroids): called once when the initial modes are known."""
    L = Loop(X, k, nmod)
    seed = SEED0
    L.cent = np.array(farthest_first(start_point), np.uint8, copy=True)
    L.memb[:] = assign_all(L.cent)
    for a in range(80):  # kmodes.pas:999-1001
        np.add.at(L.freq[:, a, :], (L.memb, L.X[:, a]), 1)
    counts = np.bincount(L.memb, minlength=k)
    for c in range(k):   # kmodes.pas:1003-1016
        if counts[c] == 0:
            for a in range(80):
                r, seed = _randint(L.n, seed)
                L.cent[c, a] = L.X[r, a]
        else:
            L.cent[c] = np.argmax(L.freq[c], axis=1)
    if on_centroids:
        on_centroids(L.cent)
    L.touched.clear()
    itr, converged, cost, total = 0, False, (1 << 64) - 1, 0
    max_iter = max_iter if max_iter > 0 else (1 << 31) - 1
    while itr <= max_iter and not converged:
        itr += 1
        moves, acc = 0, 0
        for b0 in range(0, L.n, BIN):  # KModesIter kmodes.pas:845-909
            b1 = min(L.n, b0 + BIN)
            clust, dis = assign_bin(b0, b1, L)
            L.touched.clear()
            for i in range(b0, b1):
                acc += int(dis[i - b0])
                if L.memb[i] == clust[i - b0]:
                    continue
                moves += 1
                old = int(L.memb[i])
                L.move_point(i, int(clust[i - b0]), old)
                if not (L.memb == old).any():  # CountClusterMembers(old) = 0
                    cnt = np.bincount(L.memb, minlength=k)
                    frm = int(k - 1 - np.argmax(cnt[::-1]))  # GetMaxClusterMembers: `>=`, the LAST largest cluster
                    choices = np.flatnonzero(L.memb == frm)
                    r, seed = _randint(len(choices), seed)
                    L.move_point(int(choices[r]), old, frm)
        converged = acc >= cost or moves == 0
        cost = acc
        total += moves
    return L.memb.copy(), L.cent.copy(), itr, cost, total
