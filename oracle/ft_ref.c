/* ft_ref.c — oracle restatement of the exact nearest-neighbour search that the
 * reference performs through ANN.dll with eps = 0 (main.pas:3830, 4027; FFI
 * extern.pas:63-67).  TEST INFRASTRUCTURE ONLY (see tmoracle.h).  PARITY UNPINNED.
 *
 * ANN (Mount & Arya; private float build shipped only as ANN.dll, source absent
 * from /root/reference, version not recoverable) returns, for eps = 0, the exact
 * k nearest rows under squared Euclidean distance, ascending.  A brute-force scan
 * therefore defines the same answer up to float rounding and tie order.
 */
#include "tmoracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

int tmo_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* torchrun exports OMP_NUM_THREADS=1 to its workers; the timed CPU baseline resets the team size
 * to the host's core count explicitly. */
void tmo_set_threads(int n) {
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}

/* insert (d,i) into ascending list of length k+1 (extra slot = next distance);
 * strict < keeps the lowest index among equal distances. */
static inline void ins_f64(double *bd, int32_t *bi, int k1, double d, int32_t i) {
    if (!(d < bd[k1 - 1])) return;
    int p = k1 - 1;
    while (p > 0 && d < bd[p - 1]) { bd[p] = bd[p - 1]; bi[p] = bi[p - 1]; p--; }
    bd[p] = d; bi[p] = i;
}

void tmo_knn_f64(const float *data, int64_t n, int dd, const float *q, int64_t nq,
                 int k, int32_t *idx, double *dist, double *next_dist) {
#pragma omp parallel for schedule(dynamic, 16)
    for (int64_t qi = 0; qi < nq; qi++) {
        double bd[65]; int32_t bi[65];
        int k1 = k + 1;
        for (int j = 0; j < k1; j++) { bd[j] = INFINITY; bi[j] = -1; }
        const float *qq = q + qi * dd;
        for (int64_t r = 0; r < n; r++) {
            const float *row = data + r * dd;
            double acc = 0.0;
            for (int j = 0; j < dd; j++) { double t = (double)qq[j] - (double)row[j]; acc += t * t; }
            ins_f64(bd, bi, k1, acc, (int32_t)r);
        }
        for (int j = 0; j < k; j++) { idx[qi * k + j] = bi[j]; dist[qi * k + j] = bd[j]; }
        if (next_dist) next_dist[qi] = bd[k];
    }
}

void tmo_knn_f32(const float *data, int64_t n, int dd, const float *q, int64_t nq,
                 int k, int32_t *idx, float *dist) {
#pragma omp parallel for schedule(dynamic, 16)
    for (int64_t qi = 0; qi < nq; qi++) {
        float bd[64]; int32_t bi[64];
        for (int j = 0; j < k; j++) { bd[j] = INFINITY; bi[j] = -1; }
        const float *qq = q + qi * dd;
        for (int64_t r = 0; r < n; r++) {
            const float *row = data + r * dd;
            volatile float acc = 0.0f; /* sequential, no reassociation */
            for (int j = 0; j < dd; j++) { float t = qq[j] - row[j]; acc = acc + t * t; }
            float d = acc;
            if (d < bd[k - 1]) {
                int p = k - 1;
                while (p > 0 && d < bd[p - 1]) { bd[p] = bd[p - 1]; bi[p] = bi[p - 1]; p--; }
                bd[p] = d; bi[p] = (int32_t)r;
            }
        }
        for (int j = 0; j < k; j++) { idx[qi * k + j] = bi[j]; dist[qi * k + j] = bd[j]; }
    }
}

/* Timed CPU baseline: the reference's DoFrameTiling loop (main.pas:4020-4038)
 * with the exact search done as a cache-blocked brute-force scan.  One query
 * block per thread (the reference threads per frame, main.pas:972); 8 queries
 * share each streamed dataset row so the compiler can vectorise over dd. */
#define QB 8
void tmo_nn1_blocked_f32(const float *data, int64_t n, int dd, const float *q,
                         int64_t nq, int32_t *idx, float *dist, int nthreads) {
#ifdef _OPENMP
    if (nthreads <= 0) nthreads = omp_get_max_threads();
#endif
#pragma omp parallel for schedule(dynamic, 4) num_threads(nthreads)
    for (int64_t q0 = 0; q0 < nq; q0 += QB) {
        int nb = (int)((nq - q0) < QB ? (nq - q0) : QB);
        float bd[QB]; int32_t bi[QB];
        for (int b = 0; b < QB; b++) { bd[b] = INFINITY; bi[b] = -1; }
        for (int64_t r = 0; r < n; r++) {
            const float *row = data + r * dd;
            for (int b = 0; b < nb; b++) {
                const float *qq = q + (q0 + b) * dd;
                float acc = 0.0f;
#pragma omp simd reduction(+ : acc)
                for (int j = 0; j < dd; j++) { float t = qq[j] - row[j]; acc += t * t; }
                if (acc < bd[b]) { bd[b] = acc; bi[b] = (int32_t)r; }
            }
        }
        for (int b = 0; b < nb; b++) { idx[q0 + b] = bi[b]; dist[q0 + b] = bd[b]; }
    }
}

/* ---- PrepareFrameTiling (main.pas:3791-3967) ----------------------------------------------
 * Stage 1 of a keyframe: which (palette, tile, mirror) rows enter the stage-2 dataset.
 * For every distinct (PalIdx, GlobalTileIndex) of the keyframe's tilemaps (UseOne, 3802-3853):
 *   the 8 nearest stage-1 rows of the tile's palette-index line (ann_kdtree_search_multi, 3830);
 *   hits whose err equals the previous hit's err are skipped (3835-3837);
 *   used[pal'][TRToTileIdx[idx]][TRToAttrs[idx]] := True for pal' = own palette (ftFast),
 *   every palette whose centroid distance to the own palette is < APalTol * HighestCorr
 *   (ftMedium, 3846, Corrs from BuildPaletteCorrTriangle 3855-3867), or all palettes (ftSlow).
 * ANN's order among equal distances is unspecified (kd-tree traversal); this restatement, like the
 * CUDA library, lists equal distances by ascending row index.
 * used: [npal][ntiles][4] indexed by attrs = hmir | vmir << 1.  Returns KNNSize (3942-3950). */
static double sq_dist_192(const double *a, const double *b) { /* CompareEuclideanDCTPtr main.pas:659-675 */
    double r = 0;
    for (int i = 0; i < TMO_DCT_SIZE; i++) { double d = a[i] - b[i]; r += d * d; }
    return r;
}

/* active: T^.Active per tile (NULL = all).  PrepareGlobalFT (main.pas:3763-3777) emits stage-1 rows for Active
 * tiles only and records TRToTileIdx; UseOne's query line is FTiles[GlobalTileIndex].PalPixels whatever the
 * tile's Active flag (3823-3828), and hits are mapped back through TRToTileIdx (3843-3850). */
int64_t tmo_prepare_frame_tiling_used_active(const uint8_t *palpix, int64_t ntiles, const uint8_t *active, int npal,
                                             const int32_t *item_pal, const int32_t *item_tile, int64_t nitems,
                                             int quality, double pal_tol, const double *pal_centroids,
                                             uint8_t *used) {
    static const uint8_t attrs_of[4] = {0, 1, 3, 2}; /* PrepareGlobalFT row order none,H,HV,V main.pas:3753 */
    int64_t nact = 0;
    int32_t *tr_tile = (int32_t *)malloc(sizeof(int32_t) * (size_t)(ntiles ? ntiles : 1));
    uint8_t *px = (uint8_t *)malloc((size_t)(ntiles ? ntiles : 1) * 64);
    for (int64_t t = 0; t < ntiles; t++)
        if (!active || active[t]) { memcpy(px + nact * 64, palpix + t * 64, 64); tr_tile[nact++] = (int32_t)t; }
    int64_t n = nact * 4;
    float *g = (float *)malloc(sizeof(float) * (size_t)(n ? n : 1) * 64);
    tmo_global_rows_f32(px, nact, g, NULL, NULL);
    memset(used, 0, (size_t)npal * ntiles * 4);
    uint8_t *seen = (uint8_t *)calloc((size_t)npal * ntiles, 1);
    double *corr = NULL, highest = 0.0;
    if (quality == 1) {
        corr = (double *)malloc(sizeof(double) * npal * npal);
        for (int j = 0; j < npal; j++)
            for (int i = 0; i < npal; i++) {
                double c = sq_dist_192(pal_centroids + (size_t)j * TMO_DCT_SIZE, pal_centroids + (size_t)i * TMO_DCT_SIZE);
                corr[j * npal + i] = c;
                if (!isnan(c) && c > highest) highest = c;
            }
    }
    for (int64_t it = 0; it < nitems; it++) {
        int pal = item_pal[it];
        int64_t tile = item_tile[it];
        if (seen[(size_t)pal * ntiles + tile]) continue;
        seen[(size_t)pal * ntiles + tile] = 1;
        int32_t idxs[8];
        float errs[8], line[64];
        for (int j = 0; j < 64; j++) line[j] = (float)palpix[tile * 64 + j]; /* Line = the stored tile's PalPixels */
        tmo_knn_f32(g, n, 64, line, 1, 8, idxs, errs);
        float last = INFINITY;
        for (int i = 0; i < 8; i++) {
            if (idxs[i] < 0) break; /* fewer than 8 rows */
            if (errs[i] == last) continue;
            last = errs[i];
            int64_t t2 = tr_tile[idxs[i] >> 2];
            int a2 = attrs_of[idxs[i] & 3];
            for (int p = 0; p < npal; p++) {
                int on = quality == 0 ? (p == pal)
                       : quality == 1 ? (corr[p * npal + pal] < pal_tol * highest)
                                      : 1;
                if (on) used[((size_t)p * ntiles + t2) * 4 + a2] = 1;
            }
        }
    }
    int64_t cnt = 0;
    for (size_t i = 0; i < (size_t)npal * ntiles * 4; i++) cnt += used[i];
    free(g); free(seen); free(corr); free(tr_tile); free(px);
    return cnt;
}

int64_t tmo_prepare_frame_tiling_used(const uint8_t *palpix, int64_t ntiles, int npal,
                                      const int32_t *item_pal, const int32_t *item_tile, int64_t nitems,
                                      int quality, double pal_tol, const double *pal_centroids,
                                      uint8_t *used) {
    return tmo_prepare_frame_tiling_used_active(palpix, ntiles, NULL, npal, item_pal, item_tile, nitems, quality, pal_tol,
                                                pal_centroids, used);
}

/* DoPsyV main.pas:3883-3919: rows in order pal, tile, vmir, hmir; effective mirror = flag xor tile flag */
void tmo_frame_tiling_rows(const uint8_t *used, const uint8_t *palpix, int64_t ntiles, const uint8_t *tile_flags,
                           const int32_t *pals, int npal, int pal_size, int use_wavelets, int gamma,
                           float *rows, int32_t *tr_tile, int32_t *tr_pal, uint8_t *tr_attrs) {
    int64_t di = 0;
    double dct[TMO_DCT_SIZE];
    for (int p = 0; p < npal; p++)
        for (int64_t t = 0; t < ntiles; t++)
            for (int vm = 0; vm < 2; vm++)
                for (int hm = 0; hm < 2; hm++) {
                    if (!used[((size_t)p * ntiles + t) * 4 + ((vm << 1) | hm)]) continue;
                    int fl = tile_flags ? tile_flags[t] : 0;
                    tmo_psyv(NULL, palpix + t * 64, pals + (size_t)p * pal_size, 1, use_wavelets, 0,
                             hm ^ (fl & 1), vm ^ ((fl >> 1) & 1), gamma, dct);
                    for (int j = 0; j < TMO_DCT_SIZE; j++) rows[di * TMO_DCT_SIZE + j] = (float)dct[j];
                    tr_tile[di] = (int32_t)t;
                    tr_pal[di] = p;
                    tr_attrs[di] = (uint8_t)(hm | (vm << 1));
                    di++;
                }
}

/* ---- Reindex: btnReindexClick main.pas:1199-1230 + ReindexTiles 4483-4527 ------------------------ */
typedef struct { int32_t use, tmp; } ri_tile;
static int ri_cmp(const void *a, const void *b) { /* CompareTileUseCountRev main.pas:4472-4481 */
    const ri_tile *x = (const ri_tile *)a, *y = (const ri_tile *)b;
    if (x->use != y->use) return y->use > x->use ? 1 : -1;
    return (x->tmp > y->tmp) - (x->tmp < y->tmp);
}
int64_t tmo_reindex(int32_t *tilemap, int64_t nitems, int64_t ntiles, int32_t *use_count, int32_t *idx_map,
                    int32_t *order) {
    for (int64_t t = 0; t < ntiles; t++) use_count[t] = 0;
    for (int64_t i = 0; i < nitems; i++) use_count[tilemap[i]]++;           /* 1214-1221 */
    ri_tile *act = (ri_tile *)malloc(sizeof(ri_tile) * (size_t)(ntiles ? ntiles : 1));
    int64_t cnt = 0;
    for (int64_t t = 0; t < ntiles; t++)                                     /* pack, 4497-4507 */
        if (use_count[t] > 0) { act[cnt].use = use_count[t]; act[cnt].tmp = (int32_t)t; cnt++; }
    qsort(act, (size_t)cnt, sizeof(ri_tile), ri_cmp);                        /* 4516 */
    for (int64_t t = 0; t < ntiles; t++) { idx_map[t] = -1; order[t] = -1; } /* 4510 */
    for (int64_t r = 0; r < cnt; r++) { idx_map[act[r].tmp] = (int32_t)r; order[r] = act[r].tmp; }
    for (int64_t i = 0; i < nitems; i++) tilemap[i] = idx_map[tilemap[i]];   /* 4523-4526 */
    free(act);
    return cnt;
}
