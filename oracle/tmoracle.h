/* tmoracle.h — CPU oracle for the TileMotion FrameTiling / KModes hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ may be linked, imported or
 * executed by the product path (tiler_b200/, libtmann.so).  Only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
 * use it, and only as the checker or the timed CPU baseline.
 *
 * PARITY UNPINNED: the reference (b0nefish/tiler) ships no tests, fixtures or
 * golden vectors for this path and cannot be built here (FreePascal + Win32 +
 * ANN.dll binary).  This is a restatement of the cited Pascal code; the
 * known-answer vectors in tests/ are derived from that code by hand.
 *
 * All citations are file:line under /root/reference.
 */
#ifndef TMORACLE_H
#define TMORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define TMO_TILE_W 8
#define TMO_TILE_PX 64
#define TMO_DCT_SIZE 192        /* cTileDCTSize main.pas:38 */
#define TMO_KM_FEATURES 80      /* cKModesFeatureCount kmodes.pas:15 */
#define TMO_KM_SHIFT 11         /* cDissimSubMatchingSize kmodes.pas:16 */

/* ---- psyv.c ---------------------------------------------------------- */
/* RGBToYUV main.pas:2656-2679. gamma: -1 linear, 0 -> 2.0, 1 -> 0.6 (main.pas:586) */
void tmo_rgb_to_yuv(int r, int g, int b, int gamma, double *y, double *u, double *v);
/* WaveletGS main.pas:2805-2840 (dx=dy=8, depth=2, stride 8) */
void tmo_wavelet_gs(const double *in, double *out, int dx, int dy, int depth);
/* ComputeTilePsyVisFeatures main.pas:2997-3177 with UseLAB=False.
 * rgb: 64 int32 $00BBGGRR (used when from_pal==0); palpix: 64 bytes + pal[]
 * (used when from_pal!=0). out: 192 doubles. */
void tmo_psyv(const int32_t *rgb, const uint8_t *palpix, const int32_t *pal,
              int from_pal, int use_wavelets, int qweighting, int hmirror,
              int vmirror, int gamma, double *out);
/* batch helpers (float32 output = what crosses the FFI, main.pas:3913, 4024) */
void tmo_psyv_queries_f32(const int32_t *rgb, int64_t nq, int use_wavelets,
                          int gamma, float *out);
/* stage-2 dataset rows for one palette: every tile x (vmir,hmir) in the row
 * order of DoPsyV main.pas:3902-3905; attrs = hmir | vmir<<1. */
void tmo_psyv_tileset_f32(const uint8_t *palpix, int64_t ntiles,
                          const int32_t *pal, int use_wavelets, int gamma,
                          float *out /* [ntiles*4][192] */);
void tmo_hmirror(uint8_t *palpix);   /* main.pas:3195-3209 */
void tmo_vmirror(uint8_t *palpix);   /* main.pas:3179-3193 */
/* PrepareTileMirrors main.pas:4049-4069: returns hflag | vflag<<1, flips in place */
int  tmo_prepare_tile_mirrors(uint8_t *palpix);
/* PrepareGlobalFT main.pas:3736-3780: rows (none,H,HV,V), attrs (0,1,3,2) */
void tmo_global_rows_f32(const uint8_t *palpix, int64_t ntiles, float *out /*[4n][64]*/,
                         int32_t *tr_tile, uint8_t *tr_attrs);

/* ---- ft_ref.c -------------------------------------------------------- */
/* exact k-NN, squared L2, double accumulation over float32 rows; ties -> lowest
 * index.  Also returns the (k+1)-th distance in gap2 (may be NULL) so tests
 * can build the epsilon-gap set. */
void tmo_knn_f64(const float *data, int64_t n, int dd, const float *q, int64_t nq,
                 int k, int32_t *idx, double *dist, double *next_dist);
/* float32 sequential-dimension accumulation (what ANN's annDist loop does) */
void tmo_knn_f32(const float *data, int64_t n, int dd, const float *q, int64_t nq,
                 int k, int32_t *idx, float *dist);
/* timed CPU baseline: blocked brute-force 1-NN, float32, OpenMP over queries */
void tmo_nn1_blocked_f32(const float *data, int64_t n, int dd, const float *q,
                         int64_t nq, int32_t *idx, float *dist, int nthreads);
/* PrepareFrameTiling main.pas:3791-3967: used[npal][ntiles][4] table (returns KNNSize), then the
 * stage-2 rows + side tables in DoPsyV order.  quality: 0 ftFast, 1 ftMedium, 2 ftSlow. */
int64_t tmo_prepare_frame_tiling_used(const uint8_t *palpix, int64_t ntiles, int npal,
                                      const int32_t *item_pal, const int32_t *item_tile, int64_t nitems,
                                      int quality, double pal_tol, const double *pal_centroids,
                                      uint8_t *used);
int64_t tmo_prepare_frame_tiling_used_active(const uint8_t *palpix, int64_t ntiles, const uint8_t *active, int npal,
                                             const int32_t *item_pal, const int32_t *item_tile, int64_t nitems,
                                             int quality, double pal_tol, const double *pal_centroids,
                                             uint8_t *used);
void tmo_frame_tiling_rows(const uint8_t *used, const uint8_t *palpix, int64_t ntiles, const uint8_t *tile_flags,
                           const int32_t *pals, int npal, int pal_size, int use_wavelets, int gamma,
                           float *rows, int32_t *tr_tile, int32_t *tr_pal, uint8_t *tr_attrs);
/* Reindex main.pas:1199-1230, 4483-4527: rewrites tilemap in place, returns the active tile count */
int64_t tmo_reindex(int32_t *tilemap, int64_t nitems, int64_t ntiles, int32_t *use_count, int32_t *idx_map,
                    int32_t *order);
int  tmo_max_threads(void);
void tmo_set_threads(int n);

/* ---- kdtree_ref.c ---------------------------------------------------
 * ANN's published kd-tree (standard split, bucket size bs, incremental box distance, leaf scan
 * with early exit) as created by ann_kdtree_create(pa, n, dd, 1, ANN_KD_STD) main.pas:3779, 3961
 * and searched with eps = 0 (main.pas:3830, 4027).  data is not copied. */
struct tmo_kdtree;
struct tmo_kdtree *tmo_kdtree_create(const float *data, int64_t n, int dd, int bs);
void tmo_kdtree_destroy(struct tmo_kdtree *t);
void tmo_kdtree_search(const struct tmo_kdtree *t, const float *q, int64_t nq, int k, int32_t *idx,
                       float *dist, int nthreads, int64_t *stats);

/* ---- kmodes_ref.c ---------------------------------------------------- */
/* semantics: 0 = x86-64 asm (kmodes.pas:357-412), 1 = generic (kmodes.pas:239-250) */
uint64_t tmo_dissim(const uint8_t *a, const uint8_t *b, int semantics);
/* GetMinMatchingDissim kmodes.pas:316-453 / 252-270: last index wins ties */
int32_t tmo_min_dissim(const uint8_t *rows, int64_t count, const uint8_t *item,
                       int semantics, uint64_t *best);
/* assignment pass kmodes.pas:984-997 */
void tmo_kmodes_assign(const uint8_t *X, int64_t npoints, const uint8_t *centroids,
                       int64_t k, int semantics, int32_t *clust, uint64_t *dis,
                       int nthreads);
/* UpdateMinDistance kmodes.pas:455-596: strict < */
void tmo_update_min_distance(const uint8_t *X, int64_t npoints, const uint8_t *center,
                             int semantics, uint64_t *mindist);
/* InitFarthestFirst kmodes.pas:698-776 */
void tmo_farthest_first(const uint8_t *X, int64_t npoints, int64_t k, int64_t init_point,
                        int semantics, uint8_t *centroids, int64_t *chosen);
/* representative pick of DoKModes main.pas:4231-4253 */
void tmo_kmodes_representatives(const uint8_t *X, int64_t npoints, const int32_t *labels,
                                const uint8_t *centroids, int64_t k, int semantics, int32_t *rep, int32_t *count);
/* ComputeKModes kmodes.pas:917-1060 (ANumInit <= 0: one run from start_point) incl. KModesIter 845-909 */
int tmo_compute_kmodes(const uint8_t *X, int64_t npoints, int64_t k, int64_t start_point, int nmod, int semantics,
                       int max_iter, int32_t *labels, uint8_t *centroids, uint64_t *cost_out, int64_t *moves_out);
/* WriteTileDatasetLine main.pas:4167-4183 + GetTilePalZoneThres 4142-4165 */
int  tmo_kmodes_row(const uint8_t *palpix, int pal_size, uint8_t *row80);

/* ---- smooth_ref.c ---------------------------------------------------- */
/* btnSmoothClick main.pas:1336-1370 + DoTemporalSmoothing main.pas:4071-4119; attrs bit0 H, bit1 V, bit2 Smoothed;
 * src_frame = frame whose original TileMap item the smoothed item is a copy of.  Returns -1 on bad indices. */
int tmo_temporal_smoothing(const uint8_t *palpix, int64_t ntiles, const int32_t *pals, int nkf, int npal,
                           int pal_size, const int32_t *frame_kf, int64_t nframes, int64_t npos,
                           const int32_t *tile, const int32_t *pal, const uint8_t *attrs, double strength,
                           int32_t *src_frame, int32_t *out_tile, int32_t *out_pal, uint8_t *out_attrs);

/* ---- gen.c ----------------------------------------------------------- */
/* RandInt kmodes.pas:82-86 */
uint32_t tmo_randint(uint32_t range, uint32_t *seed);
void tmo_gen_palette(uint32_t *seed, int pal_size, int32_t *pal);
void tmo_gen_tileset(uint32_t *seed, int64_t ntiles, int pal_size, uint8_t *palpix);
/* queries: planted (tile,mirror) rendered through pal + noise, 1/16 pure random */
void tmo_gen_queries(uint32_t *seed, int64_t nq, const uint8_t *palpix, int64_t ntiles,
                     const int32_t *pal, int noise, int32_t *rgb);

#ifdef __cplusplus
}
#endif
#endif
