/* tmann.h — C ABI of libtmann.so, the B200 (sm_100a) replacement for the k-NN
 * boundary of GliGli's TileMotion encoder (b0nefish/tiler).
 *
 * Part 1 re-exports, bit-compatibly, the five symbols the encoder imports from
 * ANN.dll (reference extern.pas:63-67, types extern.pas:21-39).  Part 2 adds
 * batch / fused entry points so the Pascal side can hand a whole frame (or
 * keyframe) to the GPU per call; each cites the reference code it replaces.
 *
 * Conventions: plain pointers and sizes only.  Pointers named h_* are host
 * memory, d_* are device (cudaMalloc) memory on the library's current device.
 * Part 2 calls return 0 on success or a negative TM_E* code and never abort;
 * tm_last_error() gives the message.  Part 1 keeps the legacy shape (no error
 * channel: NULL handle / index -1 on failure).  There is no CPU fallback:
 * every call fails with TM_ENODEVICE when no sm_100 device is usable.
 *
 * Threads: every entry point may be called concurrently from any number of host
 * threads over shared handles, as the encoder does (DoFrameTiling runs one frame
 * per worker thread, main.pas:972).  Device work is serialised per GPU (one stream,
 * one lock per device); frame-sized tm_frametiling_search calls copy pageable caller
 * memory through per-thread pinned staging OUTSIDE that lock, so the host copies of
 * one caller overlap the device work of another.  tm_last_error() is per thread.
 * Destroying a handle while another thread still uses it is a caller error.
 *
 * Devices: ONE process can drive several GPUs of a box (tm_set_devices, or
 * TM_DEVICES=all / "0,1,.." in the environment).  A dataset is then built once on
 * the first device and broadcast to the others over NCCL (NVLink / NVSwitch); the
 * handle the caller holds stands for all replicas.  Clip-sized host batches are cut
 * into one contiguous query shard per GPU, every GPU returns its own results into
 * the caller's buffers; frame-sized and single-query calls go to the least busy GPU,
 * so the encoder's worker threads spread over the devices without any change.
 */
#ifndef TMANN_H
#define TMANN_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#if defined(_WIN32)
#define TM_API __declspec(dllexport)
#else
#define TM_API __attribute__((visibility("default")))
#endif

/* ---------------------------------------------------------------- Part 1 */
/* Opaque handle; replaces PANNkdtree (extern.pas:36-39). */
typedef struct ANNkdtree ANNkdtree;

/* extern.pas:63 — pa: n row pointers of dd floats (rows need not be contiguous);
 * bs and split are accepted and ignored (the search is exact, like eps = 0).
 * Snapshots the rows to the device; call sites main.pas:3779 (dd=64), 3961 (dd=192). */
TM_API ANNkdtree *ann_kdtree_create(float **pa, int n, int dd, int bs, int split);
/* extern.pas:64 — NULL-safe; call sites main.pas:3784, 3983. */
TM_API void ann_kdtree_destroy(ANNkdtree *akd);
/* extern.pas:65 — exact 1-NN; returns row index (-1 on failure), *err (may be
 * NULL) = squared L2 in float32.  eps is ignored (always exact). main.pas:4027. */
TM_API int ann_kdtree_search(ANNkdtree *akd, float *q, float eps, float *err);
/* extern.pas:66 — same result as ann_kdtree_search (no call site in main.pas). */
TM_API int ann_kdtree_pri_search(ANNkdtree *akd, float *q, float eps, float *err);
/* extern.pas:67 — cnt nearest rows ascending by squared L2; returns 0 (the
 * reference's return value is garbage and ignored, main.pas:3830). */
TM_API int ann_kdtree_search_multi(ANNkdtree *akd, int *idxs, float *errs, int cnt,
                                   float *q, float eps);

/* ---------------------------------------------------------------- Part 2 */
enum {
    TM_OK = 0,
    TM_EINVAL = -1,    /* bad argument */
    TM_ENODEVICE = -2, /* no usable sm_100 GPU / CUDA failure at init */
    TM_ECUDA = -3,     /* CUDA runtime error (see tm_last_error) */
    TM_ENOMEM = -4
};

/* tm_psyv flags: the GUI options of btnDoFrameTilingClick main.pas:962-964 */
enum {
    TM_PSYV_WAVELETS = 1, /* UseWavelets (default on, main.lfm:372-379); 0 = DCT-II */
    TM_PSYV_QWEIGHT = 2   /* QWeighting (off on the FrameTiling path) */
};
/* KModes dissimilarity semantics */
enum {
    TM_KM_ASM = 0,    /* x86-64 asm that ships: kmodes.pas:357-412 */
    TM_KM_GENERIC = 1 /* documented generic form: kmodes.pas:239-250 */
};
/* search engines (tm_set_engine); AUTO picks the tensor-core path when dd % 16 == 0 */
enum { TM_ENGINE_AUTO = 0, TM_ENGINE_SIMT = 1, TM_ENGINE_TC = 2 };

TM_API const char *tm_last_error(void);
TM_API int tm_version(void);
/* Select the CUDA device for the calling process (default: current device, or
 * TM_DEVICES).  Same as tm_set_devices(&device, 1). */
TM_API int tm_set_device(int device);
/* Drive n GPUs from this process: devices[0] builds datasets, all hold replicas.  Brings up
 * one NCCL communicator over them (libnccl.so.2 is loaded at this point, not before).  May be
 * called again with another list only while no dataset handle is alive.  The reference fans
 * DoFrameTiling out over worker threads of ONE process on one shared handle (main.pas:972):
 * this is what lets that process reach every GPU of the box through the same five imports. */
TM_API int tm_set_devices(const int *devices, int n);
/* active device list (returns its length; fills at most cap entries) */
TM_API int tm_get_devices(int *devices, int cap);
/* wall time (ms) of the last NCCL communicator bring-up incl. its warm-up broadcast (0: none) */
TM_API double tm_comm_init_ms(void);
TM_API int tm_set_engine(int engine);
/* Run all library work on the caller's CUDA stream (a cudaStream_t).  From then on the
 * *_dev entry points only enqueue work on that stream and return without synchronising;
 * without it they synchronise the device on entry and their own stream on exit. */
TM_API int tm_set_stream(void *cuda_stream);
TM_API int tm_device_sync(void);

/* Page-locked host memory for buffers handed to the h_* entry points (optional: pageable
 * memory works, pinned memory lets the PCIe copies run asynchronously at full rate).
 * The Pascal side would allocate its per-frame TRGBPixels staging through these. */
TM_API void *tm_host_alloc(size_t bytes);
TM_API void tm_host_free(void *p);

/* Contiguous-row variant of ann_kdtree_create; rows may be host or device memory
 * (is_device != 0).  Replaces the per-row gather of main.pas:3757-3779. */
TM_API ANNkdtree *tm_dataset_create(const float *rows, int64_t n, int dd, int is_device);
TM_API int64_t tm_dataset_rows(const ANNkdtree *akd);
TM_API int tm_dataset_dim(const ANNkdtree *akd);
/* number of devices holding a replica of this dataset (1 without tm_set_devices) */
TM_API int tm_dataset_replicas(const ANNkdtree *akd);

/* Batch k-NN: nq queries x dd floats, contiguous; out idxs[nq*k], errs[nq*k]
 * ascending per query.  One call replaces the per-position loop of
 * DoFrameTiling main.pas:4020-4038 (k = 1) or UseOne main.pas:3830 (k = 8). */
TM_API int ann_kdtree_search_batch(ANNkdtree *akd, const float *h_q, int64_t nq, int k,
                                   int *h_idxs, float *h_errs);
TM_API int tm_search_batch_dev(ANNkdtree *akd, const float *d_q, int64_t nq, int k,
                               int *d_idxs, float *d_errs);

/* PsyV descriptors, replaces ComputeTilePsyVisFeatures main.pas:2997-3177
 * (UseLAB = False).  gamma: -1 linear, 0 or 1 = gGamma index (main.pas:586).
 * Queries: 64 int32 $00BBGGRR per tile (TRGBPixels main.pas:155), FromPal = False,
 * no mirror (main.pas:4023) -> out[nq][192] float32 (main.pas:4024-4025). */
TM_API int tm_psyv_queries(const int32_t *h_rgb, int64_t nq, int flags, int gamma, float *h_out);
TM_API int tm_psyv_queries_dev(const int32_t *d_rgb, int64_t nq, int flags, int gamma, float *d_out);

/* Stage-2 dataset straight from the tileset (DoPsyV main.pas:3883-3919 with every
 * (pal, tile, vmir, hmir) used): palpix[ntiles][64] palette indices, pals[npal][pal_size]
 * int32 colours, tile_flags[ntiles] = TTile.HMirror | VMirror<<1 (may be NULL = 0).
 * Row r = ((p*ntiles + t)*2 + vmir)*2 + hmir; TRToAttrs = hmir | vmir<<1 (main.pas:3910). */
TM_API ANNkdtree *tm_tileset_dataset_create(const uint8_t *h_palpix, int64_t ntiles,
                                            const int32_t *h_pals, int npal, int pal_size,
                                            const uint8_t *h_tile_flags, int flags, int gamma);
TM_API ANNkdtree *tm_tileset_dataset_create_dev(const uint8_t *d_palpix, int64_t ntiles,
                                                const int32_t *d_pals, int npal, int pal_size,
                                                const uint8_t *d_tile_flags, int flags, int gamma);
/* copy the dataset rows (float32) back, for inspection / parity tests; _replica reads the copy
 * held by the replica-th active device (what the NCCL broadcast delivered) */
TM_API int tm_dataset_read(const ANNkdtree *akd, float *h_rows);
TM_API int tm_dataset_read_replica(const ANNkdtree *akd, int replica, float *h_rows);

/* Fused FrameTiling hot loop (main.pas:4020-4038): RGB tiles -> descriptor ->
 * exact 1-NN in akd (dd must be 192).  out row index + squared-L2 err per query,
 * in the caller's query order.  (Batches of >= 65 536 queries are searched in a
 * colour-sorted order inside the library - similar tiles share the search kernel's
 * warps, see qorder.cu - and scattered back; TM_QORDER=0 in the environment
 * switches that off.  Results do not depend on it.) */
TM_API int tm_frametiling_search(ANNkdtree *akd, const int32_t *h_rgb, int64_t nq, int flags,
                                 int gamma, int *h_idx, float *h_err);
TM_API int tm_frametiling_search_dev(ANNkdtree *akd, const int32_t *d_rgb, int64_t nq, int flags,
                                     int gamma, int *d_idx, float *d_err);

/* Stage-1 rows of PrepareGlobalFT main.pas:3736-3780: for each tile 4 rows of 64
 * floats in order (none, H, H+V, V); tr_tile / tr_attrs (0,1,3,2) side tables. */
TM_API ANNkdtree *tm_global_dataset_create(const uint8_t *h_palpix, int64_t ntiles,
                                           int32_t *h_tr_tile, uint8_t *h_tr_attrs);
/* The same over a tile array with inactive entries (FTiles after MergeTiles and before Reindex):
 * rows exist for tiles with h_active[t] != 0 only (T^.Active, main.pas:3763-3777), tr_tile holds
 * indices into the WHOLE array, exactly like TRToTileIdx.  h_active = NULL: every tile is active.
 * tr_tile / tr_attrs must hold 4 x (number of active tiles) entries; tm_dataset_rows gives that. */
TM_API ANNkdtree *tm_global_dataset_create_active(const uint8_t *h_palpix, int64_t ntiles,
                                                  const uint8_t *h_active, int32_t *h_tr_tile,
                                                  uint8_t *h_tr_attrs);

/* One-call PrepareFrameTiling (main.pas:3791-3967) for a keyframe.  global_ds is the stage-1 handle
 * of tm_global_dataset_create(_active) over the same ntiles tiles (h_palpix is the WHOLE tile array,
 * inactive tiles included: tilemap items and the used[] table index it, main.pas:3823, 3902).  h_item_pal / h_item_tile list the
 * (PalIdx, GlobalTileIndex) of every tilemap position of the keyframe's frames (duplicates allowed, the
 * "already seen" test of main.pas:3814-3821 is done on the device).  For each distinct pair the 8
 * nearest stage-1 rows of the tile's line are taken (UseOne main.pas:3830; hits with an err equal to the
 * previous hit's are skipped, 3835-3837) and used[pal'][tile'][attrs'] is set for pal' = own palette
 * (quality 0, ftFast), palettes whose centroid distance to the own palette is < pal_tol * max
 * (quality 1, ftMedium; h_pal_centroids[npal][192] doubles = KF.PaletteCentroids, pal_tol =
 * cFTPaletteTol 0.05), or all palettes (quality 2, ftSlow).  The used rows are rendered in DoPsyV's
 * order (pal, tile, vmir, hmir; main.pas:3883-3919) and become the returned stage-2 dataset
 * (tm_dataset_rows = KNNSize); tm_keyframe_tables copies TRToTileIdx / TRToPalIdx / TRToAttrs. */
TM_API ANNkdtree *tm_keyframe_dataset_create(ANNkdtree *global_ds, const uint8_t *h_palpix, int64_t ntiles,
                                             const uint8_t *h_tile_flags, const int32_t *h_pals, int npal,
                                             int pal_size, const int32_t *h_item_pal, const int32_t *h_item_tile,
                                             int64_t nitems, int quality, double pal_tol,
                                             const double *h_pal_centroids, int flags, int gamma);
TM_API int tm_keyframe_tables(const ANNkdtree *akd, int32_t *h_tr_tile, int32_t *h_tr_pal, uint8_t *h_tr_attrs);

/* KModes assignment pass (DoGMMD kmodes.pas:825-835; loops 866-878, 984-997):
 * clust[i] = arg-min over k centroids of the matching dissimilarity, last index
 * wins ties (kmodes.pas:414-415); dis[i] (may be NULL) = that dissimilarity.
 * X[npoints][80], centroids[k][80] contiguous bytes. */
TM_API int tm_kmodes_assign(const uint8_t *h_X, int64_t npoints, const uint8_t *h_centroids,
                            int64_t k, int semantics, int32_t *h_clust, uint64_t *h_dis);
TM_API int tm_kmodes_assign_dev(const uint8_t *d_X, int64_t npoints, const uint8_t *d_centroids,
                                int64_t k, int semantics, int32_t *d_clust, uint64_t *d_dis);
/* Device-resident KModes state for the per-bin loop of KModesIter (kmodes.pas:845-909): the points are
 * uploaded once, the centroids stay on the device, the caller sends only the rows MovePointCat changed
 * (kmodes.pas:778-806) between two 960-point bins and gets one bin's (clust, dis) per call.  Same kernel,
 * same results as tm_kmodes_assign.  A bin-sized call allocates nothing, issues one or two kernel launches and no copy
 * command: results arrive in pinned host memory and the call polls a completion flag (TM_KM_POLL=0: copy + synchronise).
 * The handle remembers every point's last label and uses it as a bound for the next call on that point (a bound only:
 * results never depend on it). */
typedef struct tm_kmodes tm_kmodes;
TM_API tm_kmodes *tm_kmodes_open(const uint8_t *h_X, int64_t npoints, int semantics);
TM_API void tm_kmodes_close(tm_kmodes *km);
/* all k centroids [k][80] (after InitFarthestFirst / the initial mode computation, kmodes.pas:999-1016) */
TM_API int tm_kmodes_set_centroids(tm_kmodes *km, const uint8_t *h_centroids, int64_t k);
/* h_rows[m][80] replace the centroids h_idx[m] (duplicates: the last one wins like sequential stores) */
TM_API int tm_kmodes_update_centroids(tm_kmodes *km, const int32_t *h_idx, const uint8_t *h_rows, int64_t m);
/* DoGMMD for the points [lo, hi): h_clust[hi-lo], h_dis[hi-lo] (may be NULL) */
TM_API int tm_kmodes_assign_range(tm_kmodes *km, int64_t lo, int64_t hi, int32_t *h_clust, uint64_t *h_dis);
TM_API int tm_kmodes_read_centroids(tm_kmodes *km, uint8_t *h_centroids);

/* Allocation-free variant of tm_kmodes_assign_dev.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  It calls no cudaMalloc / cudaFree / allocating library routine and never
 * synchronises: two kernels are enqueued on cuda_stream (a cudaStream_t; NULL = the default stream) of the
 * CURRENT device, results are ready when that stream reaches them.  Its scratch (bit-sliced centroids + one
 * flag word) is a fixed device array inside the library: k <= TM_NOALLOC_MAX_CENTROIDS, larger k is rejected
 * with TM_EINVAL (tm_last_error says so); npoints is not limited.  The array costs 768.25 KB on every GPU
 * that loads the library, until exit.  All calls on one GPU share it: calls that overlap on the GPU (streams
 * without an ordering between them) race; calls on one stream, or ordered by events, are safe.  The first
 * call on a device resolves the array's device address (cudaGetSymbolAddress): make one call outside a graph
 * capture first.  Same results as tm_kmodes_assign_dev, bit for bit. */
#define TM_NOALLOC_MAX_CENTROIDS 16384
TM_API int tm_kmodes_assign_noalloc(const uint8_t *d_X, int64_t npoints, const uint8_t *d_centroids, int64_t k,
                                    int semantics, int32_t *d_clust, uint64_t *d_dis, void *cuda_stream);

/* UpdateMinDistance (kmodes.pas:455-596): mindist[i] = min(mindist[i], dissim(center, X[i])) */
TM_API int tm_kmodes_update_min_distance(const uint8_t *h_X, int64_t npoints,
                                         const uint8_t *h_center, int semantics,
                                         uint64_t *h_mindist);

/* InitFarthestFirst (kmodes.pas:698-776) on the device: centroid 0 = X[init_point]; centroid c = the
 * LAST unused point whose minimum dissimilarity to the chosen ones is largest (`>=`, 765-770).
 * h_centroids[k][80] out, h_chosen[k] (may be NULL) = the chosen point indices.  k <= npoints. */
TM_API int tm_kmodes_farthest_first(const uint8_t *h_X, int64_t npoints, int64_t k, int64_t init_point,
                                    int semantics, uint8_t *h_centroids, int64_t *h_chosen);
/* Cluster representatives (DoKModes main.pas:4231-4253): h_rep[j] = index of the member of cluster j
 * (h_labels[i] == j, point order) with the smallest dissimilarity to h_centroids[j], last member among
 * equals (GetMinMatchingDissim kmodes.pas:414-415); -1 for an empty cluster.  h_count[j] (may be NULL)
 * = member count (the reference merges only clusters with >= 2 members, main.pas:4247). */
TM_API int tm_kmodes_representatives(const uint8_t *h_X, int64_t npoints, const int32_t *h_labels,
                                     const uint8_t *h_centroids, int64_t k, int semantics, int32_t *h_rep,
                                     int32_t *h_count);

/* Reindex, the consumer of FrameTiling's output (btnReindexClick main.pas:1199-1230 + ReindexTiles
 * 4483-4527).  tilemap[nitems] = GlobalTileIndex of every tilemap item, rewritten in place to the new
 * indices.  use_count[ntiles] = UseCount per old tile; idx_map[ntiles] = old -> new index (-1 = inactive,
 * the $ffffffff of main.pas:4510); order[ntiles] = new -> old for the first *h_active entries (-1 after):
 * active tiles by UseCount descending, then old index ascending (CompareTileUseCountRev 4472-4481).
 * The output arrays may be NULL in the host variant. */
TM_API int tm_reindex(int32_t *h_tilemap, int64_t nitems, int64_t ntiles, int32_t *h_use_count,
                      int32_t *h_idx_map, int32_t *h_order, int64_t *h_active);
TM_API int tm_reindex_dev(int32_t *d_tilemap, int64_t nitems, int64_t ntiles, int32_t *d_use_count,
                          int32_t *d_idx_map, int32_t *d_order, int64_t *h_active);

/* KModes rows on the device (WriteTileDatasetLine main.pas:4167-4183 + GetTilePalZoneThres 4142-4165):
 * h_rows[ntiles][80] = 64 palette indices + 16 zone flags; h_signi[ntiles] (may be NULL) = PalSigni. */
TM_API int tm_kmodes_rows(const uint8_t *h_palpix, int64_t ntiles, int pal_size, uint8_t *h_rows, int32_t *h_signi);
/* DoFindBest of ReloadPreviousTiling (main.pas:4377-4409): like tm_kmodes_assign, but point i only sees the
 * centroids (saved tiles) j with h_cbin[j] == h_xbin[i] (bins = PalSigni, 0 <= bin < nbins), in their
 * original order; a point whose bin holds no centroid sees all of them.  clust = index into h_C. */
TM_API int tm_kmodes_assign_binned(const uint8_t *h_X, int64_t npoints, const int32_t *h_xbin, const uint8_t *h_C,
                                   int64_t k, const int32_t *h_cbin, int nbins, int semantics, int32_t *h_clust,
                                   uint64_t *h_dis);

/* Temporal smoothing of a whole clip: btnSmoothClick (main.pas:1336-1370) + DoTemporalSmoothing (main.pas:4071-4119).
 * h_palpix[ntiles][64] = FTiles[].PalPixels; h_pals[nkf][npal][pal_size] = PaletteRGB of every keyframe;
 * h_frame_kf[nframes] = keyframe of every frame; h_tile / h_pal / h_attrs [nframes][npos] = TileMap (GlobalTileIndex,
 * PalIdx, attrs bit0 HMirror, bit1 VMirror, bit2 Smoothed); strength = seTempoSmoo / 1000.  Outputs (each may be
 * NULL) [nframes][npos]: SmoothedTileMap fields, and h_src_frame = the frame whose original TileMap item the
 * smoothed item is a copy of (the Pascal side can copy the whole record, TmpIndex included, from there). */
TM_API int tm_temporal_smoothing(const uint8_t *h_palpix, int64_t ntiles, const int32_t *h_pals, int nkf, int npal,
                                 int pal_size, const int32_t *h_frame_kf, int64_t nframes, int64_t npos,
                                 const int32_t *h_tile, const int32_t *h_pal, const uint8_t *h_attrs, double strength,
                                 int32_t *h_src_frame, int32_t *h_out_tile, int32_t *h_out_pal, uint8_t *h_out_attrs);

/* counters since the last reset: kernel launches issued by this library, queries
 * that took the exact-fallback path after the tensor-core pass, and the device
 * time (ms, CUDA events) of the last search call split by phase. */
typedef struct tm_stats {
    int64_t kernel_launches;
    int64_t queries;
    int64_t fallback_queries;
    /* device time (ms, CUDA events on the library's stream) accumulated since the last
     * reset while tm_set_profiling(1) is on: descriptor kernels, whole search phase,
     * and inside it the tensor-core coarse kernel and the FP32 re-rank. */
    float desc_ms;
    float search_ms;
    float coarse_ms;
    float rerank_ms;
    int64_t coarse_launches;
    /* audit of the tensor-core engine's error model while profiling is on: the largest
     * |coarse score + ||q16||^2 - exact distance| seen on a re-ranked candidate, divided by the bound the
     * certificate assumes for it (2 e sqrt(d) + e^2 + a, csrc/search_tc.h).  Must stay below 1. */
    float max_error_ratio;
    int64_t audited_candidates;
} tm_stats;
/* tm_set_profiling: 0 off, 1 phase timing, 2 phase timing + error-model audit (slows the re-rank).
 * tm_get_stats synchronises the library's stream when profiling is on. */
TM_API int tm_set_profiling(int on);
TM_API int tm_get_stats(tm_stats *out);
TM_API int tm_reset_stats(void);

#ifdef __cplusplus
}
#endif
#endif
