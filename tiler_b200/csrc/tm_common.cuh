// tm_common.cuh — shared declarations of libtmann.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <stdio.h>
#include <atomic>
#include <mutex>
#include <vector>
#include "../../include/tmann.h"

#define TM_DCT 192
#define TM_KM_FEATURES 80
#define TM_MAX_DEV 16

// Dataset handle behind the opaque ANNkdtree*.  Layout in HBM:
//   rows_f32 [n][dd]      float32, exact copy of what crossed the FFI (re-rank / SIMT path)
//   rows_f16 [n_pad][kp]  fp16 operand of the tensor-core pass: mean-centred row, then the
//                         fp32 ||t-mu||^2 split in three fp16 parts, zero padded to kp
//   norms    [n]          float32 ||t-mu||^2 (fp32 from the unrounded centred row)
//   rows_f16s [n_pad][kp] the same operand in the k = 1 kernel's SCAN ORDER (search_tc.cu: rows sorted by norm, cut
//                         into 256-row tiles, tiles dealt in a strided order), norms_s [n_pad] their ||t~||^2
//                         (+inf for padding rows), perm [n_pad] scan position -> dataset row (-1 for padding)
//   mean     [dd]         float32 column mean (distance-invariant centring)
struct ANNkdtree {
    uint32_t magic;
    int64_t n;
    int dd;
    int device;
    float *rows_f32;
    __half *rows_f16;
    float *norms;
    __half *rows_f16s;
    float *norms_s;
    int *perm;
    __half *norm_tiles; // per 128-row tile: [||t~||^2 / S split in 3 fp16][zeros] as a no-swizzle UMMA operand image
    float norm_scale;   // S
    float *mean;
    int64_t n_pad;
    int kp;
    float max_norm;  // max ||t-mu||
    float max_dnorm; // max ||(t-mu) - fp16(t-mu)||: the rows' share of the coarse pass's rounding error
    void *tmap_rows; // CUtensorMap (host copy) for rows_f16
    std::mutex *single_mu; // serialises the legacy one-query entry points' scratch
    float *h_pin;          // pinned scratch for the legacy one-query entry points
    float *d_q1;
    int *d_i1;
    float *d_e1;
    // stage-1 (global, dd = 64) datasets: cached 8-NN of every stored tile's own line (UseOne main.pas:3830)
    int *knn8_idx;
    float *knn8_err;
    // keyframe (stage-2) datasets built by tm_keyframe_dataset_create: row -> (tile, pal, attrs) side tables
    int32_t *h_tr_tile, *h_tr_pal;
    uint8_t *h_tr_attrs;
    // stage-1 datasets over a tile list with inactive entries (tm_global_dataset_create_active): stage-1 row / 4 ->
    // index into the caller's tile array (TRToTileIdx main.pas:3763-3777); NULL = every tile is active
    int32_t *d_tr_tile;
    int64_t ntiles_all; // length of the caller's tile array (stage-1 datasets)
    // multi-device (tm_set_devices): rep[0] = this handle on the first active device, rep[i] = its replica on the
    // i-th; replicas are filled by one NCCL broadcast when the dataset is created.  owner = NULL on the primary.
    ANNkdtree *rep[TM_MAX_DEV];
    int nrep;
    ANNkdtree *owner;
    std::atomic<unsigned> *rr; // round-robin cursor for calls that go to one replica
};
#define TM_MAGIC 0x544d414eu

void tm_set_error(const char *fmt, ...);
int tm_check(cudaError_t e, const char *what);
#define TM_CUDA(x)                                  \
    do {                                            \
        int _r = tm_check((x), #x);                 \
        if (_r) return _r;                          \
    } while (0)

extern std::atomic<int64_t> g_launches, g_queries, g_fallbacks, g_coarse_launches;
extern int g_engine;
static inline void tm_count_launch(int n = 1) { g_launches.fetch_add(n, std::memory_order_relaxed); }

// ---- per-device context -------------------------------------------------------------------------------------
// One per CUDA device the library drives (tm_set_devices; default: the calling thread's current device).  All
// device work of one GPU goes through its stream and its grow-only scratch under its lock, so callers on
// different GPUs never wait for each other.  An entry point binds the calling thread to a device for the
// duration of the call (TmApiGuard in api.cu); everything below the C ABI reaches the context through tm_dev().
struct TmProfSpan { int tag; cudaEvent_t a, b; };
struct TmQScratch { // descriptor rows of a query batch (fused FrameTiling path) + coherence-order buffers (qorder.cu)
    float *desc = nullptr;
    size_t cap = 0;
    uint32_t *keys = nullptr;
    int *vals = nullptr, *perm = nullptr, *inv = nullptr, *idx_s = nullptr;
    float *err_s = nullptr;
    void *sort_tmp = nullptr;
    size_t sort_tmp_bytes = 0, cap_order = 0;
};
struct TmHostPath { // staging of the host-buffer entry points: two input slots, result buffers, a copy stream
    void *in[2] = {nullptr, nullptr};
    int *oi = nullptr;
    float *oe = nullptr;
    size_t cap_in = 0, cap_out = 0; // bytes per input slot, result elements
    cudaStream_t cs = nullptr;
    cudaEvent_t ev_in[2] = {nullptr, nullptr}, ev_done[2] = {nullptr, nullptr};
};
struct TmTcScratch { // tensor-core engine: query operand, norms, candidate lists, reject list
    __half *q16 = nullptr;
    float *qn16 = nullptr, *qne = nullptr, *qde = nullptr, *cand_s = nullptr, *cand_tau = nullptr;
    int *cand_i = nullptr, *fb_list = nullptr, *fb_count = nullptr;
    int *h_count = nullptr; // pinned
    int *chain_ready = nullptr; // chained plan: per query block, epilogue warps done (search_tc.cu)
    size_t cap_ready = 0;
    size_t cap_q = 0, cap_c = 0;
    int kp = 0;
};
// deferred handling of the queries the tensor-core certificate rejects (pipelined host calls, api_combine.inc)
struct TmDefer { int *d_list; int *h_count; int64_t cap; };
struct TmCombiner; // api_combine.inc: per-device worker that coalesces concurrent frame-sized / single-query host calls
struct TmKmScratch { // KModes: bit-sliced centroids + wide flag, partial arg-min of a centroid-split launch (grow-only)
    uint32_t *cb = nullptr;
    int *flag = nullptr;
    size_t cap = 0;
    int32_t *part_i = nullptr;
    uint32_t *part_d = nullptr;
    size_t cap_part = 0;
};
struct TmSimtScratch { float *d = nullptr; int *i = nullptr; size_t cap = 0; };
struct TmDevice {
    int device = -1, sms = 0;
    cudaStream_t stream = nullptr, own_stream = nullptr;
    bool user_stream = false;
    std::recursive_mutex mu; // serialises this device's entry points (recursive: entry points call each other)
    TmQScratch qs;
    TmHostPath hp;
    TmTcScratch tc;
    TmSimtScratch simt;
    TmKmScratch km;
    bool tc_coherent = false; // the next searches' queries are in coherence order (qorder.cu)
    TmDefer *defer = nullptr; // set around a search whose rejected queries the caller handles after completion
    std::atomic<TmCombiner *> combiner{nullptr};
    std::mutex stats_mu;
    std::vector<TmProfSpan> spans;
    float ms[4] = {0, 0, 0, 0};
    unsigned long long *audit = nullptr; // device: [0] max error ratio (float bits), [1] audited candidates
    std::atomic<int> inflight{0};        // host calls routed to this device and not yet returned
    void *nccl_comm = nullptr;           // ncclComm_t of this device in the library's communicator (>= 2 devices)
};
TmDevice *tm_dev(); // the calling thread's device context (bound by the entry point)
static inline cudaStream_t tm_stream() { return tm_dev()->stream; }
static inline int tm_num_sms() { return tm_dev()->sms; }
// true the first time it is called for `mask` on the calling thread's device (per-device one-time setup such as
// cudaFuncSetAttribute, which applies to the current device only); callers hold the device lock
static inline bool tm_first_on_device(std::atomic<uint32_t> &mask) {
    const uint32_t bit = 1u << tm_dev()->device;
    return !(mask.fetch_or(bit, std::memory_order_relaxed) & bit);
}

// Stream-ordered device memory from the driver's pool (cudaMallocAsync on the device's library stream, release
// threshold unlimited): dataset buffers and per-call temporaries are recycled without cudaMalloc / cudaFree round
// trips (cudaFree also synchronises the device).  Only for memory used on that stream.
cudaError_t tm_dev_malloc_raw(void **p, size_t bytes);
template <class T> inline cudaError_t tm_dev_malloc(T **p, size_t bytes) {
    return tm_dev_malloc_raw(reinterpret_cast<void **>(p), bytes);
}
void tm_dev_free(void *p);
int tm_prof_begin(int tag, cudaStream_t s); // returns a token (-1 when profiling is off)
// error-model audit (profiling only): device slots {max ratio as float bits, candidate count}, or nullptr
unsigned long long *tm_audit_slots();
void tm_prof_end(int token, cudaStream_t s);

// ---- descriptor kernels (psyv.cu)
int tm_launch_psyv_queries(const int32_t *d_rgb, int64_t nq, int flags, int gamma, float *d_out,
                           cudaStream_t s);
int tm_launch_psyv_tileset(const uint8_t *d_palpix, int64_t ntiles, const int32_t *d_pals, int npal,
                           int pal_size, const uint8_t *d_tile_flags, int flags, int gamma,
                           float *d_out, cudaStream_t s);
int tm_launch_global_rows(const uint8_t *d_palpix, int64_t ntiles, float *d_out, cudaStream_t s);
// tileset rows for an explicit row list: d_row_map[r] = ((pal * ntiles + tile) * 2 + vmir) * 2 + hmir
int tm_launch_psyv_tileset_rows(const uint8_t *d_palpix, int64_t ntiles, const int32_t *d_pals, int npal,
                                int pal_size, const uint8_t *d_tile_flags, int flags, int gamma,
                                const int64_t *d_row_map, int64_t nrows, float *d_out, cudaStream_t s);

// ---- keyframe preparation (keyframe.cu): PrepareFrameTiling main.pas:3791-3967 on the device
int tm_launch_kf_lines(const uint8_t *d_palpix, int64_t ntiles, float *d_lines, cudaStream_t s);
// d_tr_tile: stage-1 row / 4 -> tile index (NULL = identity, every tile Active)
int tm_launch_kf_mark(const int32_t *d_item_pal, const int32_t *d_item_tile, int64_t nitems, int64_t ntiles, int npal,
                      const int *d_knn_idx, const float *d_knn_err, const int32_t *d_tr_tile, const uint8_t *d_palmask,
                      uint32_t *d_seen, uint8_t *d_used, int *d_bad, cudaStream_t s);
// stream compaction of used[] in index order; *h_count = KNNSize (synchronises s)
int tm_kf_compact(const uint8_t *d_used, int64_t n, int64_t **d_row_map_out, int64_t *h_count, cudaStream_t s);
int tm_launch_kf_tables(const int64_t *d_row_map, int64_t nrows, int64_t ntiles, int32_t *d_tr_tile, int32_t *d_tr_pal,
                        uint8_t *d_tr_attrs, cudaStream_t s);

// ---- dataset preparation (search_tc.cu): mean, centred fp16 operand, norms, scan-order copy, TMA descriptors
int tm_prepare_dataset(ANNkdtree *h, cudaStream_t s);
int tm_dataset_tensor_maps(ANNkdtree *h);
// ---- exact SIMT search (search_simt.cu): sequential fp32 accumulation, lowest index on ties
int tm_launch_search_simt(const float *d_rows, int64_t n, int dd, const float *d_q, int64_t nq, int k,
                          int *d_idx, float *d_err, cudaStream_t s);
// exact search restricted to a list of query ids (fallback path)
int tm_launch_search_simt_list(const float *d_rows, int64_t n, int dd, const float *d_q,
                               const int *d_qlist, const int *d_qcount, int64_t max_list, int k,
                               int *d_idx, float *d_err, cudaStream_t s);
// ---- tensor-core search (search_tc.cu)
int tm_search_tc(ANNkdtree *h, const float *d_q, int64_t nq, int k, int *d_idx, float *d_err,
                 cudaStream_t s);
// ---- kmodes (kmodes.cu)
int tm_launch_kmodes_assign(const uint8_t *d_X, int64_t np, const uint8_t *d_c, int64_t k, int sem,
                            int32_t *d_clust, uint64_t *d_dis, cudaStream_t s);
int tm_kmodes_pack_centroids(TmKmScratch &c, const uint8_t *d_c, int64_t k, cudaStream_t s);
TmKmScratch &g_km_scratch(); // the calling thread's device's scratch of the one-shot KModes calls (kmodes.cu)
int tm_kmodes_update_packed(TmKmScratch &c, uint8_t *d_c, int64_t k, const uint8_t *d_rows, const int32_t *d_idx, int64_t m,
                            int *d_bad, cudaStream_t s);
// optional completion through mapped host memory (KModes handle): merged results + the call's sequence number in *m_flag
struct TmKmHostOut {
    unsigned long long *keys = nullptr; // [keys_cap] device, all ~0 between calls: (dissimilarity << 32 | ~index) per point
    int64_t keys_cap = 0;
    int32_t *prev = nullptr;            // [np] labels of these points from the previous pass (-1: none): read as a bound, rewritten
    int32_t *m_clust = nullptr;         // device-visible addresses of page-locked host buffers
    uint64_t *m_dis = nullptr;          // may be NULL
    unsigned int *d_done = nullptr;     // device counter, zero between calls
    int *m_flag = nullptr;
    int seq = 0;
};
// ho != NULL and the call is bin-sized (fewer blocks of points than two per SM, np <= ho->keys_cap): the centroid ranges
// meet in ho->keys, results go to ho->m_clust / m_dis, *ho_used = true; otherwise to d_clust / d_dis as before
int tm_kmodes_assign_packed(TmKmScratch &c, const uint8_t *d_X, int64_t np, const uint8_t *d_c, int64_t k, int sem,
                            int32_t *d_clust, uint64_t *d_dis, cudaStream_t s, const TmKmHostOut *ho = nullptr,
                            bool *ho_used = nullptr);
int tm_kmodes_inline_rows();
int tm_kmodes_update_inline(TmKmScratch &c, uint8_t *d_c, int64_t k, const uint8_t *h_rows, const int32_t *h_idx, int m,
                            cudaStream_t s);
void tm_kmodes_scratch_free(TmKmScratch &c);
int tm_launch_kmodes_farthest_first(const uint8_t *d_X, int64_t np, int64_t k, int64_t init_point, int sem,
                                    uint8_t *d_centroids, int64_t *d_chosen, unsigned long long *d_argkey,
                                    uint32_t *d_mind, uint8_t *d_used, int *d_err, cudaStream_t s);
int tm_launch_kmodes_representatives(const uint8_t *d_X, int64_t np, const int32_t *d_labels, const uint8_t *d_c, int64_t k,
                                     int sem, unsigned long long *d_best, int32_t *d_rep, int32_t *d_count, int *d_err,
                                     cudaStream_t s);
// query descriptors + (optionally) the tensor-core engine's query operand in one pass
int tm_launch_psyv_queries_operand(const int32_t *d_rgb, int64_t nq, int flags, int gamma, float *d_out,
                                   const float *d_mean, void *d_q16, float *d_qn16, float *d_qne, float *d_qde,
                                   const int *d_dst, cudaStream_t s);
// ---- coherence order of a query batch (qorder.cu): sort by a Morton code of the tile's mean colour
size_t tm_query_order_temp_bytes(int64_t nq);
int tm_launch_query_order(const int32_t *d_rgb, int64_t nq, uint32_t *d_keys /* [2][nq] */, int *d_vals /* [nq] */,
                          int *d_perm, int *d_inv, void *d_tmp, size_t tmp_bytes, cudaStream_t s);
int tm_launch_translate_list(int *d_list, const int *d_count, int64_t cap, const int *d_perm, cudaStream_t s);
int tm_launch_scatter_query_results(const int *d_perm, int64_t nq, const int *d_idx_s, const float *d_err_s, int *d_idx,
                                    float *d_err, cudaStream_t s);
// tensor-core engine with a query operand supplied by the caller: tm_tc_query_operand_buffers hands out the engine's
// scratch (padding rows cleared) or returns 1 when the engine does not apply to this (dataset, k, nq)
int tm_tc_query_operand_buffers(ANNkdtree *h, int64_t nq, void **d_q16, float **d_qn16, float **d_qne, float **d_qde,
                                cudaStream_t s);
void tm_tc_set_coherent(bool on); // hint: the next searches' queries are in coherence order (qorder.cu)
int tm_search_tc_prepared(ANNkdtree *h, const float *d_q, int64_t nq, int k, int *d_idx, float *d_err, cudaStream_t s);
// rows of dd floats (dd % 4 == 0) gathered by index list / (idx, err)[n][k] scattered back by the same list
int tm_launch_gather_rows(const float *d_rows, const int *d_list, int n, int dd, float *d_out, cudaStream_t s);
int tm_launch_scatter_results(const int *d_list, int n, int k, const int *d_ti, const float *d_te, int *d_idx, float *d_err,
                              cudaStream_t s);
int tm_launch_temporal_smoothing(const uint8_t *d_palpix, int64_t ntiles, const int32_t *d_pals, int npal, int pal_size,
                                 const int32_t *d_frame_kf, int64_t nframes, int64_t npos, const int32_t *d_tile,
                                 const int32_t *d_pal, const uint8_t *d_attrs, double strength, int32_t *d_src_frame,
                                 int32_t *d_out_tile, int32_t *d_out_pal, uint8_t *d_out_attrs, int *d_bad,
                                 cudaStream_t s);
int tm_launch_kmodes_rows(const uint8_t *d_palpix, int64_t ntiles, int pal_size, uint8_t *d_rows, int32_t *d_signi,
                          cudaStream_t s);
int tm_launch_kmodes_umd(const uint8_t *d_X, int64_t np, const uint8_t *d_center, int sem,
                         uint64_t *d_mind, cudaStream_t s);

// ---- reindex (reindex.cu): btnReindexClick main.pas:1199-1230 + ReindexTiles 4483-4527
int tm_reindex_dev_impl(int32_t *d_tilemap, int64_t n, int64_t ntiles, int32_t *d_count, int32_t *d_idx_map,
                        int32_t *d_order, int64_t *h_active, cudaStream_t s);
