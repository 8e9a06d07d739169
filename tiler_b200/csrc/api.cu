// api.cu — the C ABI of libtmann.so (see include/tmann.h for the contract and the
// reference file:line each entry point replaces).
#include "tm_common.cuh"
#include <stdarg.h>
#include <string.h>
#include <stdlib.h>
#include <chrono>
#include <algorithm>
#include <vector>

std::atomic<int64_t> g_launches{0}, g_queries{0}, g_fallbacks{0};
int g_engine = TM_ENGINE_AUTO;

namespace {
// The reference calls the library from a pool of worker threads over shared handles (DoFrameTiling, one frame per
// thread, main.pas:972).  All device work goes through one stream and grow-only scratch, so every entry point that
// touches the device runs under one process-wide lock (recursive: entry points call each other).
std::recursive_mutex g_api_mu;
#define TM_API_GUARD std::lock_guard<std::recursive_mutex> api_guard_(g_api_mu)
thread_local char t_err[512] = "";
std::mutex g_init_mu;
bool g_inited = false;
int g_device = -1, g_sms = 0;
cudaStream_t g_stream = nullptr;
std::mutex g_stats_mu;
int g_profiling = 0; // 0 off, 1 phase timing, 2 phase timing + error-model audit
struct ProfSpan { int tag; cudaEvent_t a, b; };
std::vector<ProfSpan> g_spans;
float g_ms[4] = {0, 0, 0, 0};
int64_t g_coarse_launches = 0;
unsigned long long *g_audit = nullptr; // device: [0] max error ratio (float bits), [1] audited candidates

struct DevBuf { // simple RAII device allocation
    void *p = nullptr;
    ~DevBuf() { if (p) tm_dev_free(p); }
    int alloc(size_t bytes) { return tm_check(tm_dev_malloc(&p, bytes ? bytes : 1), "cudaMalloc"); }
    template <class T> T *as() { return static_cast<T *>(p); }
};
} // namespace

void tm_set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(t_err, sizeof t_err, fmt, ap);
    va_end(ap);
}

int tm_check(cudaError_t e, const char *what) {
    if (e == cudaSuccess) return 0;
    tm_set_error("CUDA error %s: %s", what, cudaGetErrorString(e));
    return e == cudaErrorMemoryAllocation ? TM_ENOMEM : TM_ECUDA;
}

int tm_ensure_device() {
    std::lock_guard<std::mutex> lk(g_init_mu);
    if (g_inited) {
        // every host thread that enters the library must be on the library's device
        cudaSetDevice(g_device);
        return 0;
    }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        tm_set_error("no CUDA device visible: libtmann has no CPU fallback");
        return TM_ENODEVICE;
    }
    int dev = 0;
    if (g_device >= 0) dev = g_device; else cudaGetDevice(&dev);
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, dev) != cudaSuccess || prop.major != 10) {
        tm_set_error("device %d is not sm_100 (compute %d.%d): libtmann is built for sm_100a only", dev,
                     prop.major, prop.minor);
        return TM_ENODEVICE;
    }
    if (cudaSetDevice(dev) != cudaSuccess) { tm_set_error("cudaSetDevice(%d) failed", dev); return TM_ENODEVICE; }
    if (cudaStreamCreateWithFlags(&g_stream, cudaStreamNonBlocking) != cudaSuccess) {
        tm_set_error("cudaStreamCreate failed");
        return TM_ENODEVICE;
    }
    {   // keep freed blocks in the driver's pool instead of returning them to the OS at every synchronisation
        cudaMemPool_t pool;
        if (cudaDeviceGetDefaultMemPool(&pool, dev) == cudaSuccess) {
            uint64_t keep = UINT64_MAX;
            cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
        }
    }
    g_device = dev;
    g_sms = prop.multiProcessorCount;
    g_inited = true;
    return 0;
}

// profiling spans: tag 0 = descriptor, 1 = search phase, 2 = coarse tensor-core kernel, 3 = re-rank
int tm_prof_begin(int tag, cudaStream_t s) {
    if (!g_profiling) return -1;
    ProfSpan sp{tag, nullptr, nullptr};
    cudaEventCreate(&sp.a);
    cudaEventCreate(&sp.b);
    cudaEventRecord(sp.a, s);
    std::lock_guard<std::mutex> lk(g_stats_mu);
    g_spans.push_back(sp);
    if (tag == 2) g_coarse_launches++;
    return (int)g_spans.size() - 1;
}
void tm_prof_end(int token, cudaStream_t s) {
    if (token < 0) return;
    std::lock_guard<std::mutex> lk(g_stats_mu);
    cudaEventRecord(g_spans[token].b, s);
}
static void prof_collect() {
    std::lock_guard<std::mutex> lk(g_stats_mu);
    if (g_spans.empty()) return;
    cudaStreamSynchronize(g_stream);
    for (auto &sp : g_spans) {
        float ms = 0;
        if (cudaEventElapsedTime(&ms, sp.a, sp.b) == cudaSuccess) g_ms[sp.tag] += ms;
        cudaEventDestroy(sp.a);
        cudaEventDestroy(sp.b);
    }
    g_spans.clear();
}

unsigned long long *tm_audit_slots() {
    if (g_profiling < 2) return nullptr;
    if (!g_audit) {
        if (cudaMalloc(&g_audit, 16) != cudaSuccess) { g_audit = nullptr; return nullptr; }
        cudaMemsetAsync(g_audit, 0, 16, g_stream);
    }
    return g_audit;
}

cudaError_t tm_dev_malloc_raw(void **p, size_t bytes) { return cudaMallocAsync(p, bytes ? bytes : 1, g_stream); }
void tm_dev_free(void *p) {
    if (p) cudaFreeAsync(p, g_stream);
}

cudaStream_t tm_stream() { return g_stream; }
int tm_num_sms() { return g_sms; }

static int valid(const ANNkdtree *h) {
    if (!h || h->magic != TM_MAGIC) { tm_set_error("invalid dataset handle"); return TM_EINVAL; }
    return 0;
}

// The caller's stream (legacy / default stream work of torch etc.) and ours are ordered
// explicitly: device-pointer entry points synchronise the device before reading inputs that
// another stream may still be producing, and block until their own work is complete.
// After tm_set_stream(s) the library runs on the caller's stream and the device-pointer
// entry points are asynchronous on it (no host synchronisation at all).
static bool g_user_stream = false;
static int sync_in() { return g_user_stream ? 0 : tm_check(cudaDeviceSynchronize(), "sync(in)"); }
static int sync_out() { return tm_check(cudaStreamSynchronize(g_stream), "sync(out)"); }
static int sync_out_dev() { return g_user_stream ? 0 : sync_out(); }

extern "C" {

const char *tm_last_error(void) { return t_err; }
int tm_version(void) { return 100; }

int tm_set_device(int device) {
    TM_API_GUARD;
    {
        std::lock_guard<std::mutex> lk(g_init_mu);
        if (g_inited && device != g_device) {
            tm_set_error("device already initialised to %d", g_device);
            return TM_EINVAL;
        }
        g_device = device;
    }
    return tm_ensure_device();
}

int tm_set_engine(int engine) {
    if (engine < TM_ENGINE_AUTO || engine > TM_ENGINE_TC) { tm_set_error("bad engine"); return TM_EINVAL; }
    g_engine = engine;
    return 0;
}

int tm_set_stream(void *cuda_stream) {
    TM_API_GUARD;
    int r = tm_ensure_device();
    if (r) return r;
    cudaStreamSynchronize(g_stream);
    g_stream = static_cast<cudaStream_t>(cuda_stream);
    g_user_stream = true;
    return 0;
}

int tm_device_sync(void) {
    TM_API_GUARD;
    int r = tm_ensure_device();
    if (r) return r;
    return tm_check(cudaDeviceSynchronize(), "cudaDeviceSynchronize");
}

static ANNkdtree *new_handle(int64_t n, int dd) {
    ANNkdtree *h = new ANNkdtree();
    memset(h, 0, sizeof *h);
    h->magic = TM_MAGIC;
    h->n = n;
    h->dd = dd;
    h->device = g_device;
    h->single_mu = new std::mutex();
    return h;
}

static void free_handle(ANNkdtree *h) {
    if (!h) return;
    tm_dev_free(h->rows_f32); tm_dev_free(h->rows_f16); tm_dev_free(h->norms); tm_dev_free(h->rows_f16s); tm_dev_free(h->norms_s); tm_dev_free(h->perm); tm_dev_free(h->norm_tiles); tm_dev_free(h->mean);
    tm_dev_free(h->d_q1); tm_dev_free(h->d_i1); tm_dev_free(h->d_e1);
    tm_dev_free(h->knn8_idx); tm_dev_free(h->knn8_err);
    free(h->h_tr_tile); free(h->h_tr_pal); free(h->h_tr_attrs);
    if (h->h_pin) cudaFreeHost(h->h_pin);
    free(h->tmap_rows);
    delete h->single_mu;
    h->magic = 0;
    delete h;
}

// finish a handle whose rows_f32 is filled (on g_stream): build the tensor-core operand
static ANNkdtree *finish_handle(ANNkdtree *h) {
    int r = 0;
    if (!r) r = tm_prepare_dataset(h, g_stream);
    if (!r) r = sync_out();
    if (r) { free_handle(h); return nullptr; }
    return h;
}

ANNkdtree *tm_dataset_create(const float *rows, int64_t n, int dd, int is_device) {
    TM_API_GUARD;
    if (tm_ensure_device()) return nullptr;
    if (n < 0 || dd < 1 || (n > 0 && !rows)) { tm_set_error("tm_dataset_create: bad arguments"); return nullptr; }
    if (n > 0x7fffffff) { tm_set_error("tm_dataset_create: n too large"); return nullptr; }
    ANNkdtree *h = new_handle(n, dd);
    size_t bytes = sizeof(float) * (size_t)n * dd;
    if (tm_check(tm_dev_malloc(&h->rows_f32, bytes ? bytes : 4), "tm_dev_malloc(rows)")) { free_handle(h); return nullptr; }
    if (is_device && sync_in()) { free_handle(h); return nullptr; }
    if (bytes && tm_check(cudaMemcpyAsync(h->rows_f32, rows, bytes, is_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, g_stream), "copy rows")) {
        free_handle(h);
        return nullptr;
    }
    return finish_handle(h);
}

ANNkdtree *ann_kdtree_create(float **pa, int n, int dd, int bs, int split) {
    TM_API_GUARD;
    (void)bs; (void)split;
    if (n < 0 || dd < 1 || (n > 0 && !pa)) { tm_set_error("ann_kdtree_create: bad arguments"); return nullptr; }
    // rows are separately allocated Pascal dynamic arrays (main.pas:3757, 3955): gather them
    std::vector<float> flat((size_t)n * dd);
    for (int i = 0; i < n; i++) memcpy(flat.data() + (size_t)i * dd, pa[i], sizeof(float) * dd);
    return tm_dataset_create(flat.data(), n, dd, 0);
}

void ann_kdtree_destroy(ANNkdtree *akd) {
    TM_API_GUARD;
    if (!akd || akd->magic != TM_MAGIC) return; // NULL-safe like the reference wrapper
    cudaSetDevice(akd->device);
    cudaStreamSynchronize(g_stream);
    free_handle(akd);
}

int64_t tm_dataset_rows(const ANNkdtree *akd) { return valid(akd) ? -1 : akd->n; }
int tm_dataset_dim(const ANNkdtree *akd) { return valid(akd) ? -1 : akd->dd; }

int tm_dataset_read(const ANNkdtree *akd, float *h_rows) {
    TM_API_GUARD;
    int r = valid(akd);
    if (r) return r;
    if ((r = tm_ensure_device())) return r;
    TM_CUDA(cudaMemcpyAsync(h_rows, akd->rows_f32, sizeof(float) * (size_t)akd->n * akd->dd, cudaMemcpyDeviceToHost, g_stream));
    return sync_out();
}

// engine dispatch; everything on g_stream, no host sync inside
// engine choice for one batch (TM_ENGINE_*), or -1 with the error set
static int pick_engine(ANNkdtree *h, int64_t nq) {
    const bool tc_ok = (h->rows_f16 != nullptr) && h->n >= 1;
    int engine = g_engine;
    if (engine == TM_ENGINE_TC && !tc_ok) { tm_set_error("tensor-core engine unavailable for this dataset (dd=%d)", h->dd); return -1; }
    if (engine == TM_ENGINE_AUTO) engine = (tc_ok && nq >= 64) ? TM_ENGINE_TC : TM_ENGINE_SIMT;
    return engine;
}

static int search_dev(ANNkdtree *h, const float *d_q, int64_t nq, int k, int *d_idx, float *d_err, bool operand_ready = false) {
    if (k < 1 || k > 16) { tm_set_error("k must be in 1..16 (got %d)", k); return TM_EINVAL; }
    g_queries.fetch_add(nq, std::memory_order_relaxed);
    const int engine = pick_engine(h, nq);
    if (engine < 0) return TM_EINVAL;
    int tok = tm_prof_begin(1, g_stream);
    int r = (engine == TM_ENGINE_TC) ? (operand_ready ? tm_search_tc_prepared(h, d_q, nq, k, d_idx, d_err, g_stream)
                                                      : tm_search_tc(h, d_q, nq, k, d_idx, d_err, g_stream))
                                     : tm_launch_search_simt(h->rows_f32, h->n, h->dd, d_q, nq, k, d_idx, d_err, g_stream);
    tm_prof_end(tok, g_stream);
    return r;
}

int tm_search_batch_dev(ANNkdtree *akd, const float *d_q, int64_t nq, int k, int *d_idxs, float *d_errs) {
    TM_API_GUARD;
    int r = valid(akd);
    if (r) return r;
    if ((r = tm_ensure_device())) return r;
    if (nq < 0) { tm_set_error("nq < 0"); return TM_EINVAL; }
    if (nq == 0) return 0;
    if ((r = sync_in())) return r;
    if ((r = search_dev(akd, d_q, nq, k, d_idxs, d_errs))) return r;
    return sync_out_dev();
}

int ann_kdtree_search_batch(ANNkdtree *akd, const float *h_q, int64_t nq, int k, int *h_idxs, float *h_errs) {
    TM_API_GUARD;
    int r = valid(akd);
    if (r) return r;
    if ((r = tm_ensure_device())) return r;
    if (nq < 0 || k < 1) { tm_set_error("bad nq/k"); return TM_EINVAL; }
    if (nq == 0) return 0;
    DevBuf q, i, e;
    if ((r = q.alloc(sizeof(float) * (size_t)nq * akd->dd))) return r;
    if ((r = i.alloc(sizeof(int) * (size_t)nq * k))) return r;
    if ((r = e.alloc(sizeof(float) * (size_t)nq * k))) return r;
    TM_CUDA(cudaMemcpyAsync(q.p, h_q, sizeof(float) * (size_t)nq * akd->dd, cudaMemcpyHostToDevice, g_stream));
    if ((r = search_dev(akd, q.as<float>(), nq, k, i.as<int>(), e.as<float>()))) return r;
    TM_CUDA(cudaMemcpyAsync(h_idxs, i.p, sizeof(int) * (size_t)nq * k, cudaMemcpyDeviceToHost, g_stream));
    if (h_errs) TM_CUDA(cudaMemcpyAsync(h_errs, e.p, sizeof(float) * (size_t)nq * k, cudaMemcpyDeviceToHost, g_stream));
    return sync_out();
}

// legacy single-query entry points: re-entrant (per-handle mutex around the pinned scratch)
static int search_one(ANNkdtree *akd, const float *q, int k, int *idxs, float *errs) {
    if (valid(akd) || tm_ensure_device()) return -1;
    if (k < 1 || k > 16 || !q) return -1;
    std::lock_guard<std::mutex> lk(*akd->single_mu);
    if (!akd->h_pin) { // scratch of the one-query entry points, on first use (pinned allocation is slow)
        if (tm_dev_malloc(&akd->d_q1, sizeof(float) * akd->dd) || tm_dev_malloc(&akd->d_i1, sizeof(int) * 16) ||
            tm_dev_malloc(&akd->d_e1, sizeof(float) * 16) ||
            cudaMallocHost(&akd->h_pin, sizeof(float) * (akd->dd + 32)) != cudaSuccess) {
            akd->h_pin = nullptr;
            return -1;
        }
    }
    memcpy(akd->h_pin, q, sizeof(float) * akd->dd);
    if (cudaMemcpyAsync(akd->d_q1, akd->h_pin, sizeof(float) * akd->dd, cudaMemcpyHostToDevice, g_stream)) return -1;
    g_queries.fetch_add(1, std::memory_order_relaxed);
    if (tm_launch_search_simt(akd->rows_f32, akd->n, akd->dd, akd->d_q1, 1, k, akd->d_i1, akd->d_e1, g_stream)) return -1;
    float *pe = akd->h_pin + akd->dd;
    int *pi = reinterpret_cast<int *>(pe + 16);
    if (cudaMemcpyAsync(pi, akd->d_i1, sizeof(int) * k, cudaMemcpyDeviceToHost, g_stream)) return -1;
    if (cudaMemcpyAsync(pe, akd->d_e1, sizeof(float) * k, cudaMemcpyDeviceToHost, g_stream)) return -1;
    if (cudaStreamSynchronize(g_stream)) return -1;
    for (int j = 0; j < k; j++) { idxs[j] = pi[j]; if (errs) errs[j] = pe[j]; }
    return 0;
}

int ann_kdtree_search(ANNkdtree *akd, float *q, float eps, float *err) {
    TM_API_GUARD;
    (void)eps;
    int idx = -1;
    float e = 0.f; // the reference wrapper pre-sets idx = -1, dist = 0
    if (search_one(akd, q, 1, &idx, &e)) { if (err) *err = 0.f; return -1; }
    if (err) *err = e;
    return idx;
}

int ann_kdtree_pri_search(ANNkdtree *akd, float *q, float eps, float *err) { return ann_kdtree_search(akd, q, eps, err); }

int ann_kdtree_search_multi(ANNkdtree *akd, int *idxs, float *errs, int cnt, float *q, float eps) {
    TM_API_GUARD;
    (void)eps;
    if (cnt <= 0 || !idxs) return 0;
    if (cnt <= 16) { search_one(akd, q, cnt, idxs, errs); return 0; }
    tm_set_error("ann_kdtree_search_multi: cnt > 16 unsupported");
    for (int j = 0; j < cnt; j++) { idxs[j] = -1; if (errs) errs[j] = 0.f; }
    return 0;
}

// pinned host memory for the caller's frame buffers (full-rate, asynchronous PCIe copies)
void *tm_host_alloc(size_t bytes) {
    if (tm_ensure_device()) return nullptr;
    void *p = nullptr;
    if (tm_check(cudaMallocHost(&p, bytes ? bytes : 1), "cudaMallocHost")) return nullptr;
    return p;
}
void tm_host_free(void *p) {
    if (p) cudaFreeHost(p);
}

int tm_get_stats(tm_stats *out) {
    if (!out) return TM_EINVAL;
    out->kernel_launches = g_launches.load();
    out->queries = g_queries.load();
    out->fallback_queries = g_fallbacks.load();
    prof_collect();
    std::lock_guard<std::mutex> lk(g_stats_mu);
    out->desc_ms = g_ms[0];
    out->search_ms = g_ms[1];
    out->coarse_ms = g_ms[2];
    out->rerank_ms = g_ms[3];
    out->coarse_launches = g_coarse_launches;
    out->max_error_ratio = 0.f;
    out->audited_candidates = 0;
    if (g_audit) {
        unsigned long long a[2] = {0, 0};
        cudaStreamSynchronize(g_stream);
        if (cudaMemcpy(a, g_audit, 16, cudaMemcpyDeviceToHost) == cudaSuccess) {
            const unsigned int bits = (unsigned int)a[0];
            memcpy(&out->max_error_ratio, &bits, 4);
            out->audited_candidates = (int64_t)a[1];
        }
    }
    return 0;
}

int tm_reset_stats(void) {
    prof_collect();
    g_launches = 0; g_queries = 0; g_fallbacks = 0;
    std::lock_guard<std::mutex> lk(g_stats_mu);
    for (float &m : g_ms) m = 0;
    g_coarse_launches = 0;
    if (g_audit) cudaMemsetAsync(g_audit, 0, 16, g_stream);
    return 0;
}

int tm_set_profiling(int on) {
    g_profiling = on < 0 ? 0 : on;
    return 0;
}

} // extern "C"

#include "api_fused.inc"
#include "api_keyframe.inc"
