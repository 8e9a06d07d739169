// api.cu — the C ABI of libtmann.so (see include/tmann.h for the contract and the
// reference file:line each entry point replaces).
#include "tm_common.cuh"
#include <stdarg.h>
#include <string.h>
#include <stdlib.h>
#include <chrono>
#include <algorithm>
#include <vector>
#include <unordered_map>
#include <atomic>
#include <string>
#include <thread>
#include <dlfcn.h>

std::atomic<int64_t> g_launches{0}, g_queries{0}, g_fallbacks{0}, g_coarse_launches{0};
int g_engine = TM_ENGINE_AUTO;

namespace {
// The reference calls the library from a pool of worker threads over shared handles (DoFrameTiling, one frame per
// thread, main.pas:972).  Every GPU the library drives has its own context (TmDevice: stream, scratch, lock), so
// callers serialise per device, not per process; a dataset handle has one replica per active device.
thread_local char t_err[512] = "";
thread_local TmDevice *t_dev = nullptr;
std::mutex g_init_mu;
TmDevice *g_ctx[TM_MAX_DEV] = {};  // by CUDA ordinal, created on first use, never destroyed
std::vector<int> g_active;         // the devices new datasets are replicated on; [0] = primary
std::vector<int> g_requested;      // tm_set_devices / tm_set_device before first use
std::atomic<int> g_live_handles{0};
int g_profiling = 0; // 0 off, 1 phase timing, 2 phase timing + error-model audit
double g_nccl_init_ms = 0.0;

struct DevBuf { // simple RAII device allocation
    void *p = nullptr;
    ~DevBuf() { if (p) tm_dev_free(p); }
    int alloc(size_t bytes) { return tm_check(tm_dev_malloc(&p, bytes ? bytes : 1), "cudaMalloc"); }
    template <class T> T *as() { return static_cast<T *>(p); }
};

// Binds the calling thread to a device for the duration of an entry point: takes the device's lock, makes it the
// thread's CUDA device and the target of tm_dev().  Nests (entry points call each other, possibly across devices).
struct TmApiGuard {
    TmDevice *d, *prev;
    explicit TmApiGuard(TmDevice *dev) : d(dev), prev(t_dev) {
        if (!d) return;
        d->mu.lock();
        t_dev = d;
        cudaSetDevice(d->device);
    }
    ~TmApiGuard() {
        if (!d) return;
        t_dev = prev;
        if (prev && prev->device != d->device) cudaSetDevice(prev->device);
        d->mu.unlock();
    }
    TmApiGuard(const TmApiGuard &) = delete;
    TmApiGuard &operator=(const TmApiGuard &) = delete;
};
#define TM_API_GUARD TmApiGuard api_guard_(tm_primary())
#define TM_API_GUARD_H(h) TmApiGuard api_guard_(tm_device_of(h))
} // namespace

TmDevice *tm_dev() { return t_dev; }

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

namespace {
// ---- NCCL, loaded at run time (libnccl.so.2): only a process that drives >= 2 GPUs needs it ------------------
typedef struct ncclComm *ncclComm_t;
struct Nccl {
    void *lib = nullptr;
    int (*CommInitAll)(ncclComm_t *, int, const int *) = nullptr;
    int (*CommDestroy)(ncclComm_t) = nullptr;
    int (*Broadcast)(const void *, void *, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    const char *(*GetErrorString)(int) = nullptr;
} g_nccl;
constexpr int NCCL_CHAR = 0; // ncclInt8 / ncclChar

int nccl_load() {
    if (g_nccl.lib) return 0;
    const char *names[] = {getenv("TM_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
    void *lib = nullptr;
    for (const char *n : names)
        if (n && (lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL))) break;
    if (!lib) { tm_set_error("multi-device needs NCCL: cannot load libnccl.so.2 (%s)", dlerror()); return TM_ENODEVICE; }
#define TM_NCCL_SYM(field, name)                                                                         \
    if (!(*reinterpret_cast<void **>(&g_nccl.field) = dlsym(lib, name))) {                               \
        tm_set_error("libnccl lacks %s", name);                                                          \
        return TM_ENODEVICE;                                                                             \
    }
    TM_NCCL_SYM(CommInitAll, "ncclCommInitAll")
    TM_NCCL_SYM(CommDestroy, "ncclCommDestroy")
    TM_NCCL_SYM(Broadcast, "ncclBroadcast")
    TM_NCCL_SYM(GroupStart, "ncclGroupStart")
    TM_NCCL_SYM(GroupEnd, "ncclGroupEnd")
    TM_NCCL_SYM(GetErrorString, "ncclGetErrorString")
#undef TM_NCCL_SYM
    g_nccl.lib = lib;
    return 0;
}
int nccl_check(int rc, const char *what) {
    if (rc == 0) return 0;
    tm_set_error("NCCL error in %s: %s", what, g_nccl.GetErrorString ? g_nccl.GetErrorString(rc) : "?");
    return TM_ECUDA;
}

// context of one device (g_init_mu held)
TmDevice *make_ctx(int dev) {
    if (dev < 0 || dev >= TM_MAX_DEV) { tm_set_error("device %d outside [0, %d)", dev, TM_MAX_DEV); return nullptr; }
    if (g_ctx[dev]) return g_ctx[dev];
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, dev) != cudaSuccess || prop.major != 10) {
        cudaGetLastError();
        tm_set_error("device %d is not sm_100 (compute %d.%d): libtmann is built for sm_100a only", dev, prop.major, prop.minor);
        return nullptr;
    }
    int cur = -1;
    cudaGetDevice(&cur);
    if (cudaSetDevice(dev) != cudaSuccess) { tm_set_error("cudaSetDevice(%d) failed", dev); return nullptr; }
    TmDevice *d = new TmDevice();
    d->device = dev;
    d->sms = prop.multiProcessorCount;
    if (cudaStreamCreateWithFlags(&d->own_stream, cudaStreamNonBlocking) != cudaSuccess) {
        tm_set_error("cudaStreamCreate failed on device %d", dev);
        delete d;
        if (cur >= 0) cudaSetDevice(cur);
        return nullptr;
    }
    d->stream = d->own_stream;
    {   // keep freed blocks in the driver's pool instead of returning them to the OS at every synchronisation
        cudaMemPool_t pool;
        if (cudaDeviceGetDefaultMemPool(&pool, dev) == cudaSuccess) {
            uint64_t keep = UINT64_MAX;
            cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
        }
    }
    if (cur >= 0) cudaSetDevice(cur);
    g_ctx[dev] = d;
    return d;
}

void drop_comms() {
    for (TmDevice *d : g_ctx)
        if (d && d->nccl_comm) { g_nccl.CommDestroy(static_cast<ncclComm_t>(d->nccl_comm)); d->nccl_comm = nullptr; }
}

// make `devs` the active list (g_init_mu held): contexts, and for >= 2 devices one NCCL communicator over them,
// warmed with a small broadcast so that the first dataset does not pay the bring-up
int activate(const std::vector<int> &devs) {
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        tm_set_error("no CUDA device visible: libtmann has no CPU fallback");
        return TM_ENODEVICE;
    }
    std::vector<int> list = devs;
    if (list.empty()) {
        const char *env = getenv("TM_DEVICES"); // "all" or "0,1,2"
        if (env && *env) {
            if (!strcmp(env, "all")) for (int i = 0; i < ndev && i < TM_MAX_DEV; i++) list.push_back(i);
            else for (const char *p = env; *p;) { list.push_back(atoi(p)); while (*p && *p != ',') p++; if (*p) p++; }
        } else {
            int cur = 0;
            cudaGetDevice(&cur);
            list.push_back(cur);
        }
    }
    for (size_t i = 0; i < list.size(); i++) {
        if (list[i] < 0 || list[i] >= ndev) { tm_set_error("device %d not present (%d visible)", list[i], ndev); return TM_ENODEVICE; }
        for (size_t j = 0; j < i; j++) if (list[j] == list[i]) { tm_set_error("device %d listed twice", list[i]); return TM_EINVAL; }
        if (!make_ctx(list[i])) return TM_ENODEVICE;
    }
    if (list.size() > 1) {
        int r = nccl_load();
        if (r) return r;
        drop_comms();
        const auto t0 = std::chrono::steady_clock::now();
        std::vector<ncclComm_t> comms(list.size());
        if ((r = nccl_check(g_nccl.CommInitAll(comms.data(), (int)list.size(), list.data()), "ncclCommInitAll"))) return r;
        for (size_t i = 0; i < list.size(); i++) g_ctx[list[i]]->nccl_comm = comms[i];
        // warm-up: 256 bytes from the first device to all
        std::vector<void *> buf(list.size(), nullptr);
        for (size_t i = 0; i < list.size(); i++) { cudaSetDevice(list[i]); cudaMalloc(&buf[i], 256); cudaMemset(buf[i], 0, 256); }
        g_nccl.GroupStart();
        for (size_t i = 0; i < list.size(); i++)
            g_nccl.Broadcast(buf[0], buf[i], 256, NCCL_CHAR, 0, comms[i], g_ctx[list[i]]->own_stream);
        r = nccl_check(g_nccl.GroupEnd(), "ncclBroadcast(warm-up)");
        for (size_t i = 0; i < list.size(); i++) { cudaSetDevice(list[i]); cudaStreamSynchronize(g_ctx[list[i]]->own_stream); cudaFree(buf[i]); }
        cudaSetDevice(list[0]);
        g_nccl_init_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
        if (r) return r;
    } else if (g_nccl.lib) {
        drop_comms();
    }
    g_active = list;
    return 0;
}

// first active device, initialising the library on first use; nullptr (error set) when there is no usable GPU
TmDevice *tm_primary() {
    std::lock_guard<std::mutex> lk(g_init_mu);
    if (g_active.empty() && activate(g_requested)) return nullptr;
    return g_ctx[g_active[0]];
}
TmDevice *tm_device_of(const ANNkdtree *h) {
    if (h && h->magic == TM_MAGIC && h->device >= 0 && h->device < TM_MAX_DEV && g_ctx[h->device]) return g_ctx[h->device];
    return tm_primary();
}
std::vector<int> active_devices() {
    std::lock_guard<std::mutex> lk(g_init_mu);
    return g_active;
}
} // namespace

// after the guard: 0 when the calling thread is bound to a usable device
int tm_ensure_device() { return t_dev ? 0 : TM_ENODEVICE; }

// profiling spans: tag 0 = descriptor, 1 = search phase, 2 = coarse tensor-core kernel, 3 = re-rank
int tm_prof_begin(int tag, cudaStream_t s) {
    if (!g_profiling) return -1;
    TmDevice *d = t_dev;
    TmProfSpan sp{tag, nullptr, nullptr};
    cudaEventCreate(&sp.a);
    cudaEventCreate(&sp.b);
    cudaEventRecord(sp.a, s);
    std::lock_guard<std::mutex> lk(d->stats_mu);
    d->spans.push_back(sp);
    return (int)d->spans.size() - 1;
}
void tm_prof_end(int token, cudaStream_t s) {
    if (token < 0) return;
    TmDevice *d = t_dev;
    std::lock_guard<std::mutex> lk(d->stats_mu);
    cudaEventRecord(d->spans[token].b, s);
}
static void prof_collect(TmDevice *d) { // the calling thread is bound to d
    std::lock_guard<std::mutex> lk(d->stats_mu);
    if (d->spans.empty()) return;
    cudaStreamSynchronize(d->stream);
    for (auto &sp : d->spans) {
        float ms = 0;
        if (cudaEventElapsedTime(&ms, sp.a, sp.b) == cudaSuccess) d->ms[sp.tag] += ms;
        cudaEventDestroy(sp.a);
        cudaEventDestroy(sp.b);
    }
    d->spans.clear();
}

unsigned long long *tm_audit_slots() {
    if (g_profiling < 2) return nullptr;
    TmDevice *d = t_dev;
    if (!d->audit) {
        if (cudaMalloc(&d->audit, 16) != cudaSuccess) { d->audit = nullptr; return nullptr; }
        cudaMemsetAsync(d->audit, 0, 16, d->stream);
    }
    return d->audit;
}

cudaError_t tm_dev_malloc_raw(void **p, size_t bytes) { return cudaMallocAsync(p, bytes ? bytes : 1, t_dev->stream); }
void tm_dev_free(void *p) {
    if (p) cudaFreeAsync(p, t_dev->stream);
}

static int valid(const ANNkdtree *h) {
    if (!h || h->magic != TM_MAGIC) { tm_set_error("invalid dataset handle"); return TM_EINVAL; }
    return 0;
}

// The caller's stream (legacy / default stream work of torch etc.) and ours are ordered
// explicitly: device-pointer entry points synchronise the device before reading inputs that
// another stream may still be producing, and block until their own work is complete.
// After tm_set_stream(s) the library runs on the caller's stream and the device-pointer
// entry points are asynchronous on it (no host synchronisation at all).
static int sync_in() { return t_dev->user_stream ? 0 : tm_check(cudaDeviceSynchronize(), "sync(in)"); }
static int sync_out() { return tm_check(cudaStreamSynchronize(t_dev->stream), "sync(out)"); }
static int sync_out_dev() { return t_dev->user_stream ? 0 : sync_out(); }
#define g_stream (t_dev->stream)

static ANNkdtree *new_handle(int64_t n, int dd) {
    ANNkdtree *h = new ANNkdtree();
    memset(static_cast<void *>(h), 0, sizeof *h);
    h->magic = TM_MAGIC;
    h->n = n;
    h->dd = dd;
    h->device = t_dev->device;
    h->single_mu = new std::mutex();
    h->rr = new std::atomic<unsigned>(0);
    h->rep[0] = h;
    h->nrep = 1;
    g_live_handles.fetch_add(1);
    return h;
}

// one replica (or the primary itself); the calling thread is bound to its device
static void free_one(ANNkdtree *h) {
    tm_dev_free(h->rows_f32); tm_dev_free(h->rows_f16); tm_dev_free(h->norms); tm_dev_free(h->rows_f16s); tm_dev_free(h->norms_s); tm_dev_free(h->perm); tm_dev_free(h->norm_tiles); tm_dev_free(h->mean);
    tm_dev_free(h->d_q1); tm_dev_free(h->d_i1); tm_dev_free(h->d_e1);
    tm_dev_free(h->knn8_idx); tm_dev_free(h->knn8_err); tm_dev_free(h->d_tr_tile);
    free(h->h_tr_tile); free(h->h_tr_pal); free(h->h_tr_attrs);
    if (h->h_pin) cudaFreeHost(h->h_pin);
    free(h->tmap_rows);
    delete h->single_mu;
    delete h->rr;
    h->magic = 0;
    delete h;
    g_live_handles.fetch_sub(1);
}

static void free_handle(ANNkdtree *h) {
    if (!h) return;
    for (int i = 1; i < h->nrep; i++)
        if (h->rep[i]) {
            TmApiGuard g(g_ctx[h->rep[i]->device]);
            cudaStreamSynchronize(t_dev->stream);
            free_one(h->rep[i]);
        }
    TmApiGuard g(g_ctx[h->device]);
    free_one(h);
}

// Replicas of a finished dataset on the other active devices: same-size buffers there, then every buffer goes out
// in one NCCL broadcast from the primary (NVLink / NVSwitch), so the pack / sort / operand kernels run once.
static int replicate_handle(ANNkdtree *h) {
    const std::vector<int> devs = active_devices();
    if (devs.size() < 2 || devs[0] != h->device) return 0;
    const int nd = (int)devs.size();
    struct Buf { size_t off; size_t bytes; };
    const size_t n_pad = (size_t)h->n_pad, kp = (size_t)h->kp, n = (size_t)h->n, dd = (size_t)h->dd;
    std::vector<Buf> bufs;
#define TM_REP_BUF(member, nbytes) if (h->member && (size_t)(nbytes) != 0) bufs.push_back({offsetof(ANNkdtree, member), (size_t)(nbytes)})
    TM_REP_BUF(rows_f32, n * dd * 4);
    TM_REP_BUF(rows_f16, n_pad * kp * 2);
    TM_REP_BUF(norms, n_pad * 4);
    TM_REP_BUF(rows_f16s, n_pad * kp * 2);
    TM_REP_BUF(norms_s, n_pad * 4);
    TM_REP_BUF(perm, n_pad * 4);
    TM_REP_BUF(norm_tiles, (n_pad / 128) * 4096);
    TM_REP_BUF(mean, dd * 4);
    TM_REP_BUF(d_tr_tile, (n / 4) * 4);
#undef TM_REP_BUF
    auto slot = [](ANNkdtree *t, size_t off) { return reinterpret_cast<void **>(reinterpret_cast<char *>(t) + off); };
    int r = 0;
    for (int i = 1; i < nd && !r; i++) {
        TmApiGuard g(g_ctx[devs[i]]);
        ANNkdtree *c = new_handle(h->n, h->dd);
        c->owner = h;
        c->n_pad = h->n_pad; c->kp = h->kp; c->norm_scale = h->norm_scale; c->max_norm = h->max_norm; c->max_dnorm = h->max_dnorm;
        c->ntiles_all = h->ntiles_all;
        h->rep[i] = c;
        h->nrep = i + 1;
        for (const Buf &b : bufs)
            if (!r) r = tm_check(tm_dev_malloc_raw(slot(c, b.off), b.bytes), "tm_dev_malloc(replica)");
    }
    for (const Buf &b : bufs) {
        if (r) break;
        g_nccl.GroupStart();
        for (int i = 0; i < nd; i++) {
            TmDevice *d = g_ctx[devs[i]];
            g_nccl.Broadcast(*slot(h, b.off), *slot(h->rep[i], b.off), b.bytes, NCCL_CHAR, 0, static_cast<ncclComm_t>(d->nccl_comm), d->stream);
        }
        r = nccl_check(g_nccl.GroupEnd(), "ncclBroadcast(dataset)");
    }
    for (int i = 1; i < h->nrep; i++) {
        TmApiGuard g(g_ctx[devs[i]]);
        if (!r && h->rows_f16) r = tm_dataset_tensor_maps(h->rep[i]);
        int r2 = tm_check(cudaStreamSynchronize(t_dev->stream), "sync(replica)");
        if (!r) r = r2;
    }
    {
        TmApiGuard g(g_ctx[h->device]);
        int r2 = tm_check(cudaStreamSynchronize(t_dev->stream), "sync(primary)");
        if (!r) r = r2;
    }
    return r;
}

// finish a handle whose rows_f32 is filled (on the primary's stream): build the tensor-core operand, replicate
static ANNkdtree *finish_handle(ANNkdtree *h) {
    int r = 0;
    if (!r) r = tm_prepare_dataset(h, g_stream);
    if (!r) r = sync_out();
    if (!r) r = replicate_handle(h);
    if (r) { free_handle(h); return nullptr; }
    return h;
}

// replica of `h` for a call that runs on one device: the device `d_ptr` lives on when given, else the replica with
// the fewest host calls in flight (round-robin among equals)
static ANNkdtree *pick_replica(ANNkdtree *h, const void *d_ptr) {
    if (h->owner) h = h->owner;
    if (h->nrep < 2) return h;
    if (d_ptr) {
        cudaPointerAttributes pa;
        if (cudaPointerGetAttributes(&pa, d_ptr) == cudaSuccess && pa.type == cudaMemoryTypeDevice) {
            for (int i = 0; i < h->nrep; i++) if (h->rep[i]->device == pa.device) return h->rep[i];
        }
        cudaGetLastError();
        return h;
    }
    const unsigned start = h->rr->fetch_add(1, std::memory_order_relaxed);
    int best = -1, best_load = 1 << 30;
    for (int j = 0; j < h->nrep; j++) {
        const int i = (int)((start + j) % (unsigned)h->nrep);
        const int load = g_ctx[h->rep[i]->device]->inflight.load(std::memory_order_relaxed);
        if (load < best_load) { best_load = load; best = i; }
    }
    return h->rep[best];
}
struct InflightMark {
    TmDevice *d;
    explicit InflightMark(ANNkdtree *h) : d(g_ctx[h->device]) { d->inflight.fetch_add(1, std::memory_order_relaxed); }
    ~InflightMark() { d->inflight.fetch_sub(1, std::memory_order_relaxed); }
};

// Host batch over all replicas: contiguous shards (multiples of 256 queries), one worker thread per device, every
// device copies its own results straight into the caller's buffers - no collective on the result path.
template <class F> static int run_sharded(ANNkdtree *h, int64_t nq, F fn /* (replica, q0, q1) -> int */) {
    const int nd = h->nrep;
    int64_t per = (nq + nd - 1) / nd;
    per = (per + 255) & ~(int64_t)255;
    std::vector<int> rc(nd, 0);
    std::vector<std::string> msg(nd);
    std::vector<std::thread> th;
    auto work = [&](int i) {
        const int64_t q0 = std::min<int64_t>(nq, per * i), q1 = std::min<int64_t>(nq, per * (i + 1));
        if (q1 <= q0) return;
        rc[i] = fn(h->rep[i], q0, q1);
        if (rc[i]) msg[i] = t_err;
    };
    for (int i = 1; i < nd; i++) th.emplace_back(work, i);
    work(0);
    for (auto &t : th) t.join();
    for (int i = 0; i < nd; i++)
        if (rc[i]) { tm_set_error("device %d: %s", h->rep[i]->device, msg[i].c_str()); return rc[i]; }
    return 0;
}
// queries per device below which a host batch goes to ONE device (TM_SHARD_MIN in the environment: tests)
static int64_t shard_min() {
    static const int64_t v = getenv("TM_SHARD_MIN") && atoll(getenv("TM_SHARD_MIN")) > 0 ? atoll(getenv("TM_SHARD_MIN")) : (1 << 16);
    return v;
}
#define TM_SHARD_MIN shard_min()

extern "C" {

const char *tm_last_error(void) { return t_err; }
int tm_version(void) { return 200; }

int tm_set_devices(const int *devices, int n) {
    if (n < 1 || n > TM_MAX_DEV || !devices) { tm_set_error("tm_set_devices: need 1..%d devices", TM_MAX_DEV); return TM_EINVAL; }
    std::vector<int> list(devices, devices + n);
    std::lock_guard<std::mutex> lk(g_init_mu);
    if (g_active == list) return 0;
    if (g_live_handles.load() != 0) {
        tm_set_error("tm_set_devices: %d dataset handle(s) alive on the current device set", g_live_handles.load());
        return TM_EINVAL;
    }
    return activate(list);
}

int tm_set_device(int device) { return tm_set_devices(&device, 1); }

int tm_get_devices(int *devices, int cap) {
    if (!tm_primary()) return TM_ENODEVICE;
    const std::vector<int> a = active_devices();
    for (int i = 0; i < (int)a.size() && i < cap; i++) devices[i] = a[i];
    return (int)a.size();
}

double tm_comm_init_ms(void) { return g_nccl_init_ms; }

int tm_set_engine(int engine) {
    if (engine < TM_ENGINE_AUTO || engine > TM_ENGINE_TC) { tm_set_error("bad engine"); return TM_EINVAL; }
    g_engine = engine;
    return 0;
}

int tm_set_stream(void *cuda_stream) {
    TM_API_GUARD;
    int r = tm_ensure_device();
    if (r) return r;
    cudaStreamSynchronize(t_dev->stream);
    t_dev->stream = cuda_stream ? static_cast<cudaStream_t>(cuda_stream) : t_dev->own_stream;
    t_dev->user_stream = cuda_stream != nullptr;
    return 0;
}

int tm_device_sync(void) {
    if (!tm_primary()) return TM_ENODEVICE;
    int r = 0;
    for (int dev : active_devices()) {
        TmApiGuard g(g_ctx[dev]);
        int r2 = tm_check(cudaDeviceSynchronize(), "cudaDeviceSynchronize");
        if (!r) r = r2;
    }
    return r;
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
    if (bytes && tm_check(cudaMemcpyAsync(h->rows_f32, rows, bytes, cudaMemcpyDefault, g_stream), "copy rows")) {
        free_handle(h);
        return nullptr;
    }
    return finish_handle(h);
}

ANNkdtree *ann_kdtree_create(float **pa, int n, int dd, int bs, int split) {
    (void)bs; (void)split;
    if (n < 0 || dd < 1 || (n > 0 && !pa)) { tm_set_error("ann_kdtree_create: bad arguments"); return nullptr; }
    // rows are separately allocated Pascal dynamic arrays (main.pas:3757, 3955): gather them
    std::vector<float> flat((size_t)n * dd);
    for (int i = 0; i < n; i++) memcpy(flat.data() + (size_t)i * dd, pa[i], sizeof(float) * dd);
    return tm_dataset_create(flat.data(), n, dd, 0);
}

void ann_kdtree_destroy(ANNkdtree *akd) {
    if (!akd || akd->magic != TM_MAGIC) return; // NULL-safe like the reference wrapper
    if (akd->owner) akd = akd->owner;
    {
        TM_API_GUARD_H(akd);
        cudaStreamSynchronize(g_stream);
    }
    free_handle(akd);
}

int64_t tm_dataset_rows(const ANNkdtree *akd) { return valid(akd) ? -1 : akd->n; }
int tm_dataset_dim(const ANNkdtree *akd) { return valid(akd) ? -1 : akd->dd; }
int tm_dataset_replicas(const ANNkdtree *akd) { return valid(akd) ? -1 : (akd->owner ? akd->owner->nrep : akd->nrep); }

int tm_dataset_read(const ANNkdtree *akd, float *h_rows) { return tm_dataset_read_replica(akd, 0, h_rows); }

int tm_dataset_read_replica(const ANNkdtree *akd, int replica, float *h_rows) {
    int r = valid(akd);
    if (r) return r;
    if (akd->owner) akd = akd->owner;
    if (replica < 0 || replica >= akd->nrep) { tm_set_error("replica %d outside [0, %d)", replica, akd->nrep); return TM_EINVAL; }
    akd = akd->rep[replica];
    TM_API_GUARD_H(akd);
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
    int r = valid(akd);
    if (r) return r;
    akd = pick_replica(akd, d_q);
    TM_API_GUARD_H(akd);
    if ((r = tm_ensure_device())) return r;
    if (nq < 0) { tm_set_error("nq < 0"); return TM_EINVAL; }
    if (nq == 0) return 0;
    if ((r = sync_in())) return r;
    if ((r = search_dev(akd, d_q, nq, k, d_idxs, d_errs))) return r;
    return sync_out_dev();
}

// one device: the batch goes through in chunks so that the temporaries stay bounded
// Host batch on one device, in chunks of 2^20 queries.  More than one chunk: two input buffers, the upload of chunk
// i + 1 rides on the device's copy stream under the search of chunk i (C5, 768 B per query: 15 ms of PCIe per 350 ms
// chunk that were paid in full on the one stream).
static int search_batch_one(ANNkdtree *akd, const float *h_q, int64_t nq, int k, int *h_idxs, float *h_errs) {
    InflightMark mark(akd);
    TM_API_GUARD_H(akd);
    int r = tm_ensure_device();
    if (r) return r;
    const int64_t chunk = std::min<int64_t>(nq, 1 << 20);
    const bool piped = nq > chunk;
    const size_t row = sizeof(float) * (size_t)akd->dd;
    DevBuf q[2], i, e;
    if ((r = q[0].alloc(row * (size_t)chunk))) return r;
    if (piped && (r = q[1].alloc(row * (size_t)chunk))) return r;
    if ((r = i.alloc(sizeof(int) * (size_t)chunk * k))) return r;
    if ((r = e.alloc(sizeof(float) * (size_t)chunk * k))) return r;
    struct Events { // RAII: early returns must not leak them
        cudaEvent_t in[2] = {nullptr, nullptr}, done[2] = {nullptr, nullptr}, ready = nullptr;
        ~Events() { for (cudaEvent_t ev : {in[0], in[1], done[0], done[1], ready}) if (ev) cudaEventDestroy(ev); }
    } ev;
    cudaStream_t cs = g_stream;
    if (piped) {
        TmHostPath &hp = t_dev->hp;
        if (!hp.cs) TM_CUDA(cudaStreamCreateWithFlags(&hp.cs, cudaStreamNonBlocking));
        cs = hp.cs;
        for (cudaEvent_t *p : {&ev.in[0], &ev.in[1], &ev.done[0], &ev.done[1], &ev.ready}) TM_CUDA(cudaEventCreateWithFlags(p, cudaEventDisableTiming));
        TM_CUDA(cudaEventRecord(ev.ready, g_stream)); // the buffers come from the stream-ordered pool of g_stream
        TM_CUDA(cudaStreamWaitEvent(cs, ev.ready, 0));
    }
    auto upload = [&](int64_t q0, int slot) -> int {
        const int64_t c = std::min<int64_t>(chunk, nq - q0);
        if (piped) TM_CUDA(cudaStreamWaitEvent(cs, ev.done[slot], 0)); // the search that read this slot has finished (no-op at first)
        TM_CUDA(cudaMemcpyAsync(q[slot].p, h_q + q0 * akd->dd, row * (size_t)c, cudaMemcpyHostToDevice, cs));
        if (piped) TM_CUDA(cudaEventRecord(ev.in[slot], cs));
        return 0;
    };
    if ((r = upload(0, 0))) return r;
    int n = 0;
    for (int64_t q0 = 0; q0 < nq; q0 += chunk, n++) {
        const int64_t c = std::min<int64_t>(chunk, nq - q0);
        const int slot = piped ? (n & 1) : 0;
        if (piped && q0 + chunk < nq && (r = upload(q0 + chunk, slot ^ 1))) break;
        if (piped) TM_CUDA(cudaStreamWaitEvent(g_stream, ev.in[slot], 0));
        if ((r = search_dev(akd, q[slot].as<float>(), c, k, i.as<int>(), e.as<float>()))) break;
        if (piped) TM_CUDA(cudaEventRecord(ev.done[slot], g_stream));
        TM_CUDA(cudaMemcpyAsync(h_idxs + q0 * k, i.p, sizeof(int) * (size_t)c * k, cudaMemcpyDeviceToHost, g_stream));
        if (h_errs) TM_CUDA(cudaMemcpyAsync(h_errs + q0 * k, e.p, sizeof(float) * (size_t)c * k, cudaMemcpyDeviceToHost, g_stream));
    }
    if (piped) cudaStreamSynchronize(cs); // nothing of this call may outlive its buffers, on error either
    const int rs = sync_out();
    return r ? r : rs;
}

int ann_kdtree_search_batch(ANNkdtree *akd, const float *h_q, int64_t nq, int k, int *h_idxs, float *h_errs) {
    int r = valid(akd);
    if (r) return r;
    if (nq < 0 || k < 1 || k > 16) { tm_set_error("bad nq/k"); return TM_EINVAL; }
    if (nq == 0) return 0;
    if (akd->owner) akd = akd->owner;
    if (akd->nrep > 1 && nq >= TM_SHARD_MIN * akd->nrep) {
        const int dd = akd->dd;
        return run_sharded(akd, nq, [=](ANNkdtree *rep, int64_t q0, int64_t q1) {
            return search_batch_one(rep, h_q + q0 * dd, q1 - q0, k, h_idxs + q0 * k, h_errs ? h_errs + q0 * k : nullptr);
        });
    }
    return search_batch_one(pick_replica(akd, nullptr), h_q, nq, k, h_idxs, h_errs);
}

namespace {
bool combine_on();
bool combiner_idle(TmDevice *d);
int search_one_combined(ANNkdtree *akd, const float *q, int k, int *idxs, float *errs);
} // namespace

// legacy single-query entry points: re-entrant.  Concurrent callers are coalesced into one device batch by the
// device's worker (api_combine.inc); TM_COMBINE=0 serves them one by one under the device lock.
static int search_one(ANNkdtree *akd, const float *q, int k, int *idxs, float *errs) {
    if (valid(akd)) return -1;
    if (k < 1 || k > 16 || !q) return -1;
    akd = pick_replica(akd, nullptr);
    InflightMark mark(akd);
    {
        TmDevice *dv = g_ctx[akd->device];
        const bool contended = dv->inflight.load(std::memory_order_relaxed) > 1 || !combiner_idle(dv);
        if (combine_on() && akd->dd <= 256 && contended) return search_one_combined(akd, q, k, idxs, errs);
    }
    TM_API_GUARD_H(akd);
    if (tm_ensure_device()) return -1;
    std::lock_guard<std::mutex> lk(*akd->single_mu);
    if (!akd->h_pin) { // scratch of the one-query entry points, on first use (pinned allocation is slow)
        if (tm_dev_malloc(&akd->d_q1, sizeof(float) * akd->dd) || tm_dev_malloc(&akd->d_i1, sizeof(int) * 16) ||
            tm_dev_malloc(&akd->d_e1, sizeof(float) * 16) ||
            cudaMallocHost(&akd->h_pin, sizeof(float) * (akd->dd + 32)) != cudaSuccess) {
            akd->h_pin = nullptr;
            tm_set_error("single-query scratch allocation failed");
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
    (void)eps;
    int idx = -1;
    float e = 0.f; // the reference wrapper pre-sets idx = -1, dist = 0
    if (search_one(akd, q, 1, &idx, &e)) { if (err) *err = 0.f; return -1; }
    if (err) *err = e;
    return idx;
}

int ann_kdtree_pri_search(ANNkdtree *akd, float *q, float eps, float *err) { return ann_kdtree_search(akd, q, eps, err); }

int ann_kdtree_search_multi(ANNkdtree *akd, int *idxs, float *errs, int cnt, float *q, float eps) {
    (void)eps;
    if (cnt <= 0 || !idxs) return 0;
    // like ANN's wrapper, the outputs are pre-set (-1, 0): a failed call must not leave the caller's stack arrays
    // uninitialised (UseOne indexes TRToTileIdx with them, main.pas:3838)
    for (int j = 0; j < cnt; j++) { idxs[j] = -1; if (errs) errs[j] = 0.f; }
    if (cnt > 16) { tm_set_error("ann_kdtree_search_multi: cnt > 16 unsupported"); return 0; }
    search_one(akd, q, cnt, idxs, errs);
    return 0;
}

// pinned host memory for the caller's frame buffers (full-rate, asynchronous PCIe copies from every device)
void *tm_host_alloc(size_t bytes) {
    if (!tm_primary()) return nullptr;
    void *p = nullptr;
    if (tm_check(cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocPortable), "cudaHostAlloc")) return nullptr;
    return p;
}
void tm_host_free(void *p) {
    if (p) cudaFreeHost(p);
}

int tm_get_stats(tm_stats *out) {
    if (!out) return TM_EINVAL;
    memset(out, 0, sizeof *out);
    out->kernel_launches = g_launches.load();
    out->queries = g_queries.load();
    out->fallback_queries = g_fallbacks.load();
    out->coarse_launches = g_coarse_launches.load();
    for (TmDevice *d : g_ctx) {
        if (!d) continue;
        TmApiGuard g(d);
        prof_collect(d);
        std::lock_guard<std::mutex> lk(d->stats_mu);
        out->desc_ms += d->ms[0];
        out->search_ms += d->ms[1];
        out->coarse_ms += d->ms[2];
        out->rerank_ms += d->ms[3];
        if (d->audit) {
            unsigned long long a[2] = {0, 0};
            cudaStreamSynchronize(d->stream);
            if (cudaMemcpy(a, d->audit, 16, cudaMemcpyDeviceToHost) == cudaSuccess) {
                const unsigned int bits = (unsigned int)a[0];
                float ratio;
                memcpy(&ratio, &bits, 4);
                if (ratio > out->max_error_ratio) out->max_error_ratio = ratio;
                out->audited_candidates += (int64_t)a[1];
            }
        }
    }
    return 0;
}

int tm_reset_stats(void) {
    g_launches = 0; g_queries = 0; g_fallbacks = 0; g_coarse_launches = 0;
    for (TmDevice *d : g_ctx) {
        if (!d) continue;
        TmApiGuard g(d);
        prof_collect(d);
        std::lock_guard<std::mutex> lk(d->stats_mu);
        for (float &m : d->ms) m = 0;
        if (d->audit) cudaMemsetAsync(d->audit, 0, 16, d->stream);
    }
    return 0;
}

int tm_set_profiling(int on) {
    g_profiling = on < 0 ? 0 : on;
    return 0;
}

} // extern "C"

#include "api_fused.inc"
#include "api_combine.inc"
#include "api_keyframe.inc"
#include "api_kmodes.inc"
