// search_simt.cu — exact brute-force k-NN on CUDA cores (sm_100a).
//
// The correctness anchor of the library and the fallback for queries the
// tensor-core pass cannot certify.  For every (query, row) pair the squared L2
// is accumulated in float32 in dimension order with separate multiply and add
// (this file is compiled with -fmad=false), i.e. exactly the loop ANN's leaf
// scan performs for the reference's eps = 0 searches (main.pas:3830, 4027).
// Ties keep the lowest row index.
//
// Block = 256 threads = 16 x 16; a block owns 64 queries and walks 64-row tiles
// of its row range; each thread keeps a 4 x 4 register tile of running sums.
// Small query counts split the row range over gridDim.y and a second kernel
// merges the partial lists.
#include "tm_common.cuh"
#include <float.h>

namespace {

constexpr int TQ = 64, TR = 64;

template <int K>
struct TopK {
    float d[K];
    int i[K];
    __device__ __forceinline__ void init() {
#pragma unroll
        for (int j = 0; j < K; j++) { d[j] = INFINITY; i[j] = 0x7fffffff; }
    }
    // keeps (d, i) lexicographically ascending; strictly-better test keeps lowest index on ties
    __device__ __forceinline__ void push(float nd, int ni) {
        if (!(nd < d[K - 1] || (nd == d[K - 1] && ni < i[K - 1]))) return;
        d[K - 1] = nd; i[K - 1] = ni;
#pragma unroll
        for (int j = K - 1; j > 0; j--) {
            bool sw = d[j] < d[j - 1] || (d[j] == d[j - 1] && i[j] < i[j - 1]);
            if (sw) {
                float td = d[j]; d[j] = d[j - 1]; d[j - 1] = td;
                int ti = i[j]; i[j] = i[j - 1]; i[j - 1] = ti;
            }
        }
    }
};

// part_d / part_i: [nq][gridDim.y][K] partial lists (or the final output when gridDim.y == 1 and K == k)
template <int K>
__global__ void __launch_bounds__(256) simt_knn_kernel(const float *__restrict__ rows, int64_t n, int dd,
                                                        const float *__restrict__ q,
                                                        const int *__restrict__ qlist, int64_t nq,
                                                        int64_t rows_per_split, float *__restrict__ part_d,
                                                        int *__restrict__ part_i) {
    extern __shared__ float smem[];
    const int ddp = (dd + 3) & ~3;
    const int stride = ddp + 4;
    float *qs = smem;               // [TQ][stride]
    float *ts = smem + TQ * stride; // [TR][stride]
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
    const int64_t q0 = (int64_t)blockIdx.x * TQ;
    const int64_t r_begin = (int64_t)blockIdx.y * rows_per_split;
    const int64_t r_end = min(n, r_begin + rows_per_split);

    for (int e = threadIdx.x; e < TQ * ddp; e += 256) {
        int qi = e / ddp, j = e - qi * ddp;
        int64_t qg = q0 + qi;
        float v = 0.f;
        if (qg < nq && j < dd) {
            int64_t src = qlist ? (int64_t)qlist[qg] : qg;
            v = q[src * dd + j];
        }
        qs[qi * stride + j] = v;
    }
    TopK<K> best[4];
#pragma unroll
    for (int a = 0; a < 4; a++) best[a].init();

    for (int64_t r0 = r_begin; r0 < r_end; r0 += TR) {
        __syncthreads();
        for (int e = threadIdx.x; e < TR * ddp; e += 256) {
            int ri = e / ddp, j = e - ri * ddp;
            int64_t rg = r0 + ri;
            ts[ri * stride + j] = (rg < r_end && j < dd) ? __ldg(rows + rg * dd + j) : 0.f;
        }
        __syncthreads();
        float acc[4][4];
#pragma unroll
        for (int a = 0; a < 4; a++)
#pragma unroll
            for (int b = 0; b < 4; b++) acc[a][b] = 0.f;
        for (int j = 0; j < ddp; j += 4) {
            float4 qv[4], tv[4];
#pragma unroll
            for (int a = 0; a < 4; a++) qv[a] = *reinterpret_cast<const float4 *>(qs + (ty + 16 * a) * stride + j);
#pragma unroll
            for (int b = 0; b < 4; b++) tv[b] = *reinterpret_cast<const float4 *>(ts + (tx + 16 * b) * stride + j);
#pragma unroll
            for (int a = 0; a < 4; a++)
#pragma unroll
                for (int b = 0; b < 4; b++) {
                    float t0 = qv[a].x - tv[b].x; acc[a][b] = acc[a][b] + t0 * t0;
                    float t1 = qv[a].y - tv[b].y; acc[a][b] = acc[a][b] + t1 * t1;
                    float t2 = qv[a].z - tv[b].z; acc[a][b] = acc[a][b] + t2 * t2;
                    float t3 = qv[a].w - tv[b].w; acc[a][b] = acc[a][b] + t3 * t3;
                }
        }
#pragma unroll
        for (int b = 0; b < 4; b++) {
            int64_t rg = r0 + tx + 16 * b;
            if (rg < r_end) {
#pragma unroll
                for (int a = 0; a < 4; a++) best[a].push(acc[a][b], (int)rg);
            }
        }
    }
    // merge the 16 per-thread lists of each query through shared memory
    __syncthreads();
    float *md = smem;                                        // [TQ][16*K]
    int *mi = reinterpret_cast<int *>(smem + TQ * 16 * K);   // [TQ][16*K]
#pragma unroll
    for (int a = 0; a < 4; a++)
#pragma unroll
        for (int j = 0; j < K; j++) {
            md[(ty + 16 * a) * 16 * K + tx * K + j] = best[a].d[j];
            mi[(ty + 16 * a) * 16 * K + tx * K + j] = best[a].i[j];
        }
    __syncthreads();
    if (threadIdx.x < TQ) {
        int qi = threadIdx.x;
        int64_t qg = q0 + qi;
        if (qg < nq) {
            TopK<K> fin;
            fin.init();
            for (int e = 0; e < 16 * K; e++) fin.push(md[qi * 16 * K + e], mi[qi * 16 * K + e]);
            int64_t o = (qg * gridDim.y + blockIdx.y) * K;
#pragma unroll
            for (int j = 0; j < K; j++) { part_d[o + j] = fin.d[j]; part_i[o + j] = fin.i[j]; }
        }
    }
}

// merge [nq][splits][K] partial lists -> first k entries; writes to the original query slot
template <int K>
__global__ void simt_merge_kernel(const float *__restrict__ part_d, const int *__restrict__ part_i,
                                  const int *__restrict__ qlist, int64_t nq, int splits, int k,
                                  int *__restrict__ idx, float *__restrict__ err) {
    int64_t qg = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (qg >= nq) return;
    TopK<K> fin;
    fin.init();
    for (int e = 0; e < splits * K; e++) fin.push(part_d[qg * splits * K + e], part_i[qg * splits * K + e]);
    int64_t dst = qlist ? (int64_t)qlist[qg] : qg;
#pragma unroll
    for (int j = 0; j < K; j++)
        if (j < k) {
            idx[dst * k + j] = fin.i[j] == 0x7fffffff ? -1 : fin.i[j];
            err[dst * k + j] = fin.d[j];
        }
}

#define g_scratch (tm_dev()->simt) // split-merge scratch of the calling thread's device (device lock held)

template <int K>
int run(const float *rows, int64_t n, int dd, const float *q, const int *qlist, int64_t nq, int k, int *idx,
        float *err, cudaStream_t s) {
    const int ddp = (dd + 3) & ~3;
    size_t smem = (size_t)2 * TQ * (ddp + 4) * sizeof(float);
    size_t smem_merge = (size_t)TQ * 16 * K * 8;
    if (smem_merge > smem) smem = smem_merge;
    static std::atomic<uint32_t> attr_set{0}; // per template instance and device
    if (tm_first_on_device(attr_set))
        TM_CUDA(cudaFuncSetAttribute(simt_knn_kernel<K>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    int64_t qblocks = (nq + TQ - 1) / TQ;
    int64_t tiles = (n + TR - 1) / TR;
    int target = 2 * tm_num_sms();
    int64_t splits = 1;
    if (qblocks < target) splits = (target + qblocks - 1) / qblocks;
    if (splits > tiles) splits = tiles;
    if (splits > 256) splits = 256;
    if (splits < 1) splits = 1;
    int64_t rows_per_split = ((tiles + splits - 1) / splits) * TR;
    if (rows_per_split < TR) rows_per_split = TR;
    splits = (n + rows_per_split - 1) / rows_per_split;
    if (splits < 1) splits = 1;
    size_t need = (size_t)nq * splits * K;
    if (need > g_scratch.cap) {
        if (g_scratch.d) { cudaStreamSynchronize(s); cudaFree(g_scratch.d); cudaFree(g_scratch.i); g_scratch.d = nullptr; g_scratch.cap = 0; }
        TM_CUDA(cudaMalloc(&g_scratch.d, need * sizeof(float)));
        TM_CUDA(cudaMalloc(&g_scratch.i, need * sizeof(int)));
        g_scratch.cap = need;
    }
    dim3 grid((unsigned)qblocks, (unsigned)splits);
    simt_knn_kernel<K><<<grid, 256, smem, s>>>(rows, n, dd, q, qlist, nq, rows_per_split, g_scratch.d, g_scratch.i);
    TM_CUDA(cudaGetLastError());
    simt_merge_kernel<K><<<(unsigned)((nq + 127) / 128), 128, 0, s>>>(g_scratch.d, g_scratch.i, qlist, nq,
                                                                       (int)splits, k, idx, err);
    tm_count_launch(2);
    TM_CUDA(cudaGetLastError());
    // scratch is reused by the next call on the same stream order; keep the lock until queued
    return 0;
}

int dispatch(const float *rows, int64_t n, int dd, const float *q, const int *qlist, int64_t nq, int k,
             int *idx, float *err, cudaStream_t s) {
    if (nq == 0) return 0;
    if (dd < 1 || dd > 256) { tm_set_error("SIMT search supports 1 <= dd <= 256 (got %d)", dd); return TM_EINVAL; }
    if (k < 1 || k > 16) { tm_set_error("k must be in 1..16 (got %d)", k); return TM_EINVAL; }
    if (k == 1) return run<1>(rows, n, dd, q, qlist, nq, k, idx, err, s);
    if (k <= 2) return run<2>(rows, n, dd, q, qlist, nq, k, idx, err, s);
    if (k <= 4) return run<4>(rows, n, dd, q, qlist, nq, k, idx, err, s);
    if (k <= 8) return run<8>(rows, n, dd, q, qlist, nq, k, idx, err, s);
    return run<16>(rows, n, dd, q, qlist, nq, k, idx, err, s);
}

} // namespace

int tm_launch_search_simt(const float *d_rows, int64_t n, int dd, const float *d_q, int64_t nq, int k,
                          int *d_idx, float *d_err, cudaStream_t s) {
    return dispatch(d_rows, n, dd, d_q, nullptr, nq, k, d_idx, d_err, s);
}

int tm_launch_search_simt_list(const float *d_rows, int64_t n, int dd, const float *d_q, const int *d_qlist,
                               const int *d_qcount, int64_t list_len, int k, int *d_idx, float *d_err,
                               cudaStream_t s) {
    (void)d_qcount;
    return dispatch(d_rows, n, dd, d_q, d_qlist, list_len, k, d_idx, d_err, s);
}
