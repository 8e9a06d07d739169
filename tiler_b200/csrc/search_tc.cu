// search_tc.cu — tensor-core FrameTiling search (sm_100a: tcgen05 + TMEM + TMA).
//
// Replaces the per-position exact 1-NN of DoFrameTiling (reference main.pas:4020-4038,
// FFI ann_kdtree_search extern.pas:65) and the k = 8 pre-filter of UseOne (main.pas:3830).
//
// Coarse pass.  Rows and queries are centred on the dataset's column mean (distance-
// invariant) and rounded to fp16; the kernel computes, for every (query, row),
//     s~ = ||t~||^2 - 2 q~.t~  =  ||q~ - t~||^2 - ||q~||^2
// as a dense contraction on the 5th-gen tensor cores.  The kernel (search_tc2_kernel.inc) runs as persistent
// clusters of two CTAs: each CTA keeps 128 queries resident in shared memory and streams ITS 128 rows of every
// 256-row tileset tile through an 8-deep TMA ring; the leader issues tcgen05.mma.cta_group::2 (M256 N256 K16,
// fp32 accumulate) into two TMEM accumulator stages (2 x 256 columns), and eight epilogue warps per CTA read them
// back with tcgen05.ld (thread = query row, warp = lane quarter x column half) and keep each query's best rows.
// The Q x N score matrix never reaches HBM.
//
// Scan order (k = 1).  The CTA-pair kernel's k = 1 variant does not spend an MMA on ||t~||^2 (1/13 of the tensor
// time): its accumulator holds -2 q~.t~ only, and the epilogue's fast path tests
//     min(acc over 32 columns) + (smallest norm of the warp's 128 columns) < tau,
// a necessary condition that costs one add per tile; the exact norm is added only in the rare slow path.  For the
// test to be tight the rows of a tile must have nearly equal norms, so k = 1 scans a second copy of the operand in
// SCAN ORDER: rows sorted by ||t~||^2 (stable), cut into 256-row tiles, tiles dealt in a strided order so that
// every stretch of the scan (and every row split) sees the whole norm range - in plain ascending order every
// query's best score falls steadily along the scan and the slow path runs three times as often.  perm[] maps scan
// position -> dataset row; the kernel translates when it writes its candidate lists.  k > 1 keeps the dataset
// order and the norm MMA: near-identical rows (equal norms) would otherwise share one candidate list.
//
// Chained plan (tilesets beyond the L2).  One work item per query block lets the CTA pairs drift apart along the scan
// until the operand streams from HBM for most of them.  The rows are then cut into L2-sized splits, the items run
// split-major, and every item continues the candidate lists / ring and the threshold of its query block's previous
// split (TcPlan.chain, search_tc.h; hand-over through per-block ready counters): same scan, same lists, the operand
// read from HBM about once per pass instead of once per few query blocks.
//
// Exactness.  s~ + ||q~||^2 is the exact squared distance between the ROUNDED vectors up to
// fp32 summation error, so by the triangle inequality every row that is not a candidate
// has true distance >= (sqrt(tau + ||q~||^2 - a) - e)^2 with tau the final admission threshold,
// e = ||q~ - q16|| + max ||t~ - t16|| (both MEASURED: per query by the operand code, per dataset at build) and a the
// fp32 accumulation slack (search_tc.h states the model and its assumption).  rerank.cu recomputes
// the candidates in exact float32 (the reference's arithmetic), and queries whose k-th
// best exact distance is not strictly below that bound are re-run on the exact SIMT
// engine.  Results are therefore identical to the SIMT engine's, not approximately equal.
#include "tm_common.cuh"
#include "ptx_sm100.cuh"
#include "search_tc.h"
#include <cudaTypedefs.h>
#include <cub/device/device_radix_sort.cuh>
#include <math.h>
#include <algorithm>

namespace {

// ------------------------------------------------------------------ operand preparation
constexpr int PREP_ROWS = 256; // rows per partial-sum block

__global__ void colsum_partial_kernel(const float *__restrict__ rows, int64_t n, int dd, double *__restrict__ part) {
    int64_t r0 = (int64_t)blockIdx.x * PREP_ROWS;
    int64_t r1 = min(n, r0 + PREP_ROWS);
    for (int j = threadIdx.x; j < dd; j += blockDim.x) {
        double acc = 0.0;
        for (int64_t r = r0; r < r1; r++) acc += (double)rows[r * dd + j];
        part[(int64_t)blockIdx.x * dd + j] = acc;
    }
}

// one block per column: strided partial sums per thread, then a fixed-shape tree (deterministic)
__global__ void colmean_final_kernel(const double *__restrict__ part, int nblocks, int64_t n, int dd,
                                     float *__restrict__ mean) {
    __shared__ double sh[128];
    const int j = blockIdx.x;
    double acc = 0.0;
    for (int b = threadIdx.x; b < nblocks; b += 128) acc += part[(int64_t)b * dd + j];
    sh[threadIdx.x] = acc;
    __syncthreads();
    for (int o = 64; o > 0; o >>= 1) {
        if ((int)threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) mean[j] = n > 0 ? (float)(sh[0] / (double)n) : 0.f;
}

// one warp per row: centred fp16 operand (zero padded to kp), ||t~||^2 (fp32 of the rounded
// values, what the GEMM sees), exact-ish ||t'|| for the slack, range check for fp16
__global__ void operand_kernel(const float *__restrict__ rows, int64_t n, int64_t n_pad, int dd, int kp,
                               const float *__restrict__ mean, __half *__restrict__ out16,
                               float *__restrict__ norm16, float *__restrict__ norm_exact,
                               float *__restrict__ dnorm, float pad_norm,
                               float out_scale, unsigned int *__restrict__ max_norm_bits,
                               unsigned int *__restrict__ max_abs_bits) {
    const int lane = threadIdx.x & 31, wpb = blockDim.x / 32;
    float wmax_norm = 0.f, wmax_abs = 0.f, wmax_n16 = 0.f, wmax_dn = 0.f; // this warp's maxima (dataset build only)
    for (int64_t r = (int64_t)blockIdx.x * wpb + (threadIdx.x / 32); r < n_pad; r += (int64_t)gridDim.x * wpb) {
        float s16 = 0.f, sex = 0.f, sde = 0.f, mabs = 0.f;
        for (int j = lane; j < kp; j += 32) {
            float v = 0.f;
            if (r < n && j < dd) v = rows[r * dd + j] - mean[j];
            __half h = __float2half_rn(v);
            float hv = __half2float(h);
            out16[r * kp + j] = __float2half_rn(out_scale * hv); // out_scale is a power of two: exact
            s16 += hv * hv;
            sex += v * v;
            sde += (v - hv) * (v - hv); // v - hv is exact in fp32
            mabs = fmaxf(mabs, fabsf(v));
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            s16 += __shfl_xor_sync(0xffffffffu, s16, o);
            sex += __shfl_xor_sync(0xffffffffu, sex, o);
            sde += __shfl_xor_sync(0xffffffffu, sde, o);
            mabs = fmaxf(mabs, __shfl_xor_sync(0xffffffffu, mabs, o));
        }
        if (lane == 0) {
            norm16[r] = (r < n) ? s16 : pad_norm;
            if (norm_exact) norm_exact[r] = sqrtf(sex);
            if (dnorm) dnorm[r] = sqrtf(sde);
        }
        if (r < n) { // NaN-propagating maxima: a NaN row must reach the host check (its bit pattern wins atomicMax)
            const float en = sqrtf(sex), dn = sqrtf(sde);
            wmax_norm = (en != en || en > wmax_norm) ? en : wmax_norm;
            wmax_abs = (mabs != mabs || mabs > wmax_abs) ? mabs : wmax_abs;
            wmax_n16 = (s16 != s16 || s16 > wmax_n16) ? s16 : wmax_n16;
            wmax_dn = (dn != dn || dn > wmax_dn) ? dn : wmax_dn;
        }
    }
    // one set of atomics per warp (the per-row version serialised 10^5-10^6 atomics on three addresses)
    if (lane == 0) {
        if (max_norm_bits) atomicMax(max_norm_bits, __float_as_uint(wmax_norm));
        if (max_abs_bits) {
            atomicMax(max_abs_bits, __float_as_uint(wmax_abs));
            atomicMax(max_abs_bits + 1, __float_as_uint(wmax_n16));
            atomicMax(max_abs_bits + 2, __float_as_uint(wmax_dn));
        }
    }
}

// ||t~||^2 / S as three fp16 parts in K columns 0..2 of a [128 rows][16] K-major no-swizzle operand tile
// (8-row x 16-byte core matrices: byte = (k/8)*2048 + (r/8)*128 + (r%8)*16 + (k%8)*2); one extra
// tcgen05.mma against an all-[S,S,S,0..] A tile adds the norm to every accumulator column.
__global__ void normtile_kernel(const float *__restrict__ norm16, int64_t n_pad, float inv_scale,
                                __half *__restrict__ tiles) {
    int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n_pad) return;
    const float v = norm16[r] * inv_scale;
    __half h0 = __float2half_rn(v), h1 = __float2half_rn(0.f), h2 = h1;
    if (v < 60000.f) {
        const float r1 = v - __half2float(h0);
        h1 = __float2half_rn(r1);
        h2 = __float2half_rn(r1 - __half2float(h1));
    } else {
        h0 = __ushort_as_half(0x7c00); // +inf: padding rows can never be admitted
    }
    __half *t = tiles + (r / 128) * 2048;
    const int rr = (int)(r % 128);
    uint4 lo = make_uint4(0, 0, 0, 0);
    lo.x = (uint32_t)__half_as_ushort(h0) | ((uint32_t)__half_as_ushort(h1) << 16);
    lo.y = (uint32_t)__half_as_ushort(h2);
    *reinterpret_cast<uint4 *>(t + (rr / 8) * 64 + (rr % 8) * 8) = lo;                 // K core matrix 0
    *reinterpret_cast<uint4 *>(t + 1024 + (rr / 8) * 64 + (rr % 8) * 8) = make_uint4(0, 0, 0, 0); // core 1
}

__global__ void iota_kernel(int *__restrict__ v, int64_t n) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) v[i] = (int)i;
}

// one thread per 16 bytes of the fp16 operand: rows, norms and the position -> row table in scan order.
// Scan tile j holds the sorted 256-row block (j * stride) mod ntiles.
__global__ void scan_order_kernel(const int *__restrict__ sorted_row, const float *__restrict__ norm_in,
                                  const uint4 *__restrict__ in16, int64_t n, int64_t n_pad, int row_vec, int64_t stride,
                                  uint4 *__restrict__ out16, float *__restrict__ norms, int *__restrict__ perm) {
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n_pad * row_vec) return;
    const int64_t pos = e / row_vec;
    const int j = (int)(e - pos * row_vec);
    const int64_t ntiles = n_pad >> 8, tile = pos >> 8;
    const int64_t rank = (((tile * stride) % ntiles) << 8) + (pos & 255);
    const int src = sorted_row[rank];
    out16[e] = in16[(int64_t)src * row_vec + j];
    if (j == 0) {
        norms[pos] = norm_in[src];
        perm[pos] = src < n ? src : -1;
    }
}

PFN_cuTensorMapEncodeTiled_v12000 get_encode() {
    static PFN_cuTensorMapEncodeTiled_v12000 fn = nullptr;
    if (fn) return fn;
    void *p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) != cudaSuccess ||
        qres != cudaDriverEntryPointSuccess)
        return nullptr;
    fn = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(p);
    return fn;
}

// fp16 [rows][kp] row-major, box = 64 (K) x 128 (rows), SWIZZLE_128B
int make_tmap(CUtensorMap *m, const __half *base, int64_t rows, int kp) {
    auto enc = get_encode();
    if (!enc) { tm_set_error("cuTensorMapEncodeTiled unavailable"); return TM_ECUDA; }
    cuuint64_t dims[2] = {(cuuint64_t)kp, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)kp * sizeof(__half)};
    cuuint32_t box[2] = {64, 128};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<__half *>(base), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { tm_set_error("cuTensorMapEncodeTiled failed (%d)", (int)r); return TM_ECUDA; }
    return 0;
}

// norm-tile images as a 2-D tensor: fp16 [tiles * 8][256], box = 256 x 8 = one 4 KB image (used by the
// CTA-pair kernel, whose loads must be tensor copies to signal the leader CTA's barrier)
int make_tmap_norm(CUtensorMap *m, const __half *base, int64_t tiles) {
    auto enc = get_encode();
    if (!enc) { tm_set_error("cuTensorMapEncodeTiled unavailable"); return TM_ECUDA; }
    cuuint64_t dims[2] = {256, (cuuint64_t)tiles * 8};
    cuuint64_t strides[1] = {256 * sizeof(__half)};
    cuuint32_t box[2] = {256, 8};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<__half *>(base), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { tm_set_error("cuTensorMapEncodeTiled(norm tiles) failed (%d)", (int)r); return TM_ECUDA; }
    return 0;
}

} // namespace

namespace {
struct TmpBuf { // stream-ordered temporary, released on every return path
    void *p = nullptr;
    ~TmpBuf() { tm_dev_free(p); }
    template <class T> T *as() { return static_cast<T *>(p); }
    void *release() { void *q = p; p = nullptr; return q; }
};
} // namespace

// (re)build the TMA descriptors of a handle whose operand buffers exist on the calling thread's device
int tm_dataset_tensor_maps(ANNkdtree *h) {
    if (!h->tmap_rows) h->tmap_rows = malloc(3 * sizeof(CUtensorMap)); // [0] rows_f16, [1] norm tiles, [2] rows_f16s (k = 1 scan order)
    int r = make_tmap(static_cast<CUtensorMap *>(h->tmap_rows), h->rows_f16, h->n_pad, h->kp);
    if (!r) r = make_tmap_norm(static_cast<CUtensorMap *>(h->tmap_rows) + 1, h->norm_tiles, h->n_pad / 128);
    if (!r) r = make_tmap(static_cast<CUtensorMap *>(h->tmap_rows) + 2, h->rows_f16s, h->n_pad, h->kp);
    return r;
}

// Called at dataset creation (rows_f32 already queued on s).  Leaves rows_f16 == nullptr when
// the tensor-core path does not apply (dd > 192, empty set, values outside the fp16 range).
// On an error return the caller frees the handle (and with it whatever was attached to it here).
int tm_prepare_dataset(ANNkdtree *h, cudaStream_t s) {
    h->rows_f16 = nullptr;
    if (h->n < 1 || h->dd > 192) return 0;
    const int kp = ((h->dd + 63) / 64) * 64;
    const int64_t n_pad = ((h->n + 255) / 256) * 256; // whole 256-row tiles of the CTA-pair kernel
    const int nblocks = (int)((h->n + PREP_ROWS - 1) / PREP_ROWS);
    TmpBuf part_b, mx_b, r16_b;
    TM_CUDA(tm_dev_malloc_raw(&part_b.p, sizeof(double) * (size_t)nblocks * h->dd));
    TM_CUDA(tm_dev_malloc_raw(&mx_b.p, 4 * sizeof(unsigned int)));
    TM_CUDA(tm_dev_malloc(&h->mean, sizeof(float) * h->dd));
    TM_CUDA(tm_dev_malloc(&h->norms, sizeof(float) * n_pad));
    TM_CUDA(tm_dev_malloc_raw(&r16_b.p, sizeof(__half) * (size_t)n_pad * kp));
    double *part = part_b.as<double>();
    unsigned int *mx = mx_b.as<unsigned int>();
    __half *r16 = r16_b.as<__half>();
    TM_CUDA(cudaMemsetAsync(mx, 0, 4 * sizeof(unsigned int), s));
    colsum_partial_kernel<<<nblocks, 192, 0, s>>>(h->rows_f32, h->n, h->dd, part);
    colmean_final_kernel<<<h->dd, 128, 0, s>>>(part, nblocks, h->n, h->dd, h->mean);
    operand_kernel<<<(unsigned)std::min<int64_t>((n_pad + 7) / 8, 148 * 32), 256, 0, s>>>(h->rows_f32, h->n, n_pad, h->dd, kp, h->mean, r16,
                                                             h->norms, nullptr, nullptr, INFINITY, -2.f, mx, mx + 1);
    tm_count_launch(3);
    TM_CUDA(cudaGetLastError());
    unsigned int hmx[4];
    TM_CUDA(cudaMemcpyAsync(hmx, mx, sizeof hmx, cudaMemcpyDeviceToHost, s));
    TM_CUDA(cudaStreamSynchronize(s));
    float max_norm, max_abs;
    memcpy(&max_norm, &hmx[0], 4);
    memcpy(&max_abs, &hmx[1], 4);
    float max_n16;
    memcpy(&max_n16, &hmx[2], 4);
    // norm scale S (power of two, representable in fp16) so that ||t~||^2 / S fits fp16
    float S = 1.f;
    while (max_n16 / S > 32768.f && S < 32768.f) S *= 2.f;
    if (!(max_abs < 30000.f) || !(max_n16 / S < 60000.f)) return 0; // outside the fp16 envelope (or NaN): exact engine only
    TM_CUDA(tm_dev_malloc(&h->norm_tiles, (size_t)(n_pad / 128) * 4096));
    normtile_kernel<<<(unsigned)((n_pad + 255) / 256), 256, 0, s>>>(h->norms, n_pad, 1.f / S, h->norm_tiles);
    tm_count_launch();
    TM_CUDA(cudaGetLastError());
    h->norm_scale = S;
    h->n_pad = n_pad;
    h->kp = kp;
    h->max_norm = max_norm;
    memcpy(&h->max_dnorm, &hmx[3], 4);
    {   // k = 1 scan order (header comment): stable sort by ||t~||^2 - non-negative floats order like their bit
        // patterns, +inf padding last - then rows / norms / perm gathered tile by tile
        TmpBuf k1_b, v0_b, v1_b, tmp_b;
        size_t tmp_bytes = 0;
        TM_CUDA(tm_dev_malloc_raw(&k1_b.p, sizeof(uint32_t) * n_pad));
        TM_CUDA(tm_dev_malloc_raw(&v0_b.p, sizeof(int) * n_pad));
        TM_CUDA(tm_dev_malloc_raw(&v1_b.p, sizeof(int) * n_pad));
        TM_CUDA(tm_dev_malloc(&h->norms_s, sizeof(float) * n_pad));
        TM_CUDA(tm_dev_malloc(&h->perm, sizeof(int) * n_pad));
        TM_CUDA(tm_dev_malloc(&h->rows_f16s, sizeof(__half) * (size_t)n_pad * kp));
        uint32_t *k1 = k1_b.as<uint32_t>();
        int *v0 = v0_b.as<int>(), *v1 = v1_b.as<int>();
        const uint32_t *k0 = reinterpret_cast<const uint32_t *>(h->norms);
        iota_kernel<<<(unsigned)((n_pad + 255) / 256), 256, 0, s>>>(v0, n_pad);
        TM_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, k0, k1, v0, v1, (int)n_pad, 0, 32, s));
        TM_CUDA(tm_dev_malloc_raw(&tmp_b.p, tmp_bytes ? tmp_bytes : 4));
        TM_CUDA(cub::DeviceRadixSort::SortPairs(tmp_b.p, tmp_bytes, k0, k1, v0, v1, (int)n_pad, 0, 32, s));
        const int64_t ntiles = n_pad >> 8;
        int64_t stride = (int64_t)((double)ntiles * 0.6180339887) | 1; // odd, near the golden section of the tile count
        auto gcd = [](int64_t a, int64_t b) { while (b) { const int64_t t = a % b; a = b; b = t; } return a; };
        while (gcd(stride, ntiles) != 1) stride += 2;
        const int row_vec = kp / 8; // 16-byte pieces per operand row
        const int64_t total = n_pad * row_vec;
        scan_order_kernel<<<(unsigned)((total + 255) / 256), 256, 0, s>>>(v1, h->norms, reinterpret_cast<const uint4 *>(r16), h->n, n_pad,
                                                                          row_vec, stride, reinterpret_cast<uint4 *>(h->rows_f16s),
                                                                          h->norms_s, h->perm);
        tm_count_launch(2);
        TM_CUDA(cudaGetLastError());
    }
    h->rows_f16 = static_cast<__half *>(r16_b.release());
    return tm_dataset_tensor_maps(h);
}

int tm_launch_query_operand(const ANNkdtree *h, const float *d_q, int64_t nq, int64_t nq_pad, __half *d_q16,
                            float *d_qnorm16, float *d_qnorm_exact, float *d_qdnorm, cudaStream_t s) {
    operand_kernel<<<(unsigned)std::min<int64_t>((nq_pad + 7) / 8, 148 * 32), 256, 0, s>>>(d_q, nq, nq_pad, h->dd, h->kp, h->mean, d_q16,
                                                              d_qnorm16, d_qnorm_exact, d_qdnorm, 0.f, 1.f, nullptr, nullptr);
    tm_count_launch();
    return tm_check(cudaGetLastError(), "operand_kernel(q)");
}

int tm_make_query_tmap(void *tmap, const __half *d_q16, int64_t nq_pad, int kp) {
    return make_tmap(static_cast<CUtensorMap *>(tmap), d_q16, nq_pad, kp);
}

#include "search_tc_kernel.inc"
#include "search_tc2_kernel.inc"

// ------------------------------------------------------------------ host orchestration
// set by the FrameTiling path around a search whose queries are in coherence order (qorder.cu); device lock held
void tm_tc_set_coherent(bool on) { tm_dev()->tc_coherent = on; }
#define g_tc (tm_dev()->tc)

namespace {
int ensure_tc_scratch(int64_t nq_pad, int kp, size_t cand_elems, cudaStream_t s) {
    if ((size_t)nq_pad * kp > g_tc.cap_q * (size_t)(g_tc.kp ? g_tc.kp : 1) || kp != g_tc.kp || (size_t)nq_pad > g_tc.cap_q) {
        // grow-only, geometric (every step costs a synchronisation and five cudaMallocs), whole query tiles
        if (kp == g_tc.kp && (size_t)nq_pad < g_tc.cap_q + g_tc.cap_q / 2)
            nq_pad = (int64_t)(((g_tc.cap_q + g_tc.cap_q / 2 + TC_QTILE - 1) / TC_QTILE) * TC_QTILE);
        cudaStreamSynchronize(s);
        cudaFree(g_tc.q16); cudaFree(g_tc.qn16); cudaFree(g_tc.qne); cudaFree(g_tc.qde); cudaFree(g_tc.fb_list);
        g_tc.q16 = nullptr; g_tc.cap_q = 0;
        TM_CUDA(cudaMalloc(&g_tc.q16, sizeof(__half) * (size_t)nq_pad * kp));
        TM_CUDA(cudaMalloc(&g_tc.qn16, sizeof(float) * nq_pad));
        TM_CUDA(cudaMalloc(&g_tc.qne, sizeof(float) * nq_pad));
        TM_CUDA(cudaMalloc(&g_tc.qde, sizeof(float) * nq_pad));
        TM_CUDA(cudaMalloc(&g_tc.fb_list, sizeof(int) * nq_pad));
        g_tc.cap_q = nq_pad;
        g_tc.kp = kp;
    }
    if (cand_elems > g_tc.cap_c) {
        if (cand_elems < g_tc.cap_c + g_tc.cap_c / 2) cand_elems = g_tc.cap_c + g_tc.cap_c / 2;
        cudaStreamSynchronize(s);
        cudaFree(g_tc.cand_s); cudaFree(g_tc.cand_i); cudaFree(g_tc.cand_tau);
        g_tc.cand_s = nullptr; g_tc.cap_c = 0;
        TM_CUDA(cudaMalloc(&g_tc.cand_tau, sizeof(float) * cand_elems));
        TM_CUDA(cudaMalloc(&g_tc.cand_s, sizeof(float) * cand_elems));
        TM_CUDA(cudaMalloc(&g_tc.cand_i, sizeof(int) * cand_elems));
        g_tc.cap_c = cand_elems;
    }
    if (!g_tc.fb_count) {
        TM_CUDA(cudaMalloc(&g_tc.fb_count, sizeof(int)));
        TM_CUDA(cudaMallocHost(&g_tc.h_count, sizeof(int)));
    }
    return 0;
}
} // namespace

static int search_tc_impl(ANNkdtree *h, const float *d_q, int64_t nq, int k, int *d_idx, float *d_err, cudaStream_t s,
                          bool operand_ready);

int tm_search_tc(ANNkdtree *h, const float *d_q, int64_t nq, int k, int *d_idx, float *d_err, cudaStream_t s) {
    return search_tc_impl(h, d_q, nq, k, d_idx, d_err, s, false);
}

int tm_search_tc_prepared(ANNkdtree *h, const float *d_q, int64_t nq, int k, int *d_idx, float *d_err, cudaStream_t s) {
    return search_tc_impl(h, d_q, nq, k, d_idx, d_err, s, true);
}

int tm_tc_query_operand_buffers(ANNkdtree *h, int64_t nq, void **d_q16, float **d_qn16, float **d_qne, float **d_qde,
                                cudaStream_t s) {
    if (!h->rows_f16 || h->kp != h->dd || nq < 1) return 1; // fused operand rows are unpadded: kp == dd only
    const int64_t nq_pad = ((nq + TC_QTILE - 1) / TC_QTILE) * TC_QTILE;
    int r = ensure_tc_scratch(nq_pad, h->kp, 0, s);
    if (r) return r;
    if (nq_pad > nq) { // zero operand and norms for the padding queries, as operand_kernel writes them
        TM_CUDA(cudaMemsetAsync(g_tc.q16 + (size_t)nq * h->kp, 0, sizeof(__half) * (size_t)(nq_pad - nq) * h->kp, s));
        TM_CUDA(cudaMemsetAsync(g_tc.qn16 + nq, 0, sizeof(float) * (size_t)(nq_pad - nq), s));
        TM_CUDA(cudaMemsetAsync(g_tc.qne + nq, 0, sizeof(float) * (size_t)(nq_pad - nq), s));
        TM_CUDA(cudaMemsetAsync(g_tc.qde + nq, 0, sizeof(float) * (size_t)(nq_pad - nq), s));
    }
    *d_q16 = g_tc.q16;
    *d_qn16 = g_tc.qn16;
    *d_qne = g_tc.qne;
    *d_qde = g_tc.qde;
    return 0;
}

// One coarse pass + re-rank over a query batch.  Queries the certificate rejects are listed in g_tc.fb_list;
// their number is returned through *nfb_out (one host round trip).  finest_split: use the largest row split
// (most candidate lists per query) regardless of the batch size - the second chance for rejected queries.
static int tc_core(ANNkdtree *h, const float *d_q, int64_t nq, int k, int *d_idx, float *d_err, cudaStream_t s,
                   bool operand_ready, bool finest_split, int *nfb_out, int *nlists_out) {
    const int kc = (k <= 4) ? 8 : 16;
    const int nch = h->kp / 64;
    const int64_t nq_pad = ((nq + TC_QTILE - 1) / TC_QTILE) * TC_QTILE;
    const int qblocks = (int)(nq_pad / TC_QTILE);
    // the CTA-pair kernel walks 256-row tiles and keeps two candidate lists per query and row split (one per
    // accumulator column half)
    const int total_tiles = (int)(h->n_pad / T2_BN);
    // Row split.  A work item costs its tiles plus a start-up (the candidate tracking of a fresh item hits on
    // almost every group until its threshold has settled: measured 20-35 us = roughly 48 tile times), and the
    // items are dealt round-robin to the resident CTA pairs, so the pass takes rounds x (tiles per item + 48).
    // Few query blocks: split the rows so that every SM has work; many: one item per query block.
    const int units = tm_num_sms() / 2;
    const int max_split = 8;
    int nsplit = 1;
    {
        double best = 1e30;
        for (int ns = 1; ns <= max_split && ns <= (total_tiles + 3) / 4; ns++) {
            const int per = (total_tiles + ns - 1) / ns;
            const int items = qblocks * ((total_tiles + per - 1) / per);
            const double cost = (double)((items + units - 1) / units) * (per + 48);
            if (cost < best * 0.97) { best = cost; nsplit = ns; } // prefer fewer lists unless clearly better
        }
    }
    if (finest_split) nsplit = max_split;
    static const int nsplit_env = getenv("TM_TC_NSPLIT") ? atoi(getenv("TM_TC_NSPLIT")) : 0; // tuning knob
    if (nsplit_env > 0 && !finest_split) nsplit = nsplit_env;
    if (nsplit > total_tiles / 4) nsplit = total_tiles / 4;
    if (nsplit > max_split) nsplit = max_split;
    if (nsplit < 1) nsplit = 1;
    int tiles_per_split = (total_tiles + nsplit - 1) / nsplit;
    nsplit = (total_tiles + tiles_per_split - 1) / tiles_per_split;
    // k > 1 against a tileset whose operand does not fit the L2 (C5: 1 048 576 rows = 403 MB): one item per query block
    // lets the CTA pairs drift apart along the scan until every tile comes from HBM again for most of them (ncu, 1 M x
    // 1 M: 539 GB read, L2 hit rate 67 %, and the SM clock sinks under the power cap).  Chained plan: the rows are cut
    // into L2-sized splits and the items run split-major, each continuing the candidate lists and the threshold of its
    // query block's previous split (search_tc.h), so all pairs share one stretch of the operand and a query still has
    // one list per column half.  TM_TC_CHAIN=0 switches it off (A/B).
    int chain = 0;
    static const int chain_env = getenv("TM_TC_CHAIN") ? atoi(getenv("TM_TC_CHAIN")) : 1;
    static const int chain_k1 = getenv("TM_TC_CHAIN_K1") ? atoi(getenv("TM_TC_CHAIN_K1")) : 1; // k = 1 as well (A/B)
    if ((k > 1 || chain_k1) && chain_env && !finest_split && nsplit == 1 && nsplit_env <= 0 && qblocks >= 4 * units) {
        const double op_mb = (double)h->n_pad * h->kp * 2.0 / 1048576.0;
        const int target_mb = chain_env > 1 ? chain_env : 48; // operand bytes per split (TM_TC_CHAIN=<MB> to tune)
        int ns = (int)((op_mb + target_mb - 1) / target_mb);
        if (ns > 32) ns = 32; // a query keeps one list per column half whatever the number of splits
        if (ns >= 2 && total_tiles / ns >= 64) {
            chain = 1;
            nsplit = ns;
            tiles_per_split = (total_tiles + nsplit - 1) / nsplit;
            nsplit = (total_tiles + tiles_per_split - 1) / tiles_per_split;
        }
    }
    TcPlan plan{qblocks, qblocks, nsplit, tiles_per_split, nsplit, tiles_per_split, total_tiles, chain};
    // Tail (search_tc.h): with one item per block, the blocks beyond the last FULL round of CTA pairs are split finer,
    // so the last round runs tiles / ns2 + start-up instead of a whole item.  TM_TC_TAIL=0 switches it off (A/B).
    static const bool tail_on = !(getenv("TM_TC_TAIL") && atoi(getenv("TM_TC_TAIL")) == 0);
    if (tail_on && !chain && !finest_split && nsplit == 1 && nsplit_env <= 0 && qblocks > units) {
        const int full = (qblocks / units) * units, rem = qblocks - full;
        if (rem > 0) {
            int ns2 = std::min(std::min(max_split, units / rem), total_tiles / 4);
            if (ns2 >= 2) {
                int tps2 = (total_tiles + ns2 - 1) / ns2;
                ns2 = (total_tiles + tps2 - 1) / tps2;
                const double uniform = (double)((qblocks + units - 1) / units) * (total_tiles + 48);
                const double tailed = (double)(full / units) * (total_tiles + 48) + (double)((rem * ns2 + units - 1) / units) * (tps2 + 48);
                if (tailed < 0.97 * uniform) { plan.qb1 = full; plan.ns2 = ns2; plan.tps2 = tps2; }
            }
        }
    }
    const int nlists = plan.max_nlists();
    *nlists_out = nlists;

    int r = ensure_tc_scratch(nq_pad, h->kp, (size_t)plan.total_lists() * kc, s);
    if (r) return r;
    if (!operand_ready && (r = tm_launch_query_operand(h, d_q, nq, nq_pad, g_tc.q16, g_tc.qn16, g_tc.qne, g_tc.qde, s))) return r;
    CUtensorMap tq;
    if ((r = make_tmap(&tq, g_tc.q16, nq_pad, h->kp))) return r;
    const CUtensorMap &tt = *static_cast<const CUtensorMap *>(h->tmap_rows);
    const CUtensorMap &tn = static_cast<const CUtensorMap *>(h->tmap_rows)[1];
    const CUtensorMap &tts = static_cast<const CUtensorMap *>(h->tmap_rows)[2]; // k = 1: operand in scan order
    TM_CUDA(cudaMemsetAsync(g_tc.fb_count, 0, sizeof(int), s));
    if (chain) { // per query block: epilogue warps that have finished a split of it (search_tc2_kernel.inc)
        if ((size_t)qblocks > g_tc.cap_ready) {
            cudaStreamSynchronize(s);
            cudaFree(g_tc.chain_ready);
            g_tc.chain_ready = nullptr; g_tc.cap_ready = 0;
            const size_t cap = (size_t)qblocks + (size_t)qblocks / 2;
            TM_CUDA(cudaMalloc(&g_tc.chain_ready, sizeof(int) * cap));
            g_tc.cap_ready = cap;
        }
        TM_CUDA(cudaMemsetAsync(g_tc.chain_ready, 0, sizeof(int) * (size_t)qblocks, s));
    }
    // k = 1, long work items: norm-free accumulator over the operand in scan order (search_tc2_kernel.inc).  Its slow
    // path costs more per call and an item starts on a tile of one norm level, so it needs long items, longer ones when
    // the queries come in the caller's order (many slow-path calls).  Same-box A/B: sorted batch, 256 tiles per item
    // 77.2 -> 75.2 ms (C2); 1 024 tiles +4 % (C3); unsorted frame calls: 256 tiles per item 0.301 -> 0.329 ms, 340-512
    // tiles 2.24 -> 2.16 ms; 32-85 tiles per item -8 %.
    static const int nf_env = getenv("TM_TC_NF_MIN") ? atoi(getenv("TM_TC_NF_MIN")) : 0; // tuning knob
    const int nf_min = nf_env > 0 ? nf_env : (tm_dev()->tc_coherent ? 192 : 320);
    const bool normfree = k == 1 && tiles_per_split >= nf_min && h->rows_f16s != nullptr;
    g_coarse_launches.fetch_add(1, std::memory_order_relaxed);
    int tok = tm_prof_begin(2, s);
#define TC_ARGS plan, g_tc.cand_s, g_tc.cand_i, g_tc.cand_tau, g_tc.qn16, g_tc.qne, g_tc.qde, h->max_norm, h->max_dnorm, k, g_tc.chain_ready, s
#define TC_LAUNCH(NCH, KC, K1) ((K1) && normfree ? launch_tc2<NCH, KC, K1, K1>(tq, tts, tn, h->norm_scale, h->norms_s, h->perm, TC_ARGS) \
                                                : launch_tc2<NCH, KC, K1, false>(tq, tt, tn, h->norm_scale, h->norms_s, h->perm, TC_ARGS))
    if (k == 1) r = nch == 1 ? TC_LAUNCH(1, 8, true) : nch == 2 ? TC_LAUNCH(2, 8, true) : TC_LAUNCH(3, 8, true);
    else if (kc == 8) r = nch == 1 ? TC_LAUNCH(1, 8, false) : nch == 2 ? TC_LAUNCH(2, 8, false) : TC_LAUNCH(3, 8, false);
    else r = nch == 1 ? TC_LAUNCH(1, 16, false) : nch == 2 ? TC_LAUNCH(2, 16, false) : TC_LAUNCH(3, 16, false);
#undef TC_LAUNCH
#undef TC_ARGS
    tm_prof_end(tok, s);
    if (r) return r;
    tok = tm_prof_begin(3, s);
    r = tm_launch_rerank(h->rows_f32, h->dd, d_q, nq, g_tc.cand_s, g_tc.cand_i, g_tc.cand_tau, plan, kc, g_tc.qn16, g_tc.qne,
                         g_tc.qde, h->max_norm, h->max_dnorm, k, d_idx, d_err, g_tc.fb_list, g_tc.fb_count, s);
    tm_prof_end(tok, s);
    if (r) return r;
#ifdef TM_DEBUG_SWEEP
    if (tc_debug_mode()) { // the sweep modes skip candidate tracking: results are garbage, do not send them to the exact engine
        *nfb_out = 0;
        return 0;
    }
#endif
    if (TmDefer *df = tm_dev()->defer) {
        // pipelined caller (api_combine.inc): no host round trip here; the reject count and list travel with the
        // results and are handled after the batch has completed
        const int64_t nl = std::min<int64_t>(nq, df->cap);
        TM_CUDA(cudaMemcpyAsync(df->h_count, g_tc.fb_count, sizeof(int), cudaMemcpyDeviceToHost, s));
        TM_CUDA(cudaMemcpyAsync(df->d_list, g_tc.fb_list, sizeof(int) * (size_t)nl, cudaMemcpyDeviceToDevice, s));
        *nfb_out = 0;
        return 0;
    }
    TM_CUDA(cudaMemcpyAsync(g_tc.h_count, g_tc.fb_count, sizeof(int), cudaMemcpyDeviceToHost, s));
    TM_CUDA(cudaStreamSynchronize(s));
    *nfb_out = *g_tc.h_count;
    return 0;
}

namespace {
__global__ void gather_rows_kernel(const float *__restrict__ q, const int *__restrict__ list, int n, int dd,
                                   float *__restrict__ out) {
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; // one float4 each (dd % 4 == 0)
    const int per = dd >> 2;
    if (e >= (int64_t)n * per) return;
    const int i = (int)(e / per), j = (int)(e - (int64_t)i * per);
    reinterpret_cast<float4 *>(out)[e] = __ldg(reinterpret_cast<const float4 *>(q + (int64_t)list[i] * dd) + j);
}

__global__ void scatter_results_kernel(const int *__restrict__ list, int n, int k, const int *__restrict__ ti,
                                       const float *__restrict__ te, int *__restrict__ idx, float *__restrict__ err) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n * k) return;
    const int i = e / k, j = e - i * k;
    idx[(int64_t)list[i] * k + j] = ti[e];
    err[(int64_t)list[i] * k + j] = te[e];
}
} // namespace

int tm_launch_gather_rows(const float *d_rows, const int *d_list, int n, int dd, float *d_out, cudaStream_t s) {
    if (n == 0) return 0;
    const int64_t total = (int64_t)n * (dd >> 2);
    gather_rows_kernel<<<(unsigned)((total + 255) / 256), 256, 0, s>>>(d_rows, d_list, n, dd, d_out);
    tm_count_launch();
    return tm_check(cudaGetLastError(), "gather_rows_kernel");
}

int tm_launch_scatter_results(const int *d_list, int n, int k, const int *d_ti, const float *d_te, int *d_idx, float *d_err,
                              cudaStream_t s) {
    if (n == 0) return 0;
    scatter_results_kernel<<<(unsigned)(((int64_t)n * k + 255) / 256), 256, 0, s>>>(d_list, n, k, d_ti, d_te, d_idx, d_err);
    tm_count_launch();
    return tm_check(cudaGetLastError(), "scatter_results_kernel");
}

static int search_tc_impl(ANNkdtree *h, const float *d_q, int64_t nq, int k, int *d_idx, float *d_err, cudaStream_t s,
                          bool operand_ready) {
    if (!h->rows_f16 || k > 8 || ((uintptr_t)d_q & 15)) // outside the tensor-core path's envelope: exact engine
        return tm_launch_search_simt(h->rows_f32, h->n, h->dd, d_q, nq, k, d_idx, d_err, s);
    int nfb = 0, nlists0 = 0, nlists1 = 0;
    int r = tc_core(h, d_q, nq, k, d_idx, d_err, s, operand_ready, false, &nfb, &nlists0);
    if (r || nfb == 0) return r;
    static const bool trace = getenv("TM_TRACE") != nullptr;
    if (trace) fprintf(stderr, "[tmann] tc pass 1: %d of %lld queries rejected (%d lists)\n", nfb, (long long)nq, nlists0);
    // Rejected queries (candidate list too short for the rows inside the certificate's band, or an fp16
    // overflow).  Second chance on the tensor cores with the finest row split - 16 lists per query instead of
    // 2 - on a compacted copy of those queries; what is still rejected goes to the exact engine.  Worth it when
    // the exact scan of the rejects would cost more than a short tensor-core pass.
    const int finest = 16; // lists per query at the finest split
    const bool retry = (h->dd & 3) == 0 && nfb >= 64 && (double)nfb * (double)h->n >= 2.5e8 && nlists0 < finest &&
                       h->n_pad / 256 >= 32;
    if (!retry) {
        g_fallbacks.fetch_add(nfb, std::memory_order_relaxed);
        return tm_launch_search_simt_list(h->rows_f32, h->n, h->dd, d_q, g_tc.fb_list, nullptr, nfb, k, d_idx, d_err, s);
    }
    int *list0 = nullptr, *ti = nullptr;
    float *qt = nullptr, *te = nullptr;
    auto cleanup = [&]() { tm_dev_free(list0); tm_dev_free(ti); tm_dev_free(qt); tm_dev_free(te); };
    r = tm_check(tm_dev_malloc(&list0, sizeof(int) * (size_t)nfb), "tm_dev_malloc(retry list)");
    if (!r) r = tm_check(tm_dev_malloc(&qt, sizeof(float) * (size_t)nfb * h->dd), "tm_dev_malloc(retry queries)");
    if (!r) r = tm_check(tm_dev_malloc(&ti, sizeof(int) * (size_t)nfb * k), "tm_dev_malloc(retry idx)");
    if (!r) r = tm_check(tm_dev_malloc(&te, sizeof(float) * (size_t)nfb * k), "tm_dev_malloc(retry err)");
    if (r) { cleanup(); return r; }
    cudaMemcpyAsync(list0, g_tc.fb_list, sizeof(int) * (size_t)nfb, cudaMemcpyDeviceToDevice, s);
    {
        const int64_t total = (int64_t)nfb * (h->dd >> 2);
        gather_rows_kernel<<<(unsigned)((total + 255) / 256), 256, 0, s>>>(d_q, list0, nfb, h->dd, qt);
        tm_count_launch();
    }
    int nfb2 = 0;
    r = tc_core(h, qt, nfb, k, ti, te, s, false, true, &nfb2, &nlists1);
    if (trace) {
        fprintf(stderr, "[tmann] tc pass 2: %d of %d queries rejected (%d lists); first rejected queries of pass 1:", nfb2, nfb, nlists1);
        int first[16];
        const int nshow = nfb < 16 ? nfb : 16;
        cudaMemcpy(first, list0, sizeof(int) * nshow, cudaMemcpyDeviceToHost);
        for (int i = 0; i < nshow; i++) fprintf(stderr, " %d", first[i]);
        fprintf(stderr, "\n");
    }
    if (!r && nfb2 > 0) {
        g_fallbacks.fetch_add(nfb2, std::memory_order_relaxed);
        r = tm_launch_search_simt_list(h->rows_f32, h->n, h->dd, qt, g_tc.fb_list, nullptr, nfb2, k, ti, te, s);
    }
    if (!r) {
        scatter_results_kernel<<<(unsigned)(((int64_t)nfb * k + 255) / 256), 256, 0, s>>>(list0, nfb, k, ti, te, d_idx, d_err);
        tm_count_launch();
        r = tm_check(cudaGetLastError(), "scatter_results_kernel");
        if (!r) r = tm_check(cudaStreamSynchronize(s), "sync(retry)"); // the temporaries are freed below
    }
    cleanup();
    return r;
}
