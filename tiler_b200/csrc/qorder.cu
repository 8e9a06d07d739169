// qorder.cu — coherence order of a FrameTiling query batch (DoFrameTiling hot loop, reference main.pas:4020-4038).
//
// A warp of the coarse search's epilogue serves 32 consecutive queries and pays one out-of-line call whenever
// ANY of its lanes admits a row (search_tc2_kernel.inc).  Each query sets ~ln(rows) records along the scan, so a
// warp of unrelated queries makes ~32 x as many calls as one query would; queries with similar tiles have similar
// score profiles and set their records in the same groups.  Sorting the batch by a 21-bit Morton code of the
// tile's mean colour (7 bits of Y, U, V each) takes the C2 coarse pass from 80-82 to 75.4 ms on the same box -
// the same as sorting by the answer itself (scripts/query_order_experiment.py) - and makes the re-rank's row
// gathers coherent as well.  The order is internal: descriptors are computed straight into sorted positions
// (psyv_queries_kernel reads the tiles in the caller's order - coalesced, no dependent load - and stores row i at
// inv[i]) and the results are scattered back, so callers see their own order.
// Any key gives correct results; the key only decides how fast.
#include "tm_common.cuh"
#include <cub/device/device_radix_sort.cuh>
#include <stdlib.h>

namespace {

constexpr int QK_BITS = 21;

__device__ __forceinline__ uint32_t spread7(uint32_t x) { // bit i -> bit 3 i
    uint32_t o = 0;
#pragma unroll
    for (int i = 0; i < 7; i++) o |= ((x >> i) & 1u) << (3 * i);
    return o;
}

// 16 lanes per tile: one 16-byte load (4 pixels, $00BBGGRR main.pas:566-569) each, integer channel sums
__global__ void __launch_bounds__(256) query_key_kernel(const int32_t *__restrict__ rgb, int64_t nq,
                                                        uint32_t *__restrict__ keys, int *__restrict__ vals) {
    const int64_t q = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 4;
    const int l = threadIdx.x & 15;
    int sr = 0, sg = 0, sb = 0;
    if (q < nq) {
        const uint4 v = __ldg(reinterpret_cast<const uint4 *>(rgb + q * 64) + l);
        const uint32_t px[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int i = 0; i < 4; i++) {
            sr += px[i] & 0xff;
            sg += (px[i] >> 8) & 0xff;
            sb += (px[i] >> 16) & 0xff;
        }
    }
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) {
        sr += __shfl_xor_sync(0xffffffffu, sr, o);
        sg += __shfl_xor_sync(0xffffffffu, sg, o);
        sb += __shfl_xor_sync(0xffffffffu, sb, o);
    }
    if (q < nq && l == 0) {
        // mean colour in 0..255, BT.709 luma and scaled colour differences (the constants of RGBToYUV,
        // main.pas:2656-2679, in float: this is an ordering heuristic, not descriptor arithmetic)
        const float r = sr * (1.f / 64.f), g = sg * (1.f / 64.f), b = sb * (1.f / 64.f);
        const float y = 0.2126f * r + 0.7152f * g + 0.0722f * b;
        const float u = (b - y) * (0.5f / 0.9278f) + 128.f, v = (r - y) * (0.5f / 0.7874f) + 128.f;
        const uint32_t qy = (uint32_t)min(127, max(0, (int)(y * 0.5f)));
        const uint32_t qu = (uint32_t)min(127, max(0, (int)(u * 0.5f)));
        const uint32_t qv = (uint32_t)min(127, max(0, (int)(v * 0.5f)));
        keys[q] = (spread7(qy) << 2) | (spread7(qu) << 1) | spread7(qv);
        vals[q] = (int)q;
    }
}

// inverse of the sorted order: query i goes to sorted position inv[perm[p]] = p
__global__ void invert_order_kernel(const int *__restrict__ perm, int64_t n, int *__restrict__ inv) {
    const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p < n) inv[perm[p]] = (int)p;
}

__global__ void scatter_query_results_kernel(const int *__restrict__ perm, int64_t n, const int *__restrict__ idx_s,
                                             const float *__restrict__ err_s, int *__restrict__ idx,
                                             float *__restrict__ err) {
    const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    const int q = perm[p];
    idx[q] = idx_s[p];
    err[q] = err_s[p];
}

} // namespace

size_t tm_query_order_temp_bytes(int64_t nq) {
    size_t bytes = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, bytes, (const uint32_t *)nullptr, (uint32_t *)nullptr, (const int *)nullptr,
                                    (int *)nullptr, (int)nq, 0, QK_BITS);
    return bytes;
}

// d_perm[p] = index (within the batch) of the query that takes sorted position p; d_inv = its inverse
int tm_launch_query_order(const int32_t *d_rgb, int64_t nq, uint32_t *d_keys /* [2][nq] */, int *d_vals /* [nq] */,
                          int *d_perm, int *d_inv, void *d_tmp, size_t tmp_bytes, cudaStream_t s) {
    if (nq == 0) return 0;
    if (nq > 0x7fffffff) { tm_set_error("query batch too large to order"); return TM_EINVAL; }
    const int64_t threads = nq * 16;
    query_key_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, s>>>(d_rgb, nq, d_keys, d_vals);
    TM_CUDA(cudaGetLastError());
    static const int bits = getenv("TM_QORDER_BITS") ? atoi(getenv("TM_QORDER_BITS")) : QK_BITS; // tuning knob: top bits only
    const int lo = bits >= 1 && bits <= QK_BITS ? QK_BITS - bits : 0;
    TM_CUDA(cub::DeviceRadixSort::SortPairs(d_tmp, tmp_bytes, d_keys, d_keys + nq, d_vals, d_perm, (int)nq, lo, QK_BITS, s));
    invert_order_kernel<<<(unsigned)((nq + 255) / 256), 256, 0, s>>>(d_perm, nq, d_inv);
    tm_count_launch(2); // key + inversion kernels; CUB's sort kernels are library code and are not counted
    return tm_check(cudaGetLastError(), "invert_order_kernel");
}

// deferred reject list of a batch searched in coherence order: sorted position -> position in the caller's order
__global__ void translate_list_kernel(int *__restrict__ list, const int *__restrict__ count, int64_t cap,
                                      const int *__restrict__ perm) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < cap && i < *count) list[i] = perm[list[i]];
}

int tm_launch_translate_list(int *d_list, const int *d_count, int64_t cap, const int *d_perm, cudaStream_t s) {
    if (cap <= 0) return 0;
    translate_list_kernel<<<(unsigned)((cap + 255) / 256), 256, 0, s>>>(d_list, d_count, cap, d_perm);
    tm_count_launch();
    return tm_check(cudaGetLastError(), "translate_list_kernel");
}

int tm_launch_scatter_query_results(const int *d_perm, int64_t nq, const int *d_idx_s, const float *d_err_s, int *d_idx,
                                    float *d_err, cudaStream_t s) {
    if (nq == 0) return 0;
    scatter_query_results_kernel<<<(unsigned)((nq + 255) / 256), 256, 0, s>>>(d_perm, nq, d_idx_s, d_err_s, d_idx, d_err);
    tm_count_launch();
    return tm_check(cudaGetLastError(), "scatter_query_results_kernel");
}
