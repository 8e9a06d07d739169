// kmodes.cu — KModes cluster-assignment pass on the GPU (sm_100a).
//
// Replaces TKModes.DoGMMD / GetMinMatchingDissim (reference kmodes.pas:825-835,
// 316-453; loops 866-878 and 984-997) and UpdateMinDistance (kmodes.pas:455-596).
// Rows are 80 bytes (64 palette indices + 16 zone flags, main.pas:4167-4183).
// Integer work, bit-exact.  Two dissimilarity semantics:
//   TM_KM_ASM      popcnt(mismatch over 80 bytes) << 11
//                  + |d0| + 256|d1| + |d8| + 256|d9|   (pabsb of a wrapping psubb)
//                  + SAD(bytes 16..79)                  kmodes.pas:357-412
//   TM_KM_GENERIC  mismatches << 11 + SAD(bytes 0..79)  kmodes.pas:239-250
// Arg-min keeps the LAST index among equal dissimilarities (cmp/ja kmodes.pas:414-415).
//
// Each thread owns one point (20 words in registers); centroids are staged in
// shared memory in tiles and read as warp-wide broadcasts.
#include "tm_common.cuh"

namespace {

constexpr int KM_WORDS = 20;
constexpr int KM_TILE = 512; // centroids per shared-memory tile (40 KB)

__device__ __forceinline__ uint32_t nonzero_bytes(uint32_t d) {
    // 0x80 in every byte of d that is non-zero
    return (((d & 0x7f7f7f7fu) + 0x7f7f7f7fu) | d) & 0x80808080u;
}

// signed-byte abs of a wrapping byte difference (pabsb(psubb(a,b))): 0..128
__device__ __forceinline__ uint32_t pabsb_psubb(uint32_t a, uint32_t b) {
    int d = (int)(int8_t)(uint8_t)(a - b);
    return (uint32_t)(d < 0 ? -d : d);
}

template <int SEM>
__device__ __forceinline__ uint32_t dissim_words(const uint32_t (&p)[KM_WORDS], const uint32_t *__restrict__ c) {
    uint32_t mism = 0, sad = 0;
    uint32_t cw[KM_WORDS];
#pragma unroll
    for (int w = 0; w < KM_WORDS; w += 4) {
        uint4 v = *reinterpret_cast<const uint4 *>(c + w);
        cw[w] = v.x; cw[w + 1] = v.y; cw[w + 2] = v.z; cw[w + 3] = v.w;
    }
#pragma unroll
    for (int w = 0; w < KM_WORDS; w++) {
        uint32_t d = __vabsdiffu4(p[w], cw[w]);
        mism += __popc(nonzero_bytes(d));
        if (SEM == TM_KM_GENERIC || w >= 4) sad = __vsadu4(p[w], cw[w]) + sad;
    }
    if (SEM == TM_KM_ASM) {
        // bytes 0,1 (word 0) and 8,9 (word 2); bytes 2-7, 10-15 only count as mismatches
        sad += pabsb_psubb(cw[0] & 0xff, p[0] & 0xff) + (pabsb_psubb((cw[0] >> 8) & 0xff, (p[0] >> 8) & 0xff) << 8);
        sad += pabsb_psubb(cw[2] & 0xff, p[2] & 0xff) + (pabsb_psubb((cw[2] >> 8) & 0xff, (p[2] >> 8) & 0xff) << 8);
    }
    return (mism << 11) + sad;
}

template <int SEM>
__global__ void __launch_bounds__(256) kmodes_assign_kernel(const uint32_t *__restrict__ X, int64_t np,
                                                             const uint32_t *__restrict__ C, int64_t k,
                                                             int32_t *__restrict__ clust,
                                                             uint64_t *__restrict__ dis) {
    __shared__ __align__(16) uint32_t cs[KM_TILE * KM_WORDS];
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t p[KM_WORDS];
    if (i < np) {
        const uint4 *src = reinterpret_cast<const uint4 *>(X + i * KM_WORDS);
#pragma unroll
        for (int w = 0; w < 5; w++) {
            uint4 v = __ldg(src + w);
            p[4 * w] = v.x; p[4 * w + 1] = v.y; p[4 * w + 2] = v.z; p[4 * w + 3] = v.w;
        }
    } else {
#pragma unroll
        for (int w = 0; w < KM_WORDS; w++) p[w] = 0;
    }
    uint32_t best = 0xffffffffu;
    int bi = -1;
    for (int64_t c0 = 0; c0 < k; c0 += KM_TILE) {
        int cnt = (int)min((int64_t)KM_TILE, k - c0);
        __syncthreads();
        const uint4 *src = reinterpret_cast<const uint4 *>(C + c0 * KM_WORDS);
        for (int e = threadIdx.x; e < cnt * 5; e += blockDim.x) reinterpret_cast<uint4 *>(cs)[e] = __ldg(src + e);
        __syncthreads();
#pragma unroll 2
        for (int c = 0; c < cnt; c++) {
            uint32_t d = dissim_words<SEM>(p, cs + c * KM_WORDS);
            if (d <= best) { best = d; bi = (int)(c0 + c); } // last index wins ties
        }
    }
    if (i < np) {
        clust[i] = bi;
        if (dis) dis[i] = (k > 0) ? (uint64_t)best : ~(uint64_t)0;
    }
}

template <int SEM>
__global__ void kmodes_umd_kernel(const uint32_t *__restrict__ X, int64_t np, const uint32_t *__restrict__ center,
                                  uint64_t *__restrict__ mind) {
    __shared__ __align__(16) uint32_t cs[KM_WORDS];
    if (threadIdx.x < KM_WORDS) cs[threadIdx.x] = center[threadIdx.x];
    __syncthreads();
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= np) return;
    uint32_t p[KM_WORDS];
    const uint4 *src = reinterpret_cast<const uint4 *>(X + i * KM_WORDS);
#pragma unroll
    for (int w = 0; w < 5; w++) {
        uint4 v = __ldg(src + w);
        p[4 * w] = v.x; p[4 * w + 1] = v.y; p[4 * w + 2] = v.z; p[4 * w + 3] = v.w;
    }
    // asm order is dissim(item=center, list=X[i]): pabsb(psubb(list, item)) -> symmetric under abs
    uint64_t d = dissim_words<SEM>(p, cs);
    uint64_t m = mind[i];
    if (d < m) mind[i] = d; // strict: cmovb kmodes.pas:555-558
}

} // namespace

int tm_launch_kmodes_assign(const uint8_t *d_X, int64_t np, const uint8_t *d_c, int64_t k, int sem,
                            int32_t *d_clust, uint64_t *d_dis, cudaStream_t s) {
    if (sem != TM_KM_ASM && sem != TM_KM_GENERIC) { tm_set_error("unknown KModes semantics %d", sem); return TM_EINVAL; }
    if (np == 0) return 0;
    if (((uintptr_t)d_X & 15) || ((uintptr_t)d_c & 15)) { tm_set_error("KModes rows must be 16-byte aligned"); return TM_EINVAL; }
    unsigned grid = (unsigned)((np + 255) / 256);
    if (sem == TM_KM_ASM)
        kmodes_assign_kernel<TM_KM_ASM><<<grid, 256, 0, s>>>((const uint32_t *)d_X, np, (const uint32_t *)d_c, k, d_clust, d_dis);
    else
        kmodes_assign_kernel<TM_KM_GENERIC><<<grid, 256, 0, s>>>((const uint32_t *)d_X, np, (const uint32_t *)d_c, k, d_clust, d_dis);
    tm_count_launch();
    return tm_check(cudaGetLastError(), "kmodes_assign_kernel");
}

int tm_launch_kmodes_umd(const uint8_t *d_X, int64_t np, const uint8_t *d_center, int sem, uint64_t *d_mind,
                         cudaStream_t s) {
    if (sem != TM_KM_ASM && sem != TM_KM_GENERIC) { tm_set_error("unknown KModes semantics %d", sem); return TM_EINVAL; }
    if (np == 0) return 0;
    unsigned grid = (unsigned)((np + 255) / 256);
    if (sem == TM_KM_ASM)
        kmodes_umd_kernel<TM_KM_ASM><<<grid, 256, 0, s>>>((const uint32_t *)d_X, np, (const uint32_t *)d_center, d_mind);
    else
        kmodes_umd_kernel<TM_KM_GENERIC><<<grid, 256, 0, s>>>((const uint32_t *)d_X, np, (const uint32_t *)d_center, d_mind);
    tm_count_launch();
    return tm_check(cudaGetLastError(), "kmodes_umd_kernel");
}
