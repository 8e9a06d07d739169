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
#include <algorithm>

namespace {

constexpr int KM_WORDS = 20;
constexpr int KM_TILE = 512; // centroids per shared-memory tile (40 KB as byte rows, 24 KB bit-sliced)

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

// ---- bit-sliced fast path ---------------------------------------------------------------------
// With the reference's default palette size (16 colours, main.lfm:452-470) every byte of a row is < 16
// (64 palette indices + 16 zone flags 0/1).  Such a row is also stored bit-sliced: plane b (b = 0..3)
// holds bit b of all 80 values in three words.  The mismatch mask of 32 positions is then
// (a0^c0)|(a1^c1)|(a2^c2)|(a3^c3) = four LOP3, one POPC per 32 positions, i.e. ~17 instructions for the
// whole mismatch count instead of ~110 on bytes.  The mismatch count dominates the dissimilarity
// (<< 11), so the byte-wise SAD term is only evaluated (from the L1/L2-resident byte rows) for the
// few centroids whose mismatch count alone does not already exceed the running best
// (dis >= mismatches << 11).  Results are bit-identical to the byte path, which is still taken by
// blocks that see a byte >= 16 in their points or when any centroid has one.
constexpr int KM_PLANES = 4;
constexpr int KM_BWORDS = 12; // 4 planes x 3 words

// gather bit b of the 4 bytes of w into bits 0..3 (byte i -> bit i)
__device__ __forceinline__ uint32_t gather_bit(uint32_t w, int b) {
    return ((((w >> b) & 0x01010101u) * 0x01020408u) >> 24) & 0xfu;
}
// 20 byte words -> 12 bit-sliced words
__device__ __forceinline__ void bitslice_row(const uint32_t (&p)[KM_WORDS], uint32_t (&out)[KM_BWORDS]) {
#pragma unroll
    for (int b = 0; b < KM_PLANES; b++)
#pragma unroll
        for (int j = 0; j < 3; j++) {
            uint32_t v = 0;
#pragma unroll
            for (int w = 0; w < 8; w++)
                if (8 * j + w < KM_WORDS) v |= gather_bit(p[8 * j + w], b) << (4 * w);
            out[b * 3 + j] = v;
        }
}

// SAD part of the dissimilarity only (the mismatch count comes from the bit-sliced words)
template <int SEM>
__device__ __forceinline__ uint32_t sad_words(const uint32_t (&p)[KM_WORDS], const uint32_t *__restrict__ c) {
    uint32_t sad = 0;
    uint32_t cw[KM_WORDS];
#pragma unroll
    for (int w = 0; w < KM_WORDS; w += 4) {
        uint4 v = __ldg(reinterpret_cast<const uint4 *>(c + w));
        cw[w] = v.x; cw[w + 1] = v.y; cw[w + 2] = v.z; cw[w + 3] = v.w;
    }
#pragma unroll
    for (int w = (SEM == TM_KM_GENERIC ? 0 : 4); w < KM_WORDS; w++) sad = __vsadu4(p[w], cw[w]) + sad;
    if (SEM == TM_KM_ASM) {
        sad += pabsb_psubb(cw[0] & 0xff, p[0] & 0xff) + (pabsb_psubb((cw[0] >> 8) & 0xff, (p[0] >> 8) & 0xff) << 8);
        sad += pabsb_psubb(cw[2] & 0xff, p[2] & 0xff) + (pabsb_psubb((cw[2] >> 8) & 0xff, (p[2] >> 8) & 0xff) << 8);
    }
    return sad;
}

// one thread per centroid: bit-sliced copy + "some byte >= 16" flag
__global__ void kmodes_pack_centroids_kernel(const uint32_t *__restrict__ C, int64_t k, uint32_t *__restrict__ Cb,
                                             int *__restrict__ wide_flag) {
    int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= k) return;
    uint32_t p[KM_WORDS], wide = 0;
    const uint4 *src = reinterpret_cast<const uint4 *>(C + c * KM_WORDS);
#pragma unroll
    for (int w = 0; w < 5; w++) {
        uint4 v = __ldg(src + w);
        p[4 * w] = v.x; p[4 * w + 1] = v.y; p[4 * w + 2] = v.z; p[4 * w + 3] = v.w;
    }
#pragma unroll
    for (int w = 0; w < KM_WORDS; w++) wide |= p[w];
    if (wide & 0xf0f0f0f0u) atomicOr(wide_flag, 1);
    uint32_t o[KM_BWORDS];
    bitslice_row(p, o);
    uint4 *dst = reinterpret_cast<uint4 *>(Cb + c * KM_BWORDS);
    dst[0] = make_uint4(o[0], o[1], o[2], o[3]);
    dst[1] = make_uint4(o[4], o[5], o[6], o[7]);
    dst[2] = make_uint4(o[8], o[9], o[10], o[11]);
}

// KModes handle (tm_kmodes_update_centroids): rows[m][80] replace the centroids idx[m] - byte rows and bit-sliced
// copies in one pass (MovePointCat changes a handful of centroids between two 960-point bins, kmodes.pas:880-909)
__global__ void kmodes_update_centroids_kernel(const uint32_t *__restrict__ rows, const int32_t *__restrict__ idx, int64_t m,
                                               int64_t k, uint32_t *__restrict__ C, uint32_t *__restrict__ Cb,
                                               int *__restrict__ wide_flag, int *__restrict__ bad) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m) return;
    const int64_t c = idx[i];
    if (c < 0 || c >= k) { atomicOr(bad, 1); return; }
    uint32_t p[KM_WORDS], wide = 0;
#pragma unroll
    for (int w = 0; w < KM_WORDS; w++) {
        p[w] = rows[i * KM_WORDS + w];
        C[c * KM_WORDS + w] = p[w];
        wide |= p[w];
    }
    if (wide & 0xf0f0f0f0u) atomicOr(wide_flag, 1);
    uint32_t o[KM_BWORDS];
    bitslice_row(p, o);
#pragma unroll
    for (int w = 0; w < KM_BWORDS; w++) Cb[c * KM_BWORDS + w] = o[w];
}

template <int SEM>
__global__ void __launch_bounds__(256, 3) kmodes_assign_kernel(const uint32_t *__restrict__ X, int64_t np,
                                                             const uint32_t *__restrict__ C, int64_t k,
                                                             const uint32_t *__restrict__ Cb,
                                                             const int *__restrict__ cent_wide,
                                                             int32_t *__restrict__ clust,
                                                             uint64_t *__restrict__ dis, int64_t cchunk,
                                                             int32_t *__restrict__ part_i,
                                                             uint32_t *__restrict__ part_d, const TmKmHostOut bin) {
    // blockIdx.y selects a range of cchunk centroids (few points: the centroids are split over the grid and
    // merged by kmodes_merge_kernel; gridDim.y == 1: the whole set, results written directly)
    const int64_t c_lo = (int64_t)blockIdx.y * cchunk, c_hi = min(k, c_lo + cchunk);
    __shared__ __align__(16) uint32_t cs[KM_TILE * KM_WORDS]; // byte rows, or bit-sliced rows (12 words each)
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
    uint32_t wide = 0;
#pragma unroll
    for (int w = 0; w < KM_WORDS; w++) wide |= p[w];
    const bool fast = !__syncthreads_or((wide & 0xf0f0f0f0u) != 0) && *cent_wide == 0; // block-uniform
    uint32_t best = 0xffffffffu;
    int bi = -1;
    if (bin.prev && i < np) {
        // Bin-sized call with a finer centroid split: every block would start without a bound and evaluate the byte-wise
        // SAD term for each record of its own short range (some lane of a warp does so in nine iterations out of ten).
        // The point's label from the previous pass gives the bound at once - labels rarely change between two
        // KModesIter passes.  Only a bound: bi stays -1 until a centroid of THIS range reaches it (the previous
        // centroid itself does, in its own range), so results do not depend on it.
        const int pi = bin.prev[i];
        if (pi >= 0 && pi < k) best = dissim_words<SEM>(p, C + (int64_t)pi * KM_WORDS);
    }
    if (fast) {
        uint32_t pb[KM_BWORDS];
        bitslice_row(p, pb);
        for (int64_t c0 = c_lo; c0 < c_hi; c0 += KM_TILE) {
            int cnt = (int)min((int64_t)KM_TILE, c_hi - c0);
            __syncthreads();
            const uint4 *src = reinterpret_cast<const uint4 *>(Cb + c0 * KM_BWORDS);
            for (int e = threadIdx.x; e < cnt * 3; e += blockDim.x) reinterpret_cast<uint4 *>(cs)[e] = __ldg(src + e);
            __syncthreads();
#pragma unroll 4
            for (int c = 0; c < cnt; c++) {
                const uint4 *cn = reinterpret_cast<const uint4 *>(cs + c * KM_BWORDS);
                const uint4 v0 = cn[0], v1 = cn[1], v2 = cn[2];
                const uint32_t cw[KM_BWORDS] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w, v2.x, v2.y, v2.z, v2.w};
                uint32_t mism = 0;
#pragma unroll
                for (int j = 0; j < 3; j++) {
                    uint32_t m = pb[j] ^ cw[j];
                    m |= pb[3 + j] ^ cw[3 + j];
                    m |= pb[6 + j] ^ cw[6 + j];
                    m |= pb[9 + j] ^ cw[9 + j];
                    mism += __popc(m);
                }
                const uint32_t key = mism << 11;
                if (key <= best) { // otherwise dis >= key > best: cannot win, cannot tie
                    const uint32_t d = key + sad_words<SEM>(p, C + (c0 + c) * KM_WORDS);
                    if (d <= best) { best = d; bi = (int)(c0 + c); } // last index wins ties
                }
            }
        }
    } else {
        for (int64_t c0 = c_lo; c0 < c_hi; c0 += KM_TILE) {
            int cnt = (int)min((int64_t)KM_TILE, c_hi - c0);
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
    }
    if (bin.keys) {
        // KModes handle, bin-sized call (tm_kmodes_assign_range): the centroid ranges meet in one 64-bit atomicMin per
        // point - (dissimilarity, ~index): the smallest dissimilarity, the LAST index among equals - and the last
        // block of the grid to finish writes labels and dissimilarities straight into mapped host memory, re-arms
        // the keys for the next call and publishes the call's sequence number in a host flag the caller polls: no
        // merge launch, no copy command, no stream synchronisation per bin.
        if (i < np && bi >= 0)
            atomicMin(bin.keys + i, ((unsigned long long)best << 32) | (unsigned long long)(0xffffffffu - (uint32_t)bi));
        __shared__ bool last;
        __threadfence();
        __syncthreads();
        if (threadIdx.x == 0) last = atomicAdd(bin.d_done, 1u) == gridDim.x * gridDim.y - 1;
        __syncthreads();
        if (!last) return;
        __threadfence();
        for (int64_t j = threadIdx.x; j < np; j += blockDim.x) {
            const unsigned long long key = __ldcg(bin.keys + j);
            bin.keys[j] = ~0ull;
            const bool have = key != ~0ull;
            const int32_t lab = have ? (int32_t)(0xffffffffu - (uint32_t)key) : -1;
            bin.m_clust[j] = lab;
            if (bin.prev) bin.prev[j] = lab;
            if (bin.m_dis) bin.m_dis[j] = have ? (uint64_t)(key >> 32) : ~(uint64_t)0;
        }
        __threadfence_system();
        __syncthreads();
        if (threadIdx.x == 0) {
            *bin.d_done = 0;
            *reinterpret_cast<volatile int *>(bin.m_flag) = bin.seq;
        }
        return;
    }
    if (i < np) {
        if (gridDim.y > 1) {
            part_i[(int64_t)blockIdx.y * np + i] = bi;
            part_d[(int64_t)blockIdx.y * np + i] = best;
        } else {
            clust[i] = bi;
            if (dis) dis[i] = (k > 0) ? (uint64_t)best : ~(uint64_t)0;
        }
    }
}

// centroid ranges in index order; "<=" keeps the later range on ties, i.e. the last index wins (kmodes.pas:414-415)
__global__ void kmodes_merge_kernel(const int32_t *__restrict__ part_i, const uint32_t *__restrict__ part_d, int64_t np,
                                    int nparts, int32_t *__restrict__ clust, uint64_t *__restrict__ dis) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= np) return;
    uint32_t best = 0xffffffffu;
    int bi = -1;
    for (int p = 0; p < nparts; p++) {
        const int ci = part_i[(int64_t)p * np + i];
        const uint32_t d = part_d[(int64_t)p * np + i];
        if (ci >= 0 && d <= best) { best = d; bi = ci; }
    }
    clust[i] = bi;
    if (dis) dis[i] = (uint64_t)best;
}

// changed centroid rows carried in the kernel's parameter space (MovePointCat touches a handful per bin): no staging copy
constexpr int KM_INLINE_ROWS = 24;
struct KmInlineRows {
    int32_t idx[KM_INLINE_ROWS];
    uint32_t rows[KM_INLINE_ROWS][KM_WORDS];
};
__global__ void kmodes_update_inline_kernel(const __grid_constant__ KmInlineRows u, int m, int64_t k, uint32_t *__restrict__ C,
                                            uint32_t *__restrict__ Cb, int *__restrict__ wide_flag) {
    const int i = threadIdx.x;
    if (i >= m) return;
    // duplicates: the LAST row for a centroid wins, like sequential stores - earlier ones step aside
    for (int j = i + 1; j < m; j++)
        if (u.idx[j] == u.idx[i]) return;
    const int64_t c = u.idx[i];
    if (c < 0 || c >= k) return; // validated on the host
    uint32_t p[KM_WORDS], wide = 0;
#pragma unroll
    for (int w = 0; w < KM_WORDS; w++) {
        p[w] = u.rows[i][w];
        C[c * KM_WORDS + w] = p[w];
        wide |= p[w];
    }
    if (wide & 0xf0f0f0f0u) atomicOr(wide_flag, 1);
    uint32_t o[KM_BWORDS];
    bitslice_row(p, o);
#pragma unroll
    for (int w = 0; w < KM_BWORDS; w++) Cb[c * KM_BWORDS + w] = o[w];
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

// ---- farthest-first initialisation (InitFarthestFirst kmodes.pas:698-776) ------------------------
// One launch per centroid, no host round trip in between: pass c reads the point chosen by pass c-1
// from argkey[c] (device memory), records it (chosen[c], centroids[c] = X[cur], used[cur] = 1), folds
// dissim(X[cur], X[i]) into mindist[i] for EVERY point (the asm's `used` test never skips,
// kmodes.pas:567) with a strict < (555-558), and reduces the next farthest point: the LAST index among
// the unused points with the largest mindist (`>=`, 765-770) = the maximum of (mindist << 32 | i).
// Keys are stored + 1 so that 0 means "no unused point left".  X (80 B / point) stays L2-resident.
template <int SEM>
__global__ void __launch_bounds__(256) kmodes_ff_pass_kernel(const uint32_t *__restrict__ X, int64_t np, int c,
                                                              unsigned long long *__restrict__ argkey,
                                                              uint32_t *__restrict__ mind, uint8_t *__restrict__ used,
                                                              uint32_t *__restrict__ cent_out, int64_t *__restrict__ chosen,
                                                              int *__restrict__ err) {
    __shared__ __align__(16) uint32_t cs[KM_WORDS];
    __shared__ unsigned long long wbest[8];
    const unsigned long long ck = argkey[c];
    if (ck == 0) { // more centroids requested than points: the reference would index X[-1]
        if (blockIdx.x == 0 && threadIdx.x == 0) atomicOr(err, 1);
        return;
    }
    const uint32_t cur = (uint32_t)((ck - 1) & 0xffffffffull);
    if (threadIdx.x < KM_WORDS) {
        const uint32_t w = X[(int64_t)cur * KM_WORDS + threadIdx.x];
        cs[threadIdx.x] = w;
        if (blockIdx.x == 0) cent_out[(int64_t)c * KM_WORDS + threadIdx.x] = w;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) { chosen[c] = cur; used[cur] = 1; }
    __syncthreads();
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    unsigned long long key = 0;
    if (i < np) {
        uint32_t p[KM_WORDS];
        const uint4 *src = reinterpret_cast<const uint4 *>(X + i * KM_WORDS);
#pragma unroll
        for (int w = 0; w < 5; w++) {
            uint4 v = __ldg(src + w);
            p[4 * w] = v.x; p[4 * w + 1] = v.y; p[4 * w + 2] = v.z; p[4 * w + 3] = v.w;
        }
        const uint32_t d = dissim_words<SEM>(p, cs);
        uint32_t m = mind[i];
        if (d < m) { m = d; mind[i] = d; }
        if (i != cur && !used[i]) key = (((unsigned long long)m << 32) | (unsigned long long)i) + 1;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const unsigned long long other = __shfl_xor_sync(0xffffffffu, key, o);
        key = other > key ? other : key;
    }
    if ((threadIdx.x & 31) == 0) wbest[threadIdx.x >> 5] = key;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long b = wbest[0];
#pragma unroll
        for (int w = 1; w < 8; w++) b = wbest[w] > b ? wbest[w] : b;
        if (b) atomicMax(argkey + c + 1, b);
    }
}

// ---- cluster representatives (DoKModes main.pas:4231-4253) ----------------------------------------
// rep[j] = GetMinMatchingDissim(members of cluster j in point order, centroid j): minimum dissimilarity,
// LAST member among equals (kmodes.pas:414-415) = the minimum of (dis << 32 | ~i).
template <int SEM>
__global__ void __launch_bounds__(256) kmodes_rep_kernel(const uint32_t *__restrict__ X, int64_t np,
                                                          const int32_t *__restrict__ labels,
                                                          const uint32_t *__restrict__ C, int64_t k,
                                                          unsigned long long *__restrict__ best, int32_t *__restrict__ count,
                                                          int *__restrict__ err) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= np) return;
    const int32_t l = labels[i];
    if (l < 0 || l >= k) { atomicOr(err, 1); return; }
    uint32_t p[KM_WORDS];
    const uint4 *src = reinterpret_cast<const uint4 *>(X + i * KM_WORDS);
#pragma unroll
    for (int w = 0; w < 5; w++) {
        uint4 v = __ldg(src + w);
        p[4 * w] = v.x; p[4 * w + 1] = v.y; p[4 * w + 2] = v.z; p[4 * w + 3] = v.w;
    }
    const uint32_t d = dissim_words<SEM>(p, C + (int64_t)l * KM_WORDS);
    atomicMin(best + l, ((unsigned long long)d << 32) | (unsigned long long)(0xffffffffu - (uint32_t)i));
    atomicAdd(count + l, 1);
}

__global__ void kmodes_rep_finish_kernel(const unsigned long long *__restrict__ best, int64_t k, int32_t *__restrict__ rep) {
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= k) return;
    const unsigned long long b = best[j];
    rep[j] = b == ~0ull ? -1 : (int32_t)(0xffffffffu - (uint32_t)(b & 0xffffffffull));
}

// ---- KModes rows (WriteTileDatasetLine main.pas:4167-4183 + GetTilePalZoneThres 4142-4165) ---------
// row = 64 palette indices + 16 zone flags (flag z = 1 iff more than pal_size div 16 pixels fall into
// zone b*16 div pal_size == z); signi = min over zones of (64 - count) (PalSigni, the bin key of
// ReloadPreviousTiling main.pas:4393-4404).  One thread per tile, 16-byte loads / stores.
__global__ void kmodes_rows_kernel(const uint8_t *__restrict__ palpix, int64_t ntiles, int pal_size,
                                   uint8_t *__restrict__ rows, int32_t *__restrict__ signi) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= ntiles) return;
    const uint4 *src = reinterpret_cast<const uint4 *>(palpix + t * 64);
    uint4 *dst = reinterpret_cast<uint4 *>(rows + t * TM_KM_FEATURES);
    uint32_t acc[16];
#pragma unroll
    for (int z = 0; z < 16; z++) acc[z] = 0;
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const uint4 v = __ldg(src + i);
        dst[i] = v;
        const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int j = 0; j < 4; j++)
#pragma unroll
            for (int b = 0; b < 4; b++) {
                const int zone = (int)((w[j] >> (8 * b)) & 0xff) * 16 / pal_size;
#pragma unroll
                for (int z = 0; z < 16; z++) acc[z] += (zone == z); // registers only: no dynamic indexing
            }
    }
    int res = 64;
    uint32_t f[4] = {0, 0, 0, 0};
#pragma unroll
    for (int z = 0; z < 16; z++) {
        res = min(res, 64 - (int)acc[z]);
        f[z >> 2] |= (uint32_t)(acc[z] > (uint32_t)(pal_size / 16)) << (8 * (z & 3));
    }
    dst[4] = make_uint4(f[0], f[1], f[2], f[3]);
    if (signi) signi[t] = res;
}

} // namespace

// bit-sliced centroids, the "some byte >= 16" flag and the partial arg-min buffers of a centroid-split launch
// (TmKmScratch, tm_common.cuh): one set per device for the one-shot entry points, one per KModes handle
#define g_km (tm_dev()->km)

static int km_ensure_centroids(TmKmScratch &c, int64_t k, cudaStream_t s) {
    if ((size_t)k > c.cap || !c.flag) {
        cudaStreamSynchronize(s);
        cudaFree(c.cb);
        c.cb = nullptr; c.cap = 0;
        TM_CUDA(cudaMalloc(&c.cb, sizeof(uint32_t) * KM_BWORDS * (size_t)(k > 0 ? k : 1)));
        if (!c.flag) TM_CUDA(cudaMalloc(&c.flag, sizeof(int)));
        c.cap = (size_t)(k > 0 ? k : 1);
    }
    return 0;
}

// bit-sliced copy of all k centroids + wide flag
int tm_kmodes_pack_centroids(TmKmScratch &c, const uint8_t *d_c, int64_t k, cudaStream_t s) {
    int r = km_ensure_centroids(c, k, s);
    if (r) return r;
    TM_CUDA(cudaMemsetAsync(c.flag, 0, sizeof(int), s));
    if (k > 0) {
        kmodes_pack_centroids_kernel<<<(unsigned)((k + 127) / 128), 128, 0, s>>>((const uint32_t *)d_c, k, c.cb, c.flag);
        tm_count_launch();
    }
    return tm_check(cudaGetLastError(), "kmodes_pack_centroids_kernel");
}

int tm_kmodes_update_packed(TmKmScratch &c, uint8_t *d_c, int64_t k, const uint8_t *d_rows, const int32_t *d_idx, int64_t m,
                            int *d_bad, cudaStream_t s) {
    if (m == 0) return 0;
    kmodes_update_centroids_kernel<<<(unsigned)((m + 127) / 128), 128, 0, s>>>((const uint32_t *)d_rows, d_idx, m, k, (uint32_t *)d_c,
                                                                              c.cb, c.flag, d_bad);
    tm_count_launch();
    return tm_check(cudaGetLastError(), "kmodes_update_centroids_kernel");
}

int tm_kmodes_inline_rows() { return KM_INLINE_ROWS; }
// m <= KM_INLINE_ROWS changed rows from HOST memory through the kernel's parameters
int tm_kmodes_update_inline(TmKmScratch &c, uint8_t *d_c, int64_t k, const uint8_t *h_rows, const int32_t *h_idx, int m,
                            cudaStream_t s) {
    if (m == 0) return 0;
    KmInlineRows u;
    memcpy(u.idx, h_idx, sizeof(int32_t) * (size_t)m);
    memcpy(u.rows, h_rows, (size_t)m * 80);
    kmodes_update_inline_kernel<<<1, 32, 0, s>>>(u, m, k, (uint32_t *)d_c, c.cb, c.flag);
    tm_count_launch();
    return tm_check(cudaGetLastError(), "kmodes_update_inline_kernel");
}

// assignment of np points against centroids already packed in c
int tm_kmodes_assign_packed(TmKmScratch &c, const uint8_t *d_X, int64_t np, const uint8_t *d_c, int64_t k, int sem,
                            int32_t *d_clust, uint64_t *d_dis, cudaStream_t s, const TmKmHostOut *ho, bool *ho_used) {
    if (ho_used) *ho_used = false;
    if (np == 0) return 0;
    if (ho && ho->keys && np <= ho->keys_cap && k >= 1 && (np + 255) / 256 < 2 * tm_num_sms()) {
        // Bin-sized call through the handle.  The whole-set split below cuts the centroids in multiples of the 512-row
        // shared-memory tile: a 960-point bin against 16 384 centroids is then 4 x 32 = 128 blocks of 8 warps, one per
        // SM, latency-bound (96 us under ncu).  Here the ranges are as fine as one wave of 3 blocks per SM allows
        // (16-centroid granularity) and meet in an atomicMin per point instead of a partial-result array.
        const unsigned gxb = (unsigned)((np + 255) / 256);
        int64_t want = (3 * (int64_t)tm_num_sms() + gxb - 1) / gxb; // ranges per row of blocks
        if (want < 1) want = 1;
        int64_t chunk = ((k + want - 1) / want + 15) & ~(int64_t)15;
        if (chunk < 16) chunk = 16;
        const unsigned partsb = (unsigned)((k + chunk - 1) / chunk);
        const dim3 gridb(gxb, partsb);
        if (sem == TM_KM_ASM)
            kmodes_assign_kernel<TM_KM_ASM><<<gridb, 256, 0, s>>>((const uint32_t *)d_X, np, (const uint32_t *)d_c, k, c.cb, c.flag, nullptr, nullptr, chunk, nullptr, nullptr, *ho);
        else
            kmodes_assign_kernel<TM_KM_GENERIC><<<gridb, 256, 0, s>>>((const uint32_t *)d_X, np, (const uint32_t *)d_c, k, c.cb, c.flag, nullptr, nullptr, chunk, nullptr, nullptr, *ho);
        tm_count_launch();
        if (ho_used) *ho_used = true;
        return tm_check(cudaGetLastError(), "kmodes_assign_kernel(bin)");
    }
    const unsigned gx = (unsigned)((np + 255) / 256);
    // Few points (KModesIter's bins of 960, kmodes.pas:847): thread-per-point alone would occupy a handful of SMs,
    // so the centroids are split over gridDim.y in multiples of the shared-memory tile and merged afterwards.
    int parts = 1;
    const int64_t ctiles = (k + KM_TILE - 1) / KM_TILE;
    if ((int64_t)gx < 2 * tm_num_sms() && ctiles > 1) {
        parts = (int)std::min<int64_t>(ctiles, (4 * tm_num_sms() + gx - 1) / gx);
        if (parts > 256) parts = 256;
    }
    const int64_t cchunk = parts > 1 ? ((ctiles + parts - 1) / parts) * KM_TILE : (k > 0 ? k : 1);
    if (parts > 1) {
        parts = (int)((k + cchunk - 1) / cchunk);
        const size_t need = (size_t)parts * (size_t)np;
        if (need > c.cap_part) {
            cudaStreamSynchronize(s);
            cudaFree(c.part_i); cudaFree(c.part_d);
            c.part_i = nullptr; c.part_d = nullptr; c.cap_part = 0;
            TM_CUDA(cudaMalloc(&c.part_i, sizeof(int32_t) * need));
            TM_CUDA(cudaMalloc(&c.part_d, sizeof(uint32_t) * need));
            c.cap_part = need;
        }
    }
    const dim3 grid(gx, (unsigned)parts);
    const TmKmHostOut none{};
    if (sem == TM_KM_ASM)
        kmodes_assign_kernel<TM_KM_ASM><<<grid, 256, 0, s>>>((const uint32_t *)d_X, np, (const uint32_t *)d_c, k, c.cb, c.flag, d_clust, d_dis, cchunk, c.part_i, c.part_d, none);
    else
        kmodes_assign_kernel<TM_KM_GENERIC><<<grid, 256, 0, s>>>((const uint32_t *)d_X, np, (const uint32_t *)d_c, k, c.cb, c.flag, d_clust, d_dis, cchunk, c.part_i, c.part_d, none);
    if (parts > 1) {
        kmodes_merge_kernel<<<gx, 256, 0, s>>>(c.part_i, c.part_d, np, parts, d_clust, d_dis);
        tm_count_launch();
    }
    tm_count_launch();
    return tm_check(cudaGetLastError(), "kmodes_assign_kernel");
}

void tm_kmodes_scratch_free(TmKmScratch &c) {
    cudaFree(c.cb); cudaFree(c.flag); cudaFree(c.part_i); cudaFree(c.part_d);
    c = TmKmScratch();
}

TmKmScratch &g_km_scratch() { return g_km; }

int tm_launch_kmodes_assign(const uint8_t *d_X, int64_t np, const uint8_t *d_c, int64_t k, int sem,
                            int32_t *d_clust, uint64_t *d_dis, cudaStream_t s) {
    if (sem != TM_KM_ASM && sem != TM_KM_GENERIC) { tm_set_error("unknown KModes semantics %d", sem); return TM_EINVAL; }
    if (np == 0) return 0;
    if (((uintptr_t)d_X & 15) || ((uintptr_t)d_c & 15)) { tm_set_error("KModes rows must be 16-byte aligned"); return TM_EINVAL; }
    int r = tm_kmodes_pack_centroids(g_km, d_c, k, s);
    if (r) return r;
    return tm_kmodes_assign_packed(g_km, d_X, np, d_c, k, sem, d_clust, d_dis, s);
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

int tm_launch_kmodes_farthest_first(const uint8_t *d_X, int64_t np, int64_t k, int64_t init_point, int sem,
                                    uint8_t *d_centroids, int64_t *d_chosen, unsigned long long *d_argkey,
                                    uint32_t *d_mind, uint8_t *d_used, int *d_err, cudaStream_t s) {
    if (sem != TM_KM_ASM && sem != TM_KM_GENERIC) { tm_set_error("unknown KModes semantics %d", sem); return TM_EINVAL; }
    if (k == 0) return 0;
    TM_CUDA(cudaMemsetAsync(d_argkey, 0, sizeof(unsigned long long) * (size_t)(k + 1), s));
    TM_CUDA(cudaMemsetAsync(d_mind, 0xff, sizeof(uint32_t) * (size_t)np, s));
    TM_CUDA(cudaMemsetAsync(d_used, 0, (size_t)np, s));
    TM_CUDA(cudaMemsetAsync(d_err, 0, sizeof(int), s));
    const unsigned long long k0 = (unsigned long long)init_point + 1; // (mindist 0, index) + 1
    TM_CUDA(cudaMemcpyAsync(d_argkey, &k0, sizeof k0, cudaMemcpyHostToDevice, s));
    TM_CUDA(cudaStreamSynchronize(s)); // k0 is a stack variable
    const unsigned grid = (unsigned)((np + 255) / 256);
    for (int64_t c = 0; c < k; c++) {
        if (sem == TM_KM_ASM)
            kmodes_ff_pass_kernel<TM_KM_ASM><<<grid, 256, 0, s>>>((const uint32_t *)d_X, np, (int)c, d_argkey, d_mind, d_used,
                                                                 (uint32_t *)d_centroids, d_chosen, d_err);
        else
            kmodes_ff_pass_kernel<TM_KM_GENERIC><<<grid, 256, 0, s>>>((const uint32_t *)d_X, np, (int)c, d_argkey, d_mind, d_used,
                                                                     (uint32_t *)d_centroids, d_chosen, d_err);
    }
    tm_count_launch((int)k);
    return tm_check(cudaGetLastError(), "kmodes_ff_pass_kernel");
}

int tm_launch_kmodes_representatives(const uint8_t *d_X, int64_t np, const int32_t *d_labels, const uint8_t *d_c, int64_t k,
                                     int sem, unsigned long long *d_best, int32_t *d_rep, int32_t *d_count, int *d_err,
                                     cudaStream_t s) {
    if (sem != TM_KM_ASM && sem != TM_KM_GENERIC) { tm_set_error("unknown KModes semantics %d", sem); return TM_EINVAL; }
    if (k == 0) return 0;
    TM_CUDA(cudaMemsetAsync(d_best, 0xff, sizeof(unsigned long long) * (size_t)k, s));
    TM_CUDA(cudaMemsetAsync(d_count, 0, sizeof(int32_t) * (size_t)k, s));
    TM_CUDA(cudaMemsetAsync(d_err, 0, sizeof(int), s));
    if (np) {
        const unsigned grid = (unsigned)((np + 255) / 256);
        if (sem == TM_KM_ASM)
            kmodes_rep_kernel<TM_KM_ASM><<<grid, 256, 0, s>>>((const uint32_t *)d_X, np, d_labels, (const uint32_t *)d_c, k, d_best, d_count, d_err);
        else
            kmodes_rep_kernel<TM_KM_GENERIC><<<grid, 256, 0, s>>>((const uint32_t *)d_X, np, d_labels, (const uint32_t *)d_c, k, d_best, d_count, d_err);
        tm_count_launch();
    }
    kmodes_rep_finish_kernel<<<(unsigned)((k + 255) / 256), 256, 0, s>>>(d_best, k, d_rep);
    tm_count_launch();
    return tm_check(cudaGetLastError(), "kmodes_rep_kernel");
}

int tm_launch_kmodes_rows(const uint8_t *d_palpix, int64_t ntiles, int pal_size, uint8_t *d_rows, int32_t *d_signi,
                          cudaStream_t s) {
    if (pal_size < 1 || pal_size > 256) { tm_set_error("pal_size must be in 1..256"); return TM_EINVAL; }
    if (ntiles == 0) return 0;
    kmodes_rows_kernel<<<(unsigned)((ntiles + 127) / 128), 128, 0, s>>>(d_palpix, ntiles, pal_size, d_rows, d_signi);
    tm_count_launch();
    return tm_check(cudaGetLastError(), "kmodes_rows_kernel");
}

// ---- allocation-free KModes assignment (tm_kmodes_assign_noalloc) ---------------------------------------------------
// No counterpart in the reference: added for callers that cannot call
// cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same two kernels as
// tm_kmodes_assign_dev (kmodes_pack_centroids_kernel -> kmodes_assign_kernel, centroids not split over the grid), but
// the scratch - the bit-sliced copy of the centroids and the "some byte >= 16" flag - lives in a file-scope device
// array of fixed size instead of a grow-only cudaMalloc buffer.  Every GPU that loads the library holds the array
// (TM_NOALLOC_MAX_CENTROIDS x 48 B + 256 B = 768.25 KB) until the process exits.  All calls on a device share it and
// it keeps the last call's contents: the flag word is cleared at the start of every call (stream-ordered), the
// bit-sliced rows are rewritten for the k centroids the call uses.  Calls that overlap on the GPU (different streams,
// no ordering between them) race on it.
namespace {
struct __align__(256) KmNoAllocScratch {
    uint32_t cb[(size_t)TM_NOALLOC_MAX_CENTROIDS * KM_BWORDS]; // bit-sliced centroids
    int flag[64];                                              // [0] = wide flag; padded to 256 bytes
};
static __device__ __align__(256) KmNoAllocScratch g_km_noalloc; // in HOST code this name is a stand-in, not a device address

std::mutex g_km_noalloc_mu;
KmNoAllocScratch *g_km_noalloc_ptr[TM_MAX_DEV] = {}; // device address of g_km_noalloc on each device (cudaGetSymbolAddress)

// once per device, with that device current
int km_noalloc_scratch(int dev, KmNoAllocScratch **out) {
    std::lock_guard<std::mutex> lk(g_km_noalloc_mu);
    if (!g_km_noalloc_ptr[dev]) {
        void *p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_km_noalloc);
        if (e != cudaSuccess || !p) {
            tm_set_error("tm_kmodes_assign_noalloc: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
            return TM_ECUDA;
        }
        g_km_noalloc_ptr[dev] = static_cast<KmNoAllocScratch *>(p);
    }
    *out = g_km_noalloc_ptr[dev];
    return 0;
}
} // namespace

extern "C" int tm_kmodes_assign_noalloc(const uint8_t *d_X, int64_t npoints, const uint8_t *d_centroids, int64_t k, int semantics,
                                        int32_t *d_clust, uint64_t *d_dis, void *cuda_stream) {
    if (semantics != TM_KM_ASM && semantics != TM_KM_GENERIC) { tm_set_error("unknown KModes semantics %d", semantics); return TM_EINVAL; }
    if (npoints < 0 || k < 0 || (npoints && (!d_X || !d_clust)) || (k && !d_centroids)) {
        tm_set_error("tm_kmodes_assign_noalloc: bad arguments");
        return TM_EINVAL;
    }
    if (k > TM_NOALLOC_MAX_CENTROIDS) {
        tm_set_error("tm_kmodes_assign_noalloc: %lld centroids exceed the fixed scratch (TM_NOALLOC_MAX_CENTROIDS = %d); "
                     "use tm_kmodes_assign_dev", (long long)k, TM_NOALLOC_MAX_CENTROIDS);
        return TM_EINVAL;
    }
    if (npoints > 0x7fffffffll * 256) { tm_set_error("tm_kmodes_assign_noalloc: too many points"); return TM_EINVAL; }
    if (((uintptr_t)d_X & 15) || ((uintptr_t)d_centroids & 15)) { tm_set_error("KModes rows must be 16-byte aligned"); return TM_EINVAL; }
    if (npoints == 0) return 0;
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= TM_MAX_DEV) {
        cudaGetLastError();
        tm_set_error("tm_kmodes_assign_noalloc: no current CUDA device");
        return TM_ENODEVICE;
    }
    KmNoAllocScratch *sc = nullptr;
    int r = km_noalloc_scratch(dev, &sc);
    if (r) return r;
    cudaStream_t s = static_cast<cudaStream_t>(cuda_stream);
    TM_CUDA(cudaMemsetAsync(sc->flag, 0, sizeof(int), s)); // must start at zero: the pack kernel only ever sets it
    if (k > 0) {
        kmodes_pack_centroids_kernel<<<(unsigned)((k + 127) / 128), 128, 0, s>>>((const uint32_t *)d_centroids, k, sc->cb, sc->flag);
        tm_count_launch();
    }
    const dim3 grid((unsigned)((npoints + 255) / 256), 1);
    const int64_t cchunk = k > 0 ? k : 1;
    if (semantics == TM_KM_ASM)
        kmodes_assign_kernel<TM_KM_ASM><<<grid, 256, 0, s>>>((const uint32_t *)d_X, npoints, (const uint32_t *)d_centroids, k, sc->cb,
                                                             sc->flag, d_clust, d_dis, cchunk, nullptr, nullptr, TmKmHostOut{});
    else
        kmodes_assign_kernel<TM_KM_GENERIC><<<grid, 256, 0, s>>>((const uint32_t *)d_X, npoints, (const uint32_t *)d_centroids, k,
                                                                 sc->cb, sc->flag, d_clust, d_dis, cchunk, nullptr, nullptr, TmKmHostOut{});
    tm_count_launch();
    return tm_check(cudaGetLastError(), "tm_kmodes_assign_noalloc");
}
