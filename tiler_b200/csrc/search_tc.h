// search_tc.h — pieces of the tensor-core search shared between search_tc.cu and rerank.cu
#pragma once
#include "tm_common.cuh"

// ---- error model of the coarse pass, shared by the kernels' admission band and the certificate ----------
// The coarse score is  s~ = ||t16||^2 - 2 q16.t16  with q16 = fp16(q - mu), t16 = fp16(t - mu), products exact,
// fp32 accumulation.  Against the real squared distance d = ||q - t||^2:
//   | sqrt(d16) - sqrt(d) | <= e,   d16 = ||q16 - t16||^2,   e = ||dq|| + ||dt||   (triangle inequality)
//   | s~ + ||q16||^2 - d16 |  <= a                                                   (accumulation)
// e: dq = (q - mu) - q16 is measured per query, ||dt|| <= max over the rows (measured when the dataset is
//    built); 0.1 % head-room, the fp32 rounding of the centring itself (2^-24 per element, 1.3e-7 * nsum) and
//    an absolute floor are added.  Typically 0.4 x the worst case 2^-11 (||q~|| + max ||t~||) used before.
// a: 195 additions, each losing at most one ulp (2^-23, truncation) of a partial sum bounded by
//    2 ||q16|| ||t16|| + ||t16||^2 <= nsum^2  ->  2.33e-5 nsum^2; 3e-5 is used.
__host__ __device__ __forceinline__ float tc_band_e(float qdnorm, float max_dnorm, float nsum) {
    return 1.001f * (qdnorm + max_dnorm) + 1.3e-7f * nsum + 1e-6f;
}
__host__ __device__ __forceinline__ float tc_band_a(float nsum) { return 3e-5f * nsum * nsum + 1e-6f; }

// ---- work-item plan of one coarse pass ---------------------------------------------------------------------------
// A work item = (256-query block, row split).  Blocks [0, qb1) are cut into ns1 row splits of tps1 tiles, blocks
// [qb1, qblocks) - the TAIL, when the block count does not fill the last round of CTA pairs - into ns2 finer
// splits of tps2 tiles, so that the last round costs a fraction of an item instead of a whole one (a 4-frame batch
// is 225 blocks on 74 pairs: 3.04 rounds).  Every (query, split, accumulator column half) owns one candidate list;
// a query of the first part has nl1 = 2 ns1 lists, one of the tail nl2 = 2 ns2, stored back to back.
// chain != 0 (k > 1, tilesets beyond the L2, no tail): the items run SPLIT-major - every query block's split 0, then
// every block's split 1, ... - so that all CTA pairs scan the same L2-sized stretch of the operand at the same time,
// and an item CONTINUES the candidate lists and the admission threshold its query block reached in the previous split
// (search_tc2_kernel.inc): a query then has one list per column half, as with a single split.
struct TcPlan {
    int qblocks, qb1, ns1, tps1, ns2, tps2, total_tiles, chain;
    __host__ __device__ int items() const { return qb1 * ns1 + (qblocks - qb1) * ns2; }
    // item -> (query block, split, first tile, end tile)
    __host__ __device__ void decode(int item, int &qb, int &sp, int &t0, int &t1) const {
        int tps;
        if (chain) { // split-major; qb1 == qblocks
            sp = item / qblocks;
            qb = item - sp * qblocks;
            tps = tps1;
        } else if (item < qb1 * ns1) {
            qb = item / ns1;
            sp = item - qb * ns1;
            tps = tps1;
        } else {
            const int j = item - qb1 * ns1;
            qb = qb1 + j / ns2;
            sp = j - (j / ns2) * ns2;
            tps = tps2;
        }
        t0 = sp * tps;
        t1 = t0 + tps < total_tiles ? t0 + tps : total_tiles;
    }
    __host__ __device__ int nlists(int64_t q) const { return chain ? 2 : q < (int64_t)qb1 * 256 ? 2 * ns1 : 2 * ns2; }
    __host__ __device__ int64_t list_base(int64_t q) const { // index of query q's first list
        const int64_t q1 = (int64_t)qb1 * 256;
        if (chain) return q * 2;
        return q < q1 ? q * (2 * ns1) : q1 * (2 * ns1) + (q - q1) * (2 * ns2);
    }
    __host__ __device__ int64_t total_lists() const { return list_base((int64_t)qblocks * 256); }
    __host__ __device__ int max_nlists() const {
        if (chain) return 2;
        const int a = qb1 > 0 ? ns1 : 0, b = qb1 < qblocks ? ns2 : 0;
        return 2 * (a > b ? a : b);
    }
};

int tm_launch_query_operand(const ANNkdtree *h, const float *d_q, int64_t nq, int64_t nq_pad, __half *d_q16,
                            float *d_qnorm16, float *d_qnorm_exact, float *d_qdnorm, cudaStream_t s);
int tm_make_query_tmap(void *tmap, const __half *d_q16, int64_t nq_pad, int kp);

// exact float32 re-rank of the coarse candidates + certification (rerank.cu, compiled -fmad=false)
int tm_launch_rerank(const float *d_rows, int dd, const float *d_q, int64_t nq, const float *d_cand_s,
                     const int *d_cand_i, const float *d_cand_tau, const TcPlan &plan, int kc, const float *d_qnorm16, const float *d_qnorm_exact,
                     const float *d_qdnorm, float max_tnorm, float max_dnorm, int k, int *d_idx, float *d_err, int *d_fb_list, int *d_fb_count,
                     cudaStream_t s);
