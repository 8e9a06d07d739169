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

int tm_launch_query_operand(const ANNkdtree *h, const float *d_q, int64_t nq, int64_t nq_pad, __half *d_q16,
                            float *d_qnorm16, float *d_qnorm_exact, float *d_qdnorm, cudaStream_t s);
int tm_make_query_tmap(void *tmap, const __half *d_q16, int64_t nq_pad, int kp);

// exact float32 re-rank of the coarse candidates + certification (rerank.cu, compiled -fmad=false)
int tm_launch_rerank(const float *d_rows, int dd, const float *d_q, int64_t nq, const float *d_cand_s,
                     const int *d_cand_i, const float *d_cand_tau, int nsplit, int kc, const float *d_qnorm16, const float *d_qnorm_exact,
                     const float *d_qdnorm, float max_tnorm, float max_dnorm, int k, int *d_idx, float *d_err, int *d_fb_list, int *d_fb_count,
                     cudaStream_t s);
