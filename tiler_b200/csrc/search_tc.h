// search_tc.h — pieces of the tensor-core search shared between search_tc.cu and rerank.cu
#pragma once
#include "tm_common.cuh"

int tm_launch_query_operand(const ANNkdtree *h, const float *d_q, int64_t nq, int64_t nq_pad, __half *d_q16,
                            float *d_qnorm16, float *d_qnorm_exact, cudaStream_t s);
int tm_make_query_tmap(void *tmap, const __half *d_q16, int64_t nq_pad, int kp);

// exact float32 re-rank of the coarse candidates + certification (rerank.cu, compiled -fmad=false)
int tm_launch_rerank(const float *d_rows, int dd, const float *d_q, int64_t nq, const float *d_cand_s,
                     const int *d_cand_i, const float *d_cand_tau, int nsplit, int kc, const float *d_qnorm16, const float *d_qnorm_exact,
                     float max_tnorm, int k, int *d_idx, float *d_err, int *d_fb_list, int *d_fb_count,
                     cudaStream_t s);
