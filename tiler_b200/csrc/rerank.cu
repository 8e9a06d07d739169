// rerank.cu — exact float32 re-rank of the tensor-core candidates, and the certificate that
// decides which queries must be re-run on the exact engine.  Compiled with -fmad=false: the
// distance loop is the same sequential multiply / add chain as search_simt.cu, i.e. the
// arithmetic of the reference's ANN leaf scan (main.pas:4027), so certified results are
// bit-identical to a full exact scan.
#include "search_tc.h"

namespace {

constexpr int RR_QPB = 16; // queries per block, 8 lanes each

__global__ void __launch_bounds__(RR_QPB * 8) rerank_kernel(
    const float *__restrict__ rows, int dd, const float *__restrict__ q, int64_t nq,
    const float *__restrict__ cand_s, const int *__restrict__ cand_i, const float *__restrict__ cand_tau, int nsplit, int kc,
    const float *__restrict__ qn16, const float *__restrict__ qne, float max_tnorm, int k,
    int *__restrict__ idx, float *__restrict__ err, int *__restrict__ fb_list, int *__restrict__ fb_count) {
    extern __shared__ float sm[];
    const int C = nsplit * kc;
    float *sd = sm;                                        // [RR_QPB][C]
    int *si = reinterpret_cast<int *>(sm + RR_QPB * C);   // [RR_QPB][C]
    const int ql = threadIdx.x >> 3, l8 = threadIdx.x & 7;
    const int64_t qg = (int64_t)blockIdx.x * RR_QPB + ql;
    if (qg < nq) {
        const float *qq = q + qg * dd;
        // Rows scored at or above the final admission threshold are covered by the certificate's lower
        // bound exactly like rows that were never admitted: only the ones below it need exact distances.
        float tau_min = INFINITY;
        for (int sp = 0; sp < nsplit; sp++) tau_min = fminf(tau_min, cand_tau[qg * nsplit + sp]);
        for (int c = l8; c < C; c += 8) {
            int ri = cand_i[qg * C + c];
            if (!(cand_s[qg * C + c] < tau_min) && tau_min < INFINITY) ri = -1;
            float acc = INFINITY;
            if (ri >= 0) {
                const float *row = rows + (int64_t)ri * dd;
                acc = 0.f;
                if ((dd & 3) == 0) { // 128-bit loads, still one sequential add chain in dimension order
                    const float4 *q4 = reinterpret_cast<const float4 *>(qq);
                    const float4 *r4 = reinterpret_cast<const float4 *>(row);
                    for (int j = 0; j < dd / 4; j++) {
                        const float4 a = q4[j], b = __ldg(r4 + j);
                        float t0 = a.x - b.x; acc = acc + t0 * t0;
                        float t1 = a.y - b.y; acc = acc + t1 * t1;
                        float t2 = a.z - b.z; acc = acc + t2 * t2;
                        float t3 = a.w - b.w; acc = acc + t3 * t3;
                    }
                } else {
                    for (int j = 0; j < dd; j++) { // sequential, no FMA: matches search_simt.cu
                        float t = qq[j] - __ldg(row + j);
                        acc = acc + t * t;
                    }
                }
            }
            sd[ql * C + c] = acc;
            si[ql * C + c] = ri >= 0 ? ri : 0x7fffffff;
        }
    }
    __syncthreads();
    if (qg >= nq || l8 != 0) return;
    // k smallest by (distance, index); C is small
    float *d = sd + ql * C;
    int *ii = si + ql * C;
    float dk = 0.f;
    for (int r = 0; r < k; r++) {
        int bp = -1;
        float bd = INFINITY;
        int bi = 0x7fffffff;
        for (int c = 0; c < C; c++) {
            if (ii[c] == 0x7fffffff) continue;
            if (d[c] < bd || (d[c] == bd && ii[c] < bi)) { bd = d[c]; bi = ii[c]; bp = c; }
        }
        if (bp >= 0) {
            idx[qg * k + r] = bi; err[qg * k + r] = bd; ii[bp] = 0x7fffffff; dk = bd;
        } else {
            idx[qg * k + r] = -1; err[qg * k + r] = INFINITY; dk = INFINITY;
        }
    }
    // certificate: no row outside the candidate lists can be at or below the k-th exact distance
    float tau = INFINITY;
    for (int sp = 0; sp < nsplit; sp++) tau = fminf(tau, cand_tau[qg * nsplit + sp]); // admission threshold at scan end
    // NaN / overflowed fp16 query operands never certify (their scores are NaN and select nothing)
    bool ok = qn16[qg] < 3.0e38f;
    if (ok && tau < INFINITY) {
        const double nsum = (double)qne[qg] + (double)max_tnorm;
        const double a = 6e-5 * nsum * nsum + 1e-6;                   // fp32 accumulation slack
        const double e = nsum * (1.001 / 2048.0) + 1e-6;              // fp16 rounding of both vectors
        const double dt = (double)tau + (double)qn16[qg] - a;
        double root = sqrt(dt > 0.0 ? dt : 0.0) - e;
        if (root < 0.0) root = 0.0;
        const double lb = root * root * (1.0 - 2e-5);
        ok = lb > (double)dk && qn16[qg] < 3.0e38f; // NaN / overflowed fp16 operands never certify
    }
    if (!ok) fb_list[atomicAdd(fb_count, 1)] = (int)qg;
}

} // namespace

int tm_launch_rerank(const float *d_rows, int dd, const float *d_q, int64_t nq, const float *d_cand_s,
                     const int *d_cand_i, const float *d_cand_tau, int nsplit, int kc, const float *d_qnorm16, const float *d_qnorm_exact,
                     float max_tnorm, int k, int *d_idx, float *d_err, int *d_fb_list, int *d_fb_count,
                     cudaStream_t s) {
    const size_t smem = (size_t)RR_QPB * nsplit * kc * 8;
    rerank_kernel<<<(unsigned)((nq + RR_QPB - 1) / RR_QPB), RR_QPB * 8, smem, s>>>(
        d_rows, dd, d_q, nq, d_cand_s, d_cand_i, d_cand_tau, nsplit, kc, d_qnorm16, d_qnorm_exact, max_tnorm, k, d_idx, d_err,
        d_fb_list, d_fb_count);
    tm_count_launch();
    return tm_check(cudaGetLastError(), "rerank_kernel");
}
