// rerank.cu — exact float32 re-rank of the tensor-core candidates, and the certificate that
// decides which queries must be re-run on the exact engine.  Compiled with -fmad=false: the
// distance loop is the same sequential multiply / add chain as search_simt.cu, i.e. the
// arithmetic of the reference's ANN leaf scan (main.pas:4027), so certified results are
// bit-identical to a full exact scan.
#include "search_tc.h"

namespace {

constexpr int RR_THREADS = 128; // one thread per query
constexpr int RR_MAXV = 64;     // candidates below the final threshold a query may carry (more -> rejected) ...
constexpr int RR_MAXV_WIDE = 192; // ... when the lists hold more than 64 slots (finest row split, k > 4)
constexpr int RR_K = 8;         // the tensor-core path serves k <= 8

// exact float32 squared distance, one sequential add chain in dimension order, no FMA (-fmad=false):
// the arithmetic of search_simt.cu and of the reference's ANN leaf scan
__device__ __forceinline__ float exact_dist(const float *__restrict__ qq, const float *__restrict__ row, int dd) {
    float acc = 0.f;
    if ((dd & 3) == 0) {
        const float4 *q4 = reinterpret_cast<const float4 *>(qq);
        const float4 *r4 = reinterpret_cast<const float4 *>(row);
#pragma unroll 4
        for (int j = 0; j < dd / 4; j++) {
            const float4 a = q4[j], b = __ldg(r4 + j);
            float t0 = a.x - b.x; acc = acc + t0 * t0;
            float t1 = a.y - b.y; acc = acc + t1 * t1;
            float t2 = a.z - b.z; acc = acc + t2 * t2;
            float t3 = a.w - b.w; acc = acc + t3 * t3;
        }
    } else {
        for (int j = 0; j < dd; j++) {
            float t = qq[j] - __ldg(row + j);
            acc = acc + t * t;
        }
    }
    return acc;
}

__device__ __forceinline__ void cp_async16(float *smem_dst, const void *gsrc) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// Thread = query.  Pass 1 compacts the row ids of the candidates that scored below the final admission
// threshold (usually one or two of the nlists * kc slots) into a per-thread shared-memory list; pass 2
// walks that list in rounds, so a warp executes the 192-step distance chain max-over-lanes(count)
// times with every lane busy.  In a round the warp stages, 32 dimensions at a time, the matching
// 128-byte segments of its 32 query rows and 32 candidate rows in shared memory with cooperative
// 16-byte loads (8 lanes per row segment: full sectors, no L1 thrash), then every lane runs the
// sequential add chain over its own segment pair.
constexpr int RR_SEG = 32;            // dimensions per staged segment
constexpr int RR_SEG_STRIDE = 36;     // floats per staged row segment (pad: conflict-free 16-byte reads)

template <bool K1, bool STAGED>
__global__ void __launch_bounds__(RR_THREADS) rerank_kernel(
    const float *__restrict__ rows, int dd, const float *__restrict__ q, int64_t nq,
    const float *__restrict__ cand_s, const int *__restrict__ cand_i, const float *__restrict__ cand_tau, const TcPlan plan, int kc,
    const float *__restrict__ qn16, const float *__restrict__ qne, const float *__restrict__ qde, float max_tnorm,
    float max_dnorm, int k, int maxv,
    int *__restrict__ idx, float *__restrict__ err, int *__restrict__ fb_list, int *__restrict__ fb_count,
    unsigned long long *__restrict__ audit) {
    extern __shared__ uint8_t vlist_dyn[]; // [maxv][RR_THREADS] slot positions (< nlists * kc <= 256) of the live candidates
    uint8_t (*vlist)[RR_THREADS] = reinterpret_cast<uint8_t (*)[RR_THREADS]>(vlist_dyn);
    __shared__ __align__(16) float seg[STAGED ? RR_THREADS / 32 : 1][2][32 * RR_SEG_STRIDE];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t qg = (int64_t)blockIdx.x * RR_THREADS + threadIdx.x;
    const bool live = qg < nq;
    const int64_t qc = live ? qg : nq - 1; // clamped: dead lanes still help with the cooperative loads
    // this query's candidate lists (search_tc.h: a query of the plan's tail has more, shorter-range lists)
    const int nlists = plan.nlists(qc);
    const int64_t lbase = plan.list_base(qc);
    const int C = nlists * kc;
    // Rows scored at or above the final admission threshold are covered by the certificate's lower bound
    // exactly like rows that were never admitted: only the ones below it need exact distances.
    float tau = INFINITY;
    int cnt = 0;
    bool overflow = false;
    if (live) {
        for (int l = 0; l < nlists; l++) tau = fminf(tau, cand_tau[lbase + l]);
        const int4 *ci4 = reinterpret_cast<const int4 *>(cand_i + lbase * kc);
        const float4 *cs4 = reinterpret_cast<const float4 *>(cand_s + lbase * kc);
        for (int c4 = 0; c4 < C / 4; c4++) { // kc is 8 or 16: C is a multiple of 4, the lists are 16-byte aligned
            const int4 ri = ci4[c4];
            const float4 sc = cs4[c4];
            const int r[4] = {ri.x, ri.y, ri.z, ri.w};
            const float s[4] = {sc.x, sc.y, sc.z, sc.w};
#pragma unroll
            for (int e = 0; e < 4; e++)
                if (r[e] >= 0 && (s[e] < tau || !(tau < INFINITY))) {
                    if (cnt < maxv) vlist[cnt][threadIdx.x] = (uint8_t)(c4 * 4 + e); else overflow = true;
                    cnt++;
                }
        }
        if (cnt > maxv) cnt = maxv;
    }
    float bd[K1 ? 1 : RR_K];
    int bi[K1 ? 1 : RR_K];
#pragma unroll
    for (int j = 0; j < (K1 ? 1 : RR_K); j++) { bd[j] = INFINITY; bi[j] = 0x7fffffff; }
    const int rounds = STAGED ? __reduce_max_sync(0xffffffffu, cnt) : cnt;
    for (int v = 0; v < rounds; v++) {
        const bool act = v < cnt;
        const int ri = act ? __ldg(cand_i + lbase * kc + vlist[v][threadIdx.x]) : 0;
        float d;
        if (STAGED) {
            float acc = 0.f;
            float *sq = seg[warp][0], *sr = seg[warp][1];
            // rounds after the first usually serve one or two lanes: stage only their rows
            const unsigned actmask = __ballot_sync(0xffffffffu, act);
            for (int j0 = 0; j0 < dd; j0 += RR_SEG) {
                const int seg4 = min(RR_SEG, dd - j0) >> 2; // float4 per row segment (dd % 4 == 0)
                __syncwarp();
                // Asynchronous 16-byte copies straight into shared memory (LDGSTS): all sixteen of a lane are in flight at
                // once and cost no registers.  With loads through registers each iteration waited for its own two before
                // it stored them - 25 KB in flight per SM, 3.6 MB on the chip against the ~6.5 MB that 6.5 TB/s x 1 us asks
                // for (ncu: long-scoreboard stalls 13.5 per issue, 43 % of HBM).
#pragma unroll
                for (int e = lane; e < 32 * 8; e += 32) { // 8 float4 slots per row, 32 rows
                    const int rr = e >> 3, f = e & 7;
                    const int64_t qrow = __shfl_sync(0xffffffffu, qc, rr);
                    const int crow = __shfl_sync(0xffffffffu, ri, rr);
                    if (f < seg4 && ((actmask >> rr) & 1u)) {
                        cp_async16(sq + rr * RR_SEG_STRIDE + f * 4, reinterpret_cast<const float4 *>(q + qrow * dd + j0) + f);
                        cp_async16(sr + rr * RR_SEG_STRIDE + f * 4, reinterpret_cast<const float4 *>(rows + (int64_t)crow * dd + j0) + f);
                    }
                }
                cp_async_wait_all();
                __syncwarp();
                const float4 *a4 = reinterpret_cast<const float4 *>(sq + lane * RR_SEG_STRIDE);
                const float4 *b4 = reinterpret_cast<const float4 *>(sr + lane * RR_SEG_STRIDE);
#pragma unroll
                for (int f = 0; f < RR_SEG / 4; f++)
                    if (f < seg4) { // one sequential add chain in dimension order, no FMA
                        const float4 a = a4[f], b = b4[f];
                        float t0 = a.x - b.x; acc = acc + t0 * t0;
                        float t1 = a.y - b.y; acc = acc + t1 * t1;
                        float t2 = a.z - b.z; acc = acc + t2 * t2;
                        float t3 = a.w - b.w; acc = acc + t3 * t3;
                    }
            }
            d = acc;
        } else {
            d = act ? exact_dist(q + qc * dd, rows + (int64_t)ri * dd, dd) : 0.f;
        }
        if (!act) continue;
        if (audit) { // error-model audit: observed |s~ + ||q16||^2 - d| against the bound the certificate assumes
            const float sc = __ldg(cand_s + lbase * kc + vlist[v][threadIdx.x]);
            const float nsumf = qne[qg] + max_tnorm;
            const float e = tc_band_e(qde[qg], max_dnorm, nsumf), a = tc_band_a(nsumf);
            const float ratio = fabsf(sc + qn16[qg] - d) / (2.f * e * sqrtf(d) + e * e + a);
            if (ratio == ratio && __float_as_uint(ratio) > (unsigned int)audit[0])
                atomicMax(audit, (unsigned long long)__float_as_uint(ratio));
            atomicAdd(audit + 1, 1ull);
        }
        // ascending by (distance, index)
        if (K1) {
            if (d < bd[0] || (d == bd[0] && ri < bi[0])) { bd[0] = d; bi[0] = ri; }
        } else {
            float nd = d;
            int ni = ri;
#pragma unroll
            for (int j = 0; j < RR_K; j++) {
                const bool sw = nd < bd[j] || (nd == bd[j] && ni < bi[j]);
                const float td = sw ? bd[j] : nd;
                const int ti = sw ? bi[j] : ni;
                bd[j] = sw ? nd : bd[j];
                bi[j] = sw ? ni : bi[j];
                nd = td;
                ni = ti;
            }
        }
    }
    if (!live) return;
    float dk = 0.f;
#pragma unroll
    for (int j = 0; j < (K1 ? 1 : RR_K); j++)
        if (j < k) {
            const bool have = bi[j] != 0x7fffffff;
            idx[qg * k + j] = have ? bi[j] : -1;
            err[qg * k + j] = have ? bd[j] : INFINITY;
            dk = have ? bd[j] : INFINITY;
        }
    // certificate: no row outside the candidate lists can be at or below the k-th exact distance.
    // NaN / overflowed fp16 query operands never certify (their scores are NaN and select nothing).
    // A k-th distance must exist: with no candidate at all (dk = inf) there is nothing the bound could be compared
    // with - the exact engine answers (it also serves datasets with fewer than k rows).
    bool ok = !overflow && qn16[qg] < 3.0e38f && dk < INFINITY;
    if (ok && tau < INFINITY) {
        const float nsumf = qne[qg] + max_tnorm;
        const double nsum = (double)nsumf;
        const double a = (double)tc_band_a(nsumf);                        // fp32 accumulation slack (search_tc.h)
        const double e = (double)tc_band_e(qde[qg], max_dnorm, nsumf);    // fp16 rounding of both vectors
        (void)nsum;
        const double dt = (double)tau + (double)qn16[qg] - a;
        double root = sqrt(dt > 0.0 ? dt : 0.0) - e;
        if (root < 0.0) root = 0.0;
        const double lb = root * root * (1.0 - 2e-5);
        ok = lb > (double)dk;
    }
    if (!ok) fb_list[atomicAdd(fb_count, 1)] = (int)qg;
}

} // namespace

int tm_launch_rerank(const float *d_rows, int dd, const float *d_q, int64_t nq, const float *d_cand_s,
                     const int *d_cand_i, const float *d_cand_tau, const TcPlan &plan, int kc, const float *d_qnorm16, const float *d_qnorm_exact,
                     const float *d_qdnorm, float max_tnorm, float max_dnorm, int k, int *d_idx, float *d_err, int *d_fb_list, int *d_fb_count,
                     cudaStream_t s) {
    const unsigned grid = (unsigned)((nq + RR_THREADS - 1) / RR_THREADS);
    const bool staged = (dd & 3) == 0 && (((uintptr_t)d_q | (uintptr_t)d_rows) & 15) == 0;
    const int maxv = plan.max_nlists() * kc > RR_MAXV ? RR_MAXV_WIDE : RR_MAXV;
    const size_t dyn = (size_t)maxv * RR_THREADS;
    static std::atomic<uint32_t> attr{0}; // per device (tm_first_on_device)
    if (tm_first_on_device(attr)) {
        const int wide = RR_MAXV_WIDE * RR_THREADS;
        TM_CUDA(cudaFuncSetAttribute(rerank_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, wide));
        TM_CUDA(cudaFuncSetAttribute(rerank_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, wide));
        TM_CUDA(cudaFuncSetAttribute(rerank_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, wide));
        TM_CUDA(cudaFuncSetAttribute(rerank_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, wide));
    }
#define RR_LAUNCH(K1, ST) rerank_kernel<K1, ST><<<grid, RR_THREADS, dyn, s>>>(d_rows, dd, d_q, nq, d_cand_s, d_cand_i, d_cand_tau, plan, kc, d_qnorm16, d_qnorm_exact, d_qdnorm, max_tnorm, max_dnorm, k, maxv, d_idx, d_err, d_fb_list, d_fb_count, tm_audit_slots())
    if (k == 1) { if (staged) RR_LAUNCH(true, true); else RR_LAUNCH(true, false); }
    else { if (staged) RR_LAUNCH(false, true); else RR_LAUNCH(false, false); }
#undef RR_LAUNCH
    tm_count_launch();
    return tm_check(cudaGetLastError(), "rerank_kernel");
}
