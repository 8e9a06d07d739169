/* ft_ref.c — oracle restatement of the exact nearest-neighbour search that the
 * reference performs through ANN.dll with eps = 0 (main.pas:3830, 4027; FFI
 * extern.pas:63-67).  TEST INFRASTRUCTURE ONLY (see tmoracle.h).  PARITY UNPINNED.
 *
 * ANN (Mount & Arya; private float build shipped only as ANN.dll, source absent
 * from /root/reference, version not recoverable) returns, for eps = 0, the exact
 * k nearest rows under squared Euclidean distance, ascending.  A brute-force scan
 * therefore defines the same answer up to float rounding and tie order.
 */
#include "tmoracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

int tmo_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* torchrun exports OMP_NUM_THREADS=1 to its workers; the timed CPU baseline resets the team size
 * to the host's core count explicitly. */
void tmo_set_threads(int n) {
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}

/* insert (d,i) into ascending list of length k+1 (extra slot = next distance);
 * strict < keeps the lowest index among equal distances. */
static inline void ins_f64(double *bd, int32_t *bi, int k1, double d, int32_t i) {
    if (!(d < bd[k1 - 1])) return;
    int p = k1 - 1;
    while (p > 0 && d < bd[p - 1]) { bd[p] = bd[p - 1]; bi[p] = bi[p - 1]; p--; }
    bd[p] = d; bi[p] = i;
}

void tmo_knn_f64(const float *data, int64_t n, int dd, const float *q, int64_t nq,
                 int k, int32_t *idx, double *dist, double *next_dist) {
#pragma omp parallel for schedule(dynamic, 16)
    for (int64_t qi = 0; qi < nq; qi++) {
        double bd[65]; int32_t bi[65];
        int k1 = k + 1;
        for (int j = 0; j < k1; j++) { bd[j] = INFINITY; bi[j] = -1; }
        const float *qq = q + qi * dd;
        for (int64_t r = 0; r < n; r++) {
            const float *row = data + r * dd;
            double acc = 0.0;
            for (int j = 0; j < dd; j++) { double t = (double)qq[j] - (double)row[j]; acc += t * t; }
            ins_f64(bd, bi, k1, acc, (int32_t)r);
        }
        for (int j = 0; j < k; j++) { idx[qi * k + j] = bi[j]; dist[qi * k + j] = bd[j]; }
        if (next_dist) next_dist[qi] = bd[k];
    }
}

void tmo_knn_f32(const float *data, int64_t n, int dd, const float *q, int64_t nq,
                 int k, int32_t *idx, float *dist) {
#pragma omp parallel for schedule(dynamic, 16)
    for (int64_t qi = 0; qi < nq; qi++) {
        float bd[64]; int32_t bi[64];
        for (int j = 0; j < k; j++) { bd[j] = INFINITY; bi[j] = -1; }
        const float *qq = q + qi * dd;
        for (int64_t r = 0; r < n; r++) {
            const float *row = data + r * dd;
            volatile float acc = 0.0f; /* sequential, no reassociation */
            for (int j = 0; j < dd; j++) { float t = qq[j] - row[j]; acc = acc + t * t; }
            float d = acc;
            if (d < bd[k - 1]) {
                int p = k - 1;
                while (p > 0 && d < bd[p - 1]) { bd[p] = bd[p - 1]; bi[p] = bi[p - 1]; p--; }
                bd[p] = d; bi[p] = (int32_t)r;
            }
        }
        for (int j = 0; j < k; j++) { idx[qi * k + j] = bi[j]; dist[qi * k + j] = bd[j]; }
    }
}

/* Timed CPU baseline: the reference's DoFrameTiling loop (main.pas:4020-4038)
 * with the exact search done as a cache-blocked brute-force scan.  One query
 * block per thread (the reference threads per frame, main.pas:972); 8 queries
 * share each streamed dataset row so the compiler can vectorise over dd. */
#define QB 8
void tmo_nn1_blocked_f32(const float *data, int64_t n, int dd, const float *q,
                         int64_t nq, int32_t *idx, float *dist, int nthreads) {
#ifdef _OPENMP
    if (nthreads <= 0) nthreads = omp_get_max_threads();
#endif
#pragma omp parallel for schedule(dynamic, 4) num_threads(nthreads)
    for (int64_t q0 = 0; q0 < nq; q0 += QB) {
        int nb = (int)((nq - q0) < QB ? (nq - q0) : QB);
        float bd[QB]; int32_t bi[QB];
        for (int b = 0; b < QB; b++) { bd[b] = INFINITY; bi[b] = -1; }
        for (int64_t r = 0; r < n; r++) {
            const float *row = data + r * dd;
            for (int b = 0; b < nb; b++) {
                const float *qq = q + (q0 + b) * dd;
                float acc = 0.0f;
#pragma omp simd reduction(+ : acc)
                for (int j = 0; j < dd; j++) { float t = qq[j] - row[j]; acc += t * t; }
                if (acc < bd[b]) { bd[b] = acc; bi[b] = (int32_t)r; }
            }
        }
        for (int b = 0; b < nb; b++) { idx[q0 + b] = bi[b]; dist[q0 + b] = bd[b]; }
    }
}
