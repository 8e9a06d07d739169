/* kmodes_ref.c — oracle restatement of the KModes matching dissimilarity,
 * arg-min, assignment pass and farthest-first init.
 * TEST INFRASTRUCTURE ONLY (see tmoracle.h).  PARITY UNPINNED.
 * Follows /root/reference kmodes.pas: generic 239-312, x86-64 asm 316-453 and
 * 455-596, wrappers 598-612, InitFarthestFirst 698-776, assignment 825-835 /
 * 984-997; rows main.pas:4142-4183.
 */
#include "tmoracle.h"
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

static inline int iabs(int v) { return v < 0 ? -v : v; }

/* pabsb(psubb(a,b)) on one byte: wrapping 8-bit subtract, signed abs (0x80 stays 0x80).
 * kmodes.pas:363-365 */
static inline unsigned pabsb_psubb(uint8_t a, uint8_t b) {
    int8_t d = (int8_t)(uint8_t)(a - b);
    return (d == -128) ? 128u : (unsigned)(d < 0 ? -d : d);
}

uint64_t tmo_dissim(const uint8_t *a, const uint8_t *b, int semantics) {
    uint64_t mism = 0;
    for (int i = 0; i < TMO_KM_FEATURES; i++) mism += (a[i] != b[i]);
    if (semantics == 1) { /* generic kmodes.pas:239-250 */
        uint64_t sad = 0;
        for (int i = 0; i < TMO_KM_FEATURES; i++) sad += (uint64_t)iabs((int)a[i] - (int)b[i]);
        return (mism << TMO_KM_SHIFT) + sad;
    }
    /* asm kmodes.pas:357-412.  xmm11 = byte-abs of bytes 0..15 viewed as eight
     * 16-bit words; psadbw sums of bytes 16..79 land in words 0 and 4 (paddw,
     * mod 2^16); only words 0 and 4 are extracted (pextrw). */
    unsigned w0 = pabsb_psubb(a[0], b[0]) | (pabsb_psubb(a[1], b[1]) << 8);
    unsigned w4 = pabsb_psubb(a[8], b[8]) | (pabsb_psubb(a[9], b[9]) << 8);
    for (int blk = 1; blk < 5; blk++) {
        unsigned lo = 0, hi = 0;
        for (int i = 0; i < 8; i++) {
            lo += (unsigned)iabs((int)a[blk * 16 + i] - (int)b[blk * 16 + i]);
            hi += (unsigned)iabs((int)a[blk * 16 + 8 + i] - (int)b[blk * 16 + 8 + i]);
        }
        w0 = (w0 + lo) & 0xffffu;
        w4 = (w4 + hi) & 0xffffu;
    }
    return (mism << TMO_KM_SHIFT) + w0 + w4;
}

int32_t tmo_min_dissim(const uint8_t *rows, int64_t count, const uint8_t *item,
                       int semantics, uint64_t *best) {
    int32_t res = -1;
    uint64_t b = ~(uint64_t)0;
    for (int64_t i = 0; i < count; i++) {
        uint64_t d = tmo_dissim(rows + i * TMO_KM_FEATURES, item, semantics);
        if (d <= b) { b = d; res = (int32_t)i; } /* cmp/ja kmodes.pas:414-415; generic 262 */
    }
    if (best) *best = b;
    return res;
}

void tmo_kmodes_assign(const uint8_t *X, int64_t npoints, const uint8_t *centroids,
                       int64_t k, int semantics, int32_t *clust, uint64_t *dis,
                       int nthreads) {
#ifdef _OPENMP
    if (nthreads <= 0) nthreads = omp_get_max_threads();
#endif
#pragma omp parallel for schedule(static) num_threads(nthreads)
    for (int64_t i = 0; i < npoints; i++) {
        uint64_t d;
        clust[i] = tmo_min_dissim(centroids, k, X + i * TMO_KM_FEATURES, semantics, &d);
        if (dis) dis[i] = d;
    }
}

void tmo_update_min_distance(const uint8_t *X, int64_t npoints, const uint8_t *center,
                             int semantics, uint64_t *mindist) {
    /* the `used` test at kmodes.pas:567 is `test [used],0` and never skips */
    for (int64_t i = 0; i < npoints; i++) {
        uint64_t d = tmo_dissim(center, X + i * TMO_KM_FEATURES, semantics);
        if (d < mindist[i]) mindist[i] = d; /* cmovb kmodes.pas:555-558 */
    }
}

void tmo_farthest_first(const uint8_t *X, int64_t npoints, int64_t k, int64_t init_point,
                        int semantics, uint8_t *centroids, int64_t *chosen) {
    uint64_t *mind = (uint64_t *)__builtin_malloc(sizeof(uint64_t) * (size_t)npoints);
    uint8_t *used = (uint8_t *)__builtin_calloc((size_t)npoints, 1);
    for (int64_t i = 0; i < npoints; i++) mind[i] = ~(uint64_t)0;
    int64_t far = init_point;
    for (int64_t c = 0; c < k; c++) {
        if (c > 0) {
            uint64_t mx = 0;
            far = -1;
            for (int64_t i = 0; i < npoints; i++)
                if (mind[i] >= mx && !used[i]) { mx = mind[i]; far = i; } /* kmodes.pas:765-770 */
        }
        memcpy(centroids + c * TMO_KM_FEATURES, X + far * TMO_KM_FEATURES, TMO_KM_FEATURES);
        if (chosen) chosen[c] = far;
        used[far] = 1;
        tmo_update_min_distance(X, npoints, X + far * TMO_KM_FEATURES, semantics, mind);
    }
    __builtin_free(mind);
    __builtin_free(used);
}

int tmo_kmodes_row(const uint8_t *palpix, int pal_size, uint8_t *row) {
    uint8_t acc[64];
    const int zones = 16; /* sqr(cTileWidth) div 4 main.pas:4179 */
    memset(acc, 0, sizeof acc);
    memcpy(row, palpix, 64);
    for (int i = 0; i < 64; i++) acc[palpix[i] * zones / pal_size]++;
    int res = 64;
    for (int z = 0; z < zones; z++) {
        if (64 - acc[z] < res) res = 64 - acc[z];
        row[64 + z] = acc[z] > (pal_size / zones);
    }
    return res;
}

/* Per-cluster representative of DoKModes main.pas:4231-4253: for cluster j the members are gathered in
 * point order (4235-4243) and GetMinMatchingDissim(members, centroid j) picks the nearest, last among
 * equals.  rep[j] = point index or -1 (empty cluster); count[j] = members. */
void tmo_kmodes_representatives(const uint8_t *X, int64_t npoints, const int32_t *labels,
                                const uint8_t *centroids, int64_t k, int semantics, int32_t *rep, int32_t *count) {
    for (int64_t j = 0; j < k; j++) {
        uint64_t best = ~(uint64_t)0;
        int32_t bi = -1, cnt = 0;
        for (int64_t i = 0; i < npoints; i++) {
            if (labels[i] != j) continue;
            cnt++;
            uint64_t d = tmo_dissim(centroids + j * TMO_KM_FEATURES, X + i * TMO_KM_FEATURES, semantics);
            if (d <= best) { best = d; bi = (int32_t)i; }
        }
        rep[j] = bi;
        if (count) count[j] = cnt;
    }
}
