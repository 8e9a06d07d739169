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

/* ---- TKModes.ComputeKModes (kmodes.pas:917-1060) with KModesIter (845-909), MovePointCat (778-806),
 * CountClusterMembers (616-630), GetMaxClusterMembers (632-669, last cluster among equal counts: `>=`),
 * GetMaxValueIndex (149-161, first maximum: `>`), for the path DoKModes takes (main.pas:4217-4219):
 * ANumInit <= 0, i.e. ONE run started from point -ANumInit (kmodes.pas:941-945).  The ANumInit > 0 branch
 * (several starting points spaced by a power law in Single arithmetic) has no caller in main.pas and is not restated.
 * Seed = $42381337 (930); empty clusters after the first assignment take every attribute from a DIFFERENT random
 * point (RandInt per attribute, 1008-1009); inside an iteration centroids change between the 960-point bins.
 * max_iter <= 0 = MaxInt.  Returns the iteration count (NIter); labels[npoints], centroids[k][80], *cost, *moves. */
typedef struct {
    const uint8_t *X;
    int64_t n, k;
    int nmod, sem;
    int32_t *memb;
    uint8_t *cent;
    int32_t *freq; /* [k][80][nmod] */
} km_state;

static int km_max_value_index(const int32_t *arr, int n) { /* kmodes.pas:149-161 */
    int res = -1;
    int32_t best = INT32_MIN;
    for (int i = 0; i < n; i++)
        if (arr[i] > best) { best = arr[i]; res = i; }
    return res;
}

static void km_move_point(km_state *s, int64_t ipoint, int32_t to, int32_t from) { /* kmodes.pas:778-806 */
    const uint8_t *p = s->X + ipoint * TMO_KM_FEATURES;
    s->memb[ipoint] = to;
    for (int a = 0; a < TMO_KM_FEATURES; a++) {
        const int cur = p[a];
        int32_t *tc = s->freq + ((size_t)to * TMO_KM_FEATURES + a) * s->nmod;
        int32_t *fc = s->freq + ((size_t)from * TMO_KM_FEATURES + a) * s->nmod;
        tc[cur]++;
        const int32_t cur_freq = tc[cur];
        const int cen = s->cent[(size_t)to * TMO_KM_FEATURES + a];
        if (tc[cen] < cur_freq) s->cent[(size_t)to * TMO_KM_FEATURES + a] = (uint8_t)cur;
        fc[cur]--;
        if (s->cent[(size_t)from * TMO_KM_FEATURES + a] == cur)
            s->cent[(size_t)from * TMO_KM_FEATURES + a] = (uint8_t)km_max_value_index(fc, s->nmod);
    }
}

static int64_t km_count_members(const km_state *s, int32_t c) {
    int64_t r = 0;
    for (int64_t i = 0; i < s->n; i++) r += s->memb[i] == c;
    return r;
}

static int32_t km_max_cluster(const km_state *s) { /* kmodes.pas:632-669 */
    int64_t *cnt = (int64_t *)__builtin_calloc((size_t)s->k, sizeof(int64_t));
    for (int64_t i = 0; i < s->n; i++)
        if (s->memb[i] >= 0 && s->memb[i] < s->k) cnt[s->memb[i]]++;
    int64_t mc = 0;
    int32_t res = 0;
    for (int64_t c = 0; c < s->k; c++)
        if (cnt[c] >= mc) { mc = cnt[c]; res = (int32_t)c; }
    __builtin_free(cnt);
    return res;
}

/* one KModesIter (kmodes.pas:845-909); returns the number of moves */
static int64_t km_iter(km_state *s, uint32_t *seed, uint64_t *cost, int32_t *clust, uint64_t *dis, int32_t *choices) {
    const int64_t bin_size = 960; /* cBinSize kmodes.pas:847 */
    int64_t moves = 0;
    uint64_t acc = 0;
    for (int64_t b0 = 0; b0 < s->n; b0 += bin_size) {
        const int64_t b1 = b0 + bin_size < s->n ? b0 + bin_size : s->n;
        tmo_kmodes_assign(s->X + b0 * TMO_KM_FEATURES, b1 - b0, s->cent, s->k, s->sem, clust + b0, dis + b0, 0);
        for (int64_t i = b0; i < b1; i++) {
            acc += dis[i];
            if (s->memb[i] == clust[i]) continue;
            moves++;
            const int32_t old = s->memb[i];
            km_move_point(s, i, clust[i], old);
            if (km_count_members(s, old) == 0) { /* re-seed the emptied cluster from the largest one (890-905) */
                const int32_t from = km_max_cluster(s);
                int64_t cnt = 0;
                for (int64_t j = 0; j < s->n; j++)
                    if (s->memb[j] == from) choices[cnt++] = (int32_t)j;
                const int32_t r = choices[tmo_randint((uint32_t)cnt, seed)];
                km_move_point(s, r, old, from);
            }
        }
    }
    *cost = acc;
    return moves;
}

int tmo_compute_kmodes(const uint8_t *X, int64_t npoints, int64_t k, int64_t start_point, int nmod, int semantics,
                       int max_iter, int32_t *labels, uint8_t *centroids, uint64_t *cost_out, int64_t *moves_out) {
    km_state s = {X, npoints, k, nmod, semantics, labels, centroids, NULL};
    uint32_t seed = 0x42381337u; /* kmodes.pas:930 */
    s.freq = (int32_t *)__builtin_calloc((size_t)k * TMO_KM_FEATURES * nmod, sizeof(int32_t));
    int32_t *clust = (int32_t *)__builtin_malloc(sizeof(int32_t) * (size_t)(npoints ? npoints : 1));
    uint64_t *dis = (uint64_t *)__builtin_malloc(sizeof(uint64_t) * (size_t)(npoints ? npoints : 1));
    int32_t *choices = (int32_t *)__builtin_malloc(sizeof(int32_t) * (size_t)(npoints ? npoints : 1));
    tmo_farthest_first(X, npoints, k, start_point, semantics, centroids, NULL);             /* 968 */
    tmo_kmodes_assign(X, npoints, centroids, k, semantics, labels, NULL, 0);                /* 984-997 */
    for (int64_t i = 0; i < npoints; i++)                                                    /* 999-1001 */
        for (int a = 0; a < TMO_KM_FEATURES; a++)
            s.freq[((size_t)labels[i] * TMO_KM_FEATURES + a) * nmod + X[i * TMO_KM_FEATURES + a]]++;
    for (int64_t c = 0; c < k; c++) {                                                        /* 1003-1016 */
        if (km_count_members(&s, (int32_t)c) == 0) {
            for (int a = 0; a < TMO_KM_FEATURES; a++)
                centroids[c * TMO_KM_FEATURES + a] = X[(int64_t)tmo_randint((uint32_t)npoints, &seed) * TMO_KM_FEATURES + a];
        } else {
            for (int a = 0; a < TMO_KM_FEATURES; a++)
                centroids[c * TMO_KM_FEATURES + a] =
                    (uint8_t)km_max_value_index(s.freq + ((size_t)c * TMO_KM_FEATURES + a) * nmod, nmod);
        }
    }
    int itr = 0, converged = 0;
    uint64_t cost = ~(uint64_t)0, ncost = 0;
    int64_t total = 0;
    if (max_iter <= 0) max_iter = INT32_MAX;
    while (itr <= max_iter && !converged) {                                                  /* 1023-1039 */
        itr++;
        const int64_t moves = km_iter(&s, &seed, &ncost, clust, dis, choices);
        converged = ncost >= cost || moves == 0;
        cost = ncost;
        total += moves;
    }
    if (cost_out) *cost_out = cost;
    if (moves_out) *moves_out = total;
    __builtin_free(s.freq); __builtin_free(clust); __builtin_free(dis); __builtin_free(choices);
    return itr;
}
