/* kdtree_ref.c — oracle restatement of the search structure the reference uses through
 * ANN.dll: ann_kdtree_create(pa, n, dd, bs = 1, ANN_KD_STD) and annkSearch with eps = 0
 * (call sites main.pas:3779, 3961 create; 3830, 4027 search; FFI extern.pas:63-67).
 * TEST INFRASTRUCTURE ONLY (see tmoracle.h).  PARITY UNPINNED.
 *
 * ANN (D. Mount & S. Arya, "ANN: A Library for Approximate Nearest Neighbor Searching") is a
 * third-party dependency shipped only as a Win64 binary; its source is not under
 * /root/reference and the version is not recoverable from the DLL.  This file restates the
 * library's PUBLISHED algorithm, written from its description, not from its source:
 *   build   bounding box of all points; recursively split on the dimension of maximum point
 *           spread at the median (the "standard" kd split rule), cut value midway between
 *           the two middle points; leaves hold <= bs points; every split node remembers the
 *           box extent along its cut dimension;
 *   search  depth-first, nearer child first, with the incremental box-distance update of
 *           Arya & Mount: entering the far child replaces the old offset along the cut
 *           dimension by the distance to the cutting plane; the far child is visited only
 *           while box_dist * (1+eps)^2 < current k-th best; leaf points are scanned with a
 *           float32 sum in dimension order that stops as soon as it exceeds the k-th best.
 * Coordinates and distances are float32 (TANNFloat = Single, extern.pas:30), distances are
 * squared.  With eps = 0 the result is the exact k-NN set; tie order depends on traversal.
 *
 * It serves two purposes: (1) an independent cross-check of the brute-force oracle
 * (tests/test_oracle_known_answers.py), (2) the "ANN exact search" variant of the timed CPU
 * baseline.
 */
#include "tmoracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

typedef struct kd_node {
    int32_t cut_dim;        /* -1 for a leaf */
    float cut_val;
    float lo_bound, hi_bound; /* box extent along cut_dim at this node */
    int32_t child[2];       /* node indices (split) */
    int32_t first, count;   /* leaf: range in perm[] */
} kd_node;

struct tmo_kdtree {
    const float *data; /* not owned: the caller keeps rows alive, like ANN (main.pas:3784-3788) */
    int64_t n;
    int dd, bs;
    int32_t *perm;
    kd_node *nodes;
    int32_t n_nodes, cap_nodes;
    float *box_lo, *box_hi;
    int32_t root;
};

static int32_t new_node(struct tmo_kdtree *t) {
    if (t->n_nodes == t->cap_nodes) {
        t->cap_nodes = t->cap_nodes ? t->cap_nodes * 2 : 1024;
        t->nodes = (kd_node *)realloc(t->nodes, sizeof(kd_node) * (size_t)t->cap_nodes);
    }
    return t->n_nodes++;
}

#define PT(t, i, d) ((t)->data[(int64_t)(i) * (t)->dd + (d)])

/* dimension with the largest (max - min) over the points of perm[first .. first+cnt) */
static int max_spread_dim(const struct tmo_kdtree *t, int32_t first, int32_t cnt) {
    int best = 0;
    float best_spread = -1.f;
    for (int d = 0; d < t->dd; d++) {
        float mn = PT(t, t->perm[first], d), mx = mn;
        for (int32_t i = 1; i < cnt; i++) {
            float v = PT(t, t->perm[first + i], d);
            if (v < mn) mn = v;
            if (v > mx) mx = v;
        }
        if (mx - mn > best_spread) { best_spread = mx - mn; best = d; }
    }
    return best;
}

/* quickselect: afterwards perm[first .. first+n_lo) <= perm[first+n_lo ..) along dim d, and
 * perm[first+n_lo-1] is the largest of the low side, perm[first+n_lo] the smallest of the high side */
static void median_split(struct tmo_kdtree *t, int32_t first, int32_t cnt, int d, int32_t n_lo) {
    int32_t *p = t->perm + first;
    int32_t l = 0, r = cnt - 1;
    while (l < r) {
        int32_t i = (l + r) / 2, tmp;
        if (PT(t, p[i], d) > PT(t, p[r], d)) { tmp = p[i]; p[i] = p[r]; p[r] = tmp; }
        tmp = p[l]; p[l] = p[i]; p[i] = tmp;
        float c = PT(t, p[l], d);
        i = l;
        int32_t k = r;
        for (;;) {
            while (PT(t, p[++i], d) < c) {}
            while (PT(t, p[--k], d) > c) {}
            if (i < k) { tmp = p[i]; p[i] = p[k]; p[k] = tmp; } else break;
        }
        tmp = p[l]; p[l] = p[k]; p[k] = tmp;
        if (k > n_lo) r = k - 1;
        else if (k < n_lo) l = k + 1;
        else break;
    }
    if (n_lo > 0) { /* move the maximum of the low side next to the cut */
        float c = PT(t, p[0], d);
        int32_t k = 0;
        for (int32_t i = 1; i < n_lo; i++)
            if (PT(t, p[i], d) > c) { c = PT(t, p[i], d); k = i; }
        int32_t tmp = p[n_lo - 1]; p[n_lo - 1] = p[k]; p[k] = tmp;
    }
}

static int32_t build_rec(struct tmo_kdtree *t, int32_t first, int32_t cnt) {
    int32_t id = new_node(t);
    if (cnt <= t->bs) {
        kd_node *nd = &t->nodes[id];
        nd->cut_dim = -1; nd->first = first; nd->count = cnt;
        nd->cut_val = nd->lo_bound = nd->hi_bound = 0.f; nd->child[0] = nd->child[1] = -1;
        return id;
    }
    int cd = max_spread_dim(t, first, cnt);
    int32_t n_lo = cnt / 2;
    median_split(t, first, cnt, cd, n_lo);
    float cv = (PT(t, t->perm[first + n_lo - 1], cd) + PT(t, t->perm[first + n_lo], cd)) / 2.f;
    float lv = t->box_lo[cd], hv = t->box_hi[cd];
    t->box_hi[cd] = cv;
    int32_t lo = build_rec(t, first, n_lo);
    t->box_hi[cd] = hv;
    t->box_lo[cd] = cv;
    int32_t hi = build_rec(t, first + n_lo, cnt - n_lo);
    t->box_lo[cd] = lv;
    kd_node *nd = &t->nodes[id]; /* nodes may have been reallocated by the recursion */
    nd->cut_dim = cd; nd->cut_val = cv; nd->lo_bound = lv; nd->hi_bound = hv;
    nd->child[0] = lo; nd->child[1] = hi; nd->first = 0; nd->count = 0;
    return id;
}

struct tmo_kdtree *tmo_kdtree_create(const float *data, int64_t n, int dd, int bs) {
    struct tmo_kdtree *t = (struct tmo_kdtree *)calloc(1, sizeof *t);
    t->data = data; t->n = n; t->dd = dd; t->bs = bs < 1 ? 1 : bs;
    t->perm = (int32_t *)malloc(sizeof(int32_t) * (size_t)(n > 0 ? n : 1));
    for (int64_t i = 0; i < n; i++) t->perm[i] = (int32_t)i;
    t->box_lo = (float *)malloc(sizeof(float) * (size_t)dd);
    t->box_hi = (float *)malloc(sizeof(float) * (size_t)dd);
    for (int d = 0; d < dd; d++) {
        float mn = n ? data[d] : 0.f, mx = mn;
        for (int64_t i = 1; i < n; i++) {
            float v = data[i * dd + d];
            if (v < mn) mn = v;
            if (v > mx) mx = v;
        }
        t->box_lo[d] = mn; t->box_hi[d] = mx;
    }
    t->root = build_rec(t, 0, (int32_t)n);
    return t;
}

void tmo_kdtree_destroy(struct tmo_kdtree *t) {
    if (!t) return;
    free(t->perm); free(t->nodes); free(t->box_lo); free(t->box_hi); free(t);
}

typedef struct {
    const struct tmo_kdtree *t;
    const float *q;
    int k, cnt;
    float *bd;     /* ascending, k entries */
    int32_t *bi;
    int64_t visited; /* leaf points whose distance loop was entered */
    int64_t dims;    /* dimension steps executed in those loops */
} kd_search;

static inline float kth(const kd_search *s) { return s->cnt < s->k ? INFINITY : s->bd[s->k - 1]; }

static inline void kd_insert(kd_search *s, float d, int32_t i) {
    int p = s->cnt < s->k ? s->cnt++ : s->k - 1;
    while (p > 0 && d < s->bd[p - 1]) { s->bd[p] = s->bd[p - 1]; s->bi[p] = s->bi[p - 1]; p--; }
    s->bd[p] = d; s->bi[p] = i;
}

static void search_rec(kd_search *s, int32_t id, float box_dist) {
    const kd_node *nd = &s->t->nodes[id];
    if (nd->cut_dim < 0) {
        float min_dist = kth(s);
        const int dd = s->t->dd;
        for (int32_t e = 0; e < nd->count; e++) {
            int32_t pi = s->t->perm[nd->first + e];
            const float *p = s->t->data + (int64_t)pi * dd;
            float dist = 0.f;
            int d;
            for (d = 0; d < dd; d++) {
                float tt = s->q[d] - p[d];
                dist = dist + tt * tt;
                if (dist > min_dist) break;
            }
            s->visited++;
            s->dims += d < dd ? d + 1 : dd;
            if (d >= dd) { kd_insert(s, dist, pi); min_dist = kth(s); }
        }
        return;
    }
    float cut_diff = s->q[nd->cut_dim] - nd->cut_val;
    if (cut_diff < 0.f) {
        search_rec(s, nd->child[0], box_dist);
        float box_diff = nd->lo_bound - s->q[nd->cut_dim];
        if (box_diff < 0.f) box_diff = 0.f;
        box_dist = box_dist + (cut_diff * cut_diff - box_diff * box_diff);
        if (box_dist < kth(s)) search_rec(s, nd->child[1], box_dist); /* (1+eps)^2 = 1 */
    } else {
        search_rec(s, nd->child[1], box_dist);
        float box_diff = s->q[nd->cut_dim] - nd->hi_bound;
        if (box_diff < 0.f) box_diff = 0.f;
        box_dist = box_dist + (cut_diff * cut_diff - box_diff * box_diff);
        if (box_dist < kth(s)) search_rec(s, nd->child[0], box_dist);
    }
}

/* k-NN for a batch of queries, OpenMP over queries (the reference searches one handle from
 * many MTProcs threads, main.pas:972 -> 4027).  stats (may be NULL): [0] leaf points visited,
 * [1] dimension steps, summed over the batch. */
void tmo_kdtree_search(const struct tmo_kdtree *t, const float *q, int64_t nq, int k, int32_t *idx,
                       float *dist, int nthreads, int64_t *stats) {
    int64_t tot_v = 0, tot_d = 0;
#ifdef _OPENMP
    if (nthreads <= 0) nthreads = omp_get_max_threads();
#endif
#pragma omp parallel for schedule(dynamic, 16) num_threads(nthreads) reduction(+ : tot_v, tot_d)
    for (int64_t qi = 0; qi < nq; qi++) {
        float bd[64];
        int32_t bi[64];
        kd_search s;
        s.t = t; s.q = q + qi * t->dd; s.k = k; s.cnt = 0; s.bd = bd; s.bi = bi; s.visited = 0; s.dims = 0;
        float box_dist = 0.f; /* distance from q to the root bounding box */
        for (int d = 0; d < t->dd; d++) {
            float v = s.q[d];
            if (v < t->box_lo[d]) { float e = t->box_lo[d] - v; box_dist += e * e; }
            else if (v > t->box_hi[d]) { float e = v - t->box_hi[d]; box_dist += e * e; }
        }
        if (t->n > 0) search_rec(&s, t->root, box_dist);
        for (int j = 0; j < k; j++) {
            idx[qi * k + j] = j < s.cnt ? bi[j] : -1;
            dist[qi * k + j] = j < s.cnt ? bd[j] : 0.f;
        }
        tot_v += s.visited; tot_d += s.dims;
    }
    if (stats) { stats[0] = tot_v; stats[1] = tot_d; }
}
