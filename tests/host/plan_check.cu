// Host-side check of the coarse pass's work-item plan (tiler_b200/csrc/search_tc.h: TcPlan), compiled and run by
// tests/test_plan_host.py with nvcc as plain host code: every (query block, row split) is served exactly once, the
// tile ranges of a block's splits partition the tileset, the candidate lists of different (query, split, column half)
// never overlap, and a chained plan runs split-major with one list per column half.
#include "../../tiler_b200/csrc/search_tc.h"
#include <cstdio>
#include <cstdlib>
#include <map>
#include <set>
#include <vector>

static int fails = 0;
#define CHECK(c, ...) do { if (!(c)) { fails++; std::printf("FAIL %s:%d: ", __FILE__, __LINE__); std::printf(__VA_ARGS__); std::printf("\n"); } } while (0)

static void check_plan(const TcPlan &p, const char *name) {
    const int items = p.items();
    std::map<std::pair<int, int>, int> seen; // (qb, sp) -> item
    std::vector<std::vector<std::pair<int, int>>> ranges(p.qblocks);
    int last_sp = -1;
    for (int it = 0; it < items; it++) {
        int qb, sp, t0, t1;
        p.decode(it, qb, sp, t0, t1);
        CHECK(qb >= 0 && qb < p.qblocks, "%s: item %d block %d", name, it, qb);
        CHECK(t0 >= 0 && t0 < t1 && t1 <= p.total_tiles, "%s: item %d tiles [%d, %d)", name, it, t0, t1);
        CHECK(seen.insert({{qb, sp}, it}).second, "%s: (block %d, split %d) served twice", name, qb, sp);
        ranges[qb].push_back({t0, t1});
        if (p.chain) { // split-major, blocks ascending inside a split; the predecessor is exactly qblocks items earlier
            CHECK(sp >= last_sp, "%s: item %d goes back to split %d", name, it, sp);
            CHECK(it == sp * p.qblocks + qb, "%s: item %d is not (split %d, block %d) in split-major order", name, it, sp, qb);
            last_sp = sp;
        }
    }
    for (int qb = 0; qb < p.qblocks; qb++) { // a block's splits partition [0, total_tiles)
        int covered = 0, prev_end = 0;
        for (auto r : ranges[qb]) { CHECK(r.first == prev_end, "%s: block %d gap at tile %d", name, qb, prev_end); prev_end = r.second; covered += r.second - r.first; }
        CHECK(covered == p.total_tiles && prev_end == p.total_tiles, "%s: block %d covers %d of %d tiles", name, qb, covered, p.total_tiles);
    }
    // candidate lists: the kernel writes list_base(q) + (chain ? 0 : sp * 2) + h; rerank reads nlists(q) from list_base(q)
    std::set<int64_t> used;
    for (int qb = 0; qb < p.qblocks; qb++)
        for (int64_t q : {(int64_t)qb * 256, (int64_t)qb * 256 + 255}) {
            const int nsp = (int)ranges[qb].size();
            CHECK(p.nlists(q) == (p.chain ? 2 : 2 * nsp), "%s: query %lld has %d lists for %d splits", name, (long long)q, p.nlists(q), nsp);
            CHECK(p.nlists(q) <= p.max_nlists(), "%s: max_nlists", name);
            for (int sp = 0; sp < nsp; sp++)
                for (int h = 0; h < 2; h++) {
                    const int64_t l = p.list_base(q) + (p.chain ? 0 : sp * 2) + h;
                    CHECK(l >= p.list_base(q) && l < p.list_base(q) + p.nlists(q), "%s: list outside the query's range", name);
                    CHECK(l < p.total_lists(), "%s: list %lld beyond total %lld", name, (long long)l, (long long)p.total_lists());
                    if (!p.chain || sp == 0) CHECK(used.insert(l).second, "%s: list %lld used twice", name, (long long)l);
                }
        }
    CHECK(p.list_base((int64_t)p.qblocks * 256) == p.total_lists(), "%s: total_lists", name);
}

int main() {
    // one item per block; uniform split; tail plan (225 blocks on 74 pairs, tail of 3 blocks in 8 finer splits); chained
    check_plan(TcPlan{300, 300, 1, 256, 1, 256, 256, 0}, "single");
    check_plan(TcPlan{40, 40, 4, 64, 4, 64, 256, 0}, "split4");
    check_plan(TcPlan{40, 40, 3, 86, 3, 86, 256, 0}, "split3 ragged");
    check_plan(TcPlan{225, 222, 1, 256, 8, 32, 256, 0}, "tail");
    check_plan(TcPlan{225, 222, 1, 1000, 7, 143, 1000, 0}, "tail ragged");
    check_plan(TcPlan{4050, 4050, 8, 512, 8, 512, 4096, 1}, "chain c5");
    check_plan(TcPlan{600, 600, 3, 342, 3, 342, 1024, 1}, "chain c3 ragged");
    std::printf(fails ? "%d failures\n" : "ok\n", fails);
    return fails ? 1 : 0;
}
