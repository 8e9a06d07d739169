// host_threads_bench.cpp — the encoder's call patterns against libtmann.so from NATIVE threads (no Python, no GIL):
//   * one frame per tm_frametiling_search call from T worker threads (DoFrameTiling, main.pas:972, 4005-4038),
//     pageable caller memory, like Pascal dynamic arrays;
//   * one ann_kdtree_search per tilemap position from T threads (the unchanged encoder, main.pas:4027).
// Prints one JSON line per case.  Build: g++ -O2 -std=c++17 scripts/host_threads_bench.cpp -o build/host_threads_bench -ldl -lpthread
// Run:   build/host_threads_bench tiler_b200/libtmann.so [tiles=16384] [positions=14400] [devices=1]
#include <dlfcn.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <atomic>
#include <chrono>
#include <thread>
#include <vector>

struct ANNkdtree;
typedef ANNkdtree *(*create_fn)(const uint8_t *, int64_t, const int32_t *, int, int, const uint8_t *, int, int);
typedef int (*ft_fn)(ANNkdtree *, const int32_t *, int64_t, int, int, int *, float *);
typedef int (*search_fn)(ANNkdtree *, float *, float, float *);
typedef int (*psyv_fn)(const int32_t *, int64_t, int, int, float *);
typedef void (*destroy_fn)(ANNkdtree *);
typedef int (*setdev_fn)(const int *, int);
typedef const char *(*err_fn)(void);

static uint32_t rnd(uint32_t range, uint32_t &seed) { // RandInt, kmodes.pas:82-86
    seed = seed * 0x08088405u + 1u;
    return (uint32_t)(((uint64_t)seed * range) >> 32);
}

int main(int argc, char **argv) {
    if (argc < 2) { fprintf(stderr, "usage: %s libtmann.so [tiles] [positions] [devices]\n", argv[0]); return 2; }
    const int ntiles = argc > 2 ? atoi(argv[2]) : 16384, pos = argc > 3 ? atoi(argv[3]) : 14400, ndev = argc > 4 ? atoi(argv[4]) : 1;
    void *lib = dlopen(argv[1], RTLD_NOW);
    if (!lib) { fprintf(stderr, "dlopen: %s\n", dlerror()); return 2; }
    auto create = (create_fn)dlsym(lib, "tm_tileset_dataset_create");
    auto ft = (ft_fn)dlsym(lib, "tm_frametiling_search");
    auto search = (search_fn)dlsym(lib, "ann_kdtree_search");
    auto psyv = (psyv_fn)dlsym(lib, "tm_psyv_queries");
    auto destroy = (destroy_fn)dlsym(lib, "ann_kdtree_destroy");
    auto setdev = (setdev_fn)dlsym(lib, "tm_set_devices");
    auto lasterr = (err_fn)dlsym(lib, "tm_last_error");
    if (ndev > 1) {
        std::vector<int> d(ndev);
        for (int i = 0; i < ndev; i++) d[i] = i;
        if (setdev(d.data(), ndev)) { fprintf(stderr, "tm_set_devices: %s\n", lasterr()); return 1; }
    }
    uint32_t seed = 0x42381337u;
    int32_t pal[16];
    for (int i = 0; i < 16; i++) pal[i] = (int32_t)rnd(1u << 24, seed);
    std::vector<uint8_t> tiles((size_t)ntiles * 64);
    for (int t = 0; t < ntiles; t++) {
        const int base = rnd(16, seed), gx = (int)rnd(7, seed) - 3, gy = (int)rnd(7, seed) - 3, div = 2 + rnd(6, seed);
        for (int y = 0; y < 8; y++)
            for (int x = 0; x < 8; x++) {
                int v = base + (gx * (x - 4) + gy * (y - 4)) / div + (int)rnd(3, seed) - 1;
                tiles[(size_t)t * 64 + y * 8 + x] = (uint8_t)(v < 0 ? 0 : v > 15 ? 15 : v);
            }
    }
    const int nframes = 64;
    std::vector<int32_t> rgb((size_t)nframes * pos * 64);
    for (size_t q = 0; q < (size_t)nframes * pos; q++) {
        const int t = rnd(ntiles, seed);
        for (int i = 0; i < 64; i++) {
            const int32_t c = pal[tiles[(size_t)t * 64 + i]];
            int ch[3];
            for (int k = 0; k < 3; k++) {
                int v = ((c >> (8 * k)) & 255) + (int)rnd(13, seed) - 6;
                ch[k] = v < 0 ? 0 : v > 255 ? 255 : v;
            }
            rgb[q * 64 + i] = (ch[2] << 16) | (ch[1] << 8) | ch[0];
        }
    }
    ANNkdtree *tree = create(tiles.data(), ntiles, pal, 1, 16, nullptr, 1, -1);
    if (!tree) { fprintf(stderr, "create: %s\n", lasterr()); return 1; }
    std::vector<int> idx((size_t)nframes * pos);
    std::vector<float> err((size_t)nframes * pos);
    using clk = std::chrono::steady_clock;
    // persistent worker threads (an encoder's thread pool): every thread makes untimed calls first, the clock runs
    // between two barriers
    struct Barrier {
        std::atomic<int> count{0}, gen{0};
        int n;
        explicit Barrier(int n_) : n(n_) {}
        void wait() {
            const int g = gen.load();
            if (count.fetch_add(1) + 1 == n) { count.store(0); gen.fetch_add(1); }
            else while (gen.load() == g) std::this_thread::yield();
        }
    };
    for (int T : {1, 2, 4, 8, 16}) {
        const int rounds = T == 1 ? 2 : 12;
        std::atomic<int> bad{0};
        Barrier bar(T + 1);
        auto body = [&](int w, int reps) {
            for (int r = 0; r < reps; r++)
                for (int f = w; f < nframes; f += T)
                    if (ft(tree, rgb.data() + (size_t)f * pos * 64, pos, 1, -1, idx.data() + (size_t)f * pos, err.data() + (size_t)f * pos)) bad++;
        };
        std::vector<std::thread> th;
        for (int w = 0; w < T; w++)
            th.emplace_back([&, w] {
                body(w, 2); // warm: this thread's staging, the worker's scratch at the batch sizes this thread count produces
                bar.wait();
                body(w, rounds);
                bar.wait();
            });
        bar.wait();
        const auto t0 = clk::now();
        bar.wait();
        const double dt = std::chrono::duration<double>(clk::now() - t0).count();
        for (auto &t : th) t.join();
        const double calls = (double)nframes * rounds;
        printf("{\"case\": \"frame_calls\", \"threads\": %d, \"devices\": %d, \"queries_per_call\": %d, \"rows\": %d, \"calls\": %.0f, "
               "\"ms_per_call\": %.4f, \"queries_per_s\": %.0f, \"errors\": %d}\n",
               T, ndev, pos, ntiles * 4, calls, 1e3 * dt / calls, calls * pos / dt, bad.load());
        fflush(stdout);
    }
    // single-query calls: descriptors of the first 4096 query tiles
    const int nq1 = 4096;
    std::vector<float> desc((size_t)nq1 * 192);
    if (psyv(rgb.data(), nq1, 1, -1, desc.data())) { fprintf(stderr, "psyv: %s\n", lasterr()); return 1; }
    for (int T : {1, 16, 32}) {
        const int per = T == 1 ? 512 : 1024;
        std::atomic<int> bad{0};
        Barrier bar(T + 1);
        auto body = [&](int w, int n) {
            float e;
            for (int j = 0; j < n; j++)
                if (search(tree, desc.data() + (size_t)((w * 131 + j) % nq1) * 192, 0.f, &e) < 0) bad++;
        };
        std::vector<std::thread> th;
        for (int w = 0; w < T; w++)
            th.emplace_back([&, w] {
                body(w, 16);
                bar.wait();
                body(w, per);
                bar.wait();
            });
        bar.wait();
        const auto t0 = clk::now();
        bar.wait();
        const double dt = std::chrono::duration<double>(clk::now() - t0).count();
        for (auto &t : th) t.join();
        printf("{\"case\": \"single_query_calls\", \"threads\": %d, \"devices\": %d, \"rows\": %d, \"calls\": %d, \"calls_per_s\": %.0f, "
               "\"us_per_call\": %.2f, \"errors\": %d}\n",
               T, ndev, ntiles * 4, T * per, T * per / dt, 1e6 * dt / (T * per), bad.load());
        fflush(stdout);
    }
    destroy(tree);
    return 0;
}
