// reindex.cu — the Reindex step that consumes FrameTiling's output (reference btnReindexClick
// main.pas:1199-1230 + ReindexTiles 4483-4527) on the device: use-count histogram of every tilemap
// item's GlobalTileIndex, inactive tiles dropped, active tiles ordered by (UseCount descending,
// old index ascending) (CompareTileUseCountRev 4472-4481: a strict total order, so the result does not
// depend on the sort algorithm), tilemap indices rewritten through the old -> new map.
// Integer work, bit-exact; HBM-bound: 4 B read + 4 B written per tilemap item, twice.
#include "tm_common.cuh"
#include <cub/device/device_radix_sort.cuh>

namespace {

__global__ void ri_hist_kernel(const int32_t *__restrict__ tm, int64_t n, int64_t ntiles, int32_t *__restrict__ count,
                               int *__restrict__ err) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const int32_t t = tm[i];
        if (t < 0 || t >= ntiles) { atomicOr(err, 1); continue; }
        atomicAdd(count + t, 1);
    }
}

__global__ void ri_keys_kernel(const int32_t *__restrict__ count, int64_t ntiles, unsigned long long *__restrict__ keys,
                               int *__restrict__ n_active) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= ntiles) return;
    const uint32_t c = (uint32_t)count[t];
    keys[t] = ((unsigned long long)(0xffffffffu - c) << 32) | (unsigned long long)t; // count desc, index asc
    if (c) atomicAdd(n_active, 1);
}

__global__ void ri_map_kernel(const unsigned long long *__restrict__ sorted, const int32_t *__restrict__ count,
                              int64_t ntiles, int32_t *__restrict__ idx_map, int32_t *__restrict__ order) {
    const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= ntiles) return;
    const int32_t old = (int32_t)(sorted[r] & 0xffffffffull);
    const bool active = count[old] != 0;
    idx_map[old] = active ? (int32_t)r : -1; // IdxMap is pre-filled with $ffffffff, main.pas:4510
    order[r] = active ? old : -1;
}

__global__ void ri_remap_kernel(int32_t *__restrict__ tm, int64_t n, const int32_t *__restrict__ idx_map) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) tm[i] = idx_map[tm[i]];
}

} // namespace

// d_tilemap[n] in/out; d_count[ntiles], d_idx_map[ntiles], d_order[ntiles] out; *h_active = tiles kept
int tm_reindex_dev_impl(int32_t *d_tilemap, int64_t n, int64_t ntiles, int32_t *d_count, int32_t *d_idx_map,
                        int32_t *d_order, int64_t *h_active, cudaStream_t s) {
    *h_active = 0;
    if (ntiles == 0) return n ? (tm_set_error("tm_reindex: tilemap items but no tiles"), TM_EINVAL) : 0;
    unsigned long long *keys = nullptr, *sorted = nullptr;
    int *flags = nullptr; // [0] error, [1] active count
    void *tmp = nullptr;
    size_t tmp_bytes = 0;
    int r = tm_check(cudaMalloc(&keys, sizeof(unsigned long long) * (size_t)ntiles), "cudaMalloc");
    if (!r) r = tm_check(cudaMalloc(&sorted, sizeof(unsigned long long) * (size_t)ntiles), "cudaMalloc");
    if (!r) r = tm_check(cudaMalloc(&flags, 2 * sizeof(int)), "cudaMalloc");
    if (!r) r = tm_check(cudaMemsetAsync(flags, 0, 2 * sizeof(int), s), "memset");
    if (!r) r = tm_check(cudaMemsetAsync(d_count, 0, sizeof(int32_t) * (size_t)ntiles, s), "memset");
    const unsigned gb = (unsigned)std::min<int64_t>((n + 255) / 256, (int64_t)tm_num_sms() * 16);
    if (!r && n) {
        ri_hist_kernel<<<gb, 256, 0, s>>>(d_tilemap, n, ntiles, d_count, flags);
        tm_count_launch();
    }
    if (!r) {
        ri_keys_kernel<<<(unsigned)((ntiles + 255) / 256), 256, 0, s>>>(d_count, ntiles, keys, flags + 1);
        tm_count_launch();
        cub::DeviceRadixSort::SortKeys(nullptr, tmp_bytes, keys, sorted, (int)ntiles, 0, 64, s);
        r = tm_check(cudaMalloc(&tmp, tmp_bytes ? tmp_bytes : 1), "cudaMalloc(sort)");
    }
    if (!r) r = tm_check(cub::DeviceRadixSort::SortKeys(tmp, tmp_bytes, keys, sorted, (int)ntiles, 0, 64, s), "SortKeys");
    int h_flags[2] = {0, 0};
    if (!r) {
        tm_count_launch(3);
        ri_map_kernel<<<(unsigned)((ntiles + 255) / 256), 256, 0, s>>>(sorted, d_count, ntiles, d_idx_map, d_order);
        tm_count_launch();
        r = tm_check(cudaMemcpyAsync(h_flags, flags, sizeof h_flags, cudaMemcpyDeviceToHost, s), "D2H");
    }
    if (!r) r = tm_check(cudaStreamSynchronize(s), "sync(reindex)");
    if (!r && h_flags[0]) { tm_set_error("tm_reindex: tilemap index outside [0, ntiles)"); r = TM_EINVAL; }
    if (!r && n) {
        ri_remap_kernel<<<gb, 256, 0, s>>>(d_tilemap, n, d_idx_map);
        tm_count_launch();
        r = tm_check(cudaGetLastError(), "ri_remap_kernel");
    }
    if (!r) r = tm_check(cudaStreamSynchronize(s), "sync(reindex)"); // scratch is freed below
    cudaFree(tmp); cudaFree(flags); cudaFree(sorted); cudaFree(keys);
    if (!r) *h_active = h_flags[1];
    return r;
}
