// keyframe.cu — PrepareFrameTiling (reference main.pas:3791-3967) on the device.
//
// UseOne (main.pas:3802-3853): for every distinct (PalIdx, GlobalTileIndex) of the keyframe's
// tilemaps, the 8 nearest stage-1 rows of the tile's own palette-index line (the k = 8 search is
// done once per stored tile by the tensor-core engine and cached on the stage-1 handle); hits
// whose err equals the previous hit's are skipped (3835-3837); used[pal'][tile'][attrs'] is set for
// the palettes the quality mode selects (3841-3851).  Then the used table is compacted in the row
// order of DoPsyV (pal, tile, vmir, hmir; 3902-3905) into the row list the descriptor kernel renders.
// Integer / byte work, bit-exact; equal distances are listed by ascending row index (ANN's tie order
// is traversal-dependent and unspecified).
#include "tm_common.cuh"
#include <cub/device/device_scan.cuh>
#include <cub/iterator/transform_input_iterator.cuh>

namespace {

__global__ void kf_mark_kernel(const int32_t *__restrict__ item_pal, const int32_t *__restrict__ item_tile,
                               int64_t nitems, int64_t ntiles, int npal, const int *__restrict__ knn_idx,
                               const float *__restrict__ knn_err, const int32_t *__restrict__ tr_tile,
                               const uint8_t *__restrict__ palmask,
                               uint32_t *__restrict__ seen, uint8_t *__restrict__ used, int *__restrict__ bad) {
    const int64_t it = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (it >= nitems) return;
    const int pal = item_pal[it];
    const int64_t tile = item_tile[it];
    if (pal < 0 || pal >= npal || tile < 0 || tile >= ntiles) { atomicOr(bad, 1); return; }
    // "already seen (pal, tile)" test of main.pas:3814-3821 (a spin lock there, an atomic bit here)
    const int64_t bit = (int64_t)pal * ntiles + tile;
    const uint32_t m = 1u << (bit & 31);
    if (atomicOr(seen + (bit >> 5), m) & m) return;
    const uint8_t *mask = palmask + (int64_t)pal * npal;
    float last = INFINITY;
    for (int i = 0; i < 8; i++) {
        const int idx = knn_idx[tile * 8 + i];
        if (idx < 0) break; // fewer than 8 stage-1 rows
        const float e = knn_err[tile * 8 + i];
        if (e == last) continue; // main.pas:3835-3837
        last = e;
        // stage-1 rows exist for Active tiles only (PrepareGlobalFT main.pas:3763-3777): TRToTileIdx maps back
        const int64_t t2 = tr_tile ? tr_tile[idx >> 2] : idx >> 2;
        const int a2 = (0x2310 >> ((idx & 3) * 4)) & 3; // row order none,H,HV,V -> attrs 0,1,3,2 (main.pas:3753)
        for (int p = 0; p < npal; p++)
            if (mask[p]) used[((int64_t)p * ntiles + t2) * 4 + a2] = 1;
    }
}

// UseOne's query line: the stored tile's 64 palette indices as floats (main.pas:3823-3828)
__global__ void kf_lines_kernel(const uint8_t *__restrict__ palpix, int64_t n, float *__restrict__ out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = (float)palpix[i];
}

struct ByteToInt {
    __host__ __device__ __forceinline__ int operator()(const uint8_t &v) const { return v ? 1 : 0; }
};

__global__ void kf_scatter_kernel(const uint8_t *__restrict__ used, const int *__restrict__ pos, int64_t n,
                                  int64_t *__restrict__ row_map) {
    const int64_t u = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (u < n && used[u]) row_map[pos[u]] = u;
}

__global__ void kf_tables_kernel(const int64_t *__restrict__ row_map, int64_t nrows, int64_t ntiles,
                                 int32_t *__restrict__ tr_tile, int32_t *__restrict__ tr_pal,
                                 uint8_t *__restrict__ tr_attrs) {
    const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= nrows) return;
    const int64_t u = row_map[r];
    tr_tile[r] = (int32_t)((u >> 2) % ntiles);           // main.pas:3908
    tr_pal[r] = (int32_t)((u >> 2) / ntiles);            // main.pas:3909
    tr_attrs[r] = (uint8_t)(u & 3);                      // Ord(hmir) or Ord(vmir) shl 1, main.pas:3910
}

} // namespace

int tm_launch_kf_lines(const uint8_t *d_palpix, int64_t ntiles, float *d_lines, cudaStream_t s) {
    if (ntiles == 0) return 0;
    const int64_t n = ntiles * 64;
    kf_lines_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(d_palpix, n, d_lines);
    tm_count_launch();
    return tm_check(cudaGetLastError(), "kf_lines_kernel");
}

int tm_launch_kf_mark(const int32_t *d_item_pal, const int32_t *d_item_tile, int64_t nitems, int64_t ntiles, int npal,
                      const int *d_knn_idx, const float *d_knn_err, const int32_t *d_tr_tile, const uint8_t *d_palmask,
                      uint32_t *d_seen, uint8_t *d_used, int *d_bad, cudaStream_t s) {
    if (nitems == 0) return 0;
    kf_mark_kernel<<<(unsigned)((nitems + 255) / 256), 256, 0, s>>>(d_item_pal, d_item_tile, nitems, ntiles, npal, d_knn_idx,
                                                                   d_knn_err, d_tr_tile, d_palmask, d_seen, d_used, d_bad);
    tm_count_launch();
    return tm_check(cudaGetLastError(), "kf_mark_kernel");
}

int tm_kf_compact(const uint8_t *d_used, int64_t n, int64_t **d_row_map_out, int64_t *h_count, cudaStream_t s) {
    *d_row_map_out = nullptr;
    *h_count = 0;
    if (n == 0) return 0;
    if (n > 0x7fffffff) { tm_set_error("used table too large"); return TM_EINVAL; }
    int *pos = nullptr;
    void *tmp = nullptr;
    size_t tmp_bytes = 0;
    TM_CUDA(tm_dev_malloc(&pos, sizeof(int) * (size_t)n));
    cub::TransformInputIterator<int, ByteToInt, const uint8_t *> in(d_used, ByteToInt());
    cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, in, pos, (int)n, s);
    int r = tm_check(tm_dev_malloc(&tmp, tmp_bytes ? tmp_bytes : 1), "tm_dev_malloc(scan)");
    if (!r) r = tm_check(cub::DeviceScan::ExclusiveSum(tmp, tmp_bytes, in, pos, (int)n, s), "ExclusiveSum");
    tm_count_launch(2);
    int last_pos = 0;
    uint8_t last_flag = 0;
    if (!r) r = tm_check(cudaMemcpyAsync(&last_pos, pos + (n - 1), sizeof(int), cudaMemcpyDeviceToHost, s), "D2H count");
    if (!r) r = tm_check(cudaMemcpyAsync(&last_flag, d_used + (n - 1), 1, cudaMemcpyDeviceToHost, s), "D2H flag");
    if (!r) r = tm_check(cudaStreamSynchronize(s), "sync(compact)");
    const int64_t cnt = (int64_t)last_pos + (last_flag ? 1 : 0);
    int64_t *row_map = nullptr;
    if (!r) r = tm_check(tm_dev_malloc(&row_map, sizeof(int64_t) * (size_t)(cnt ? cnt : 1)), "tm_dev_malloc(row map)");
    if (!r && cnt) {
        kf_scatter_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(d_used, pos, n, row_map);
        tm_count_launch();
        r = tm_check(cudaGetLastError(), "kf_scatter_kernel");
        if (!r) r = tm_check(cudaStreamSynchronize(s), "sync(scatter)"); // pos / tmp are freed below
    }
    tm_dev_free(tmp);
    tm_dev_free(pos);
    if (r) { tm_dev_free(row_map); return r; }
    *d_row_map_out = row_map;
    *h_count = cnt;
    return 0;
}

int tm_launch_kf_tables(const int64_t *d_row_map, int64_t nrows, int64_t ntiles, int32_t *d_tr_tile, int32_t *d_tr_pal,
                        uint8_t *d_tr_attrs, cudaStream_t s) {
    if (nrows == 0) return 0;
    kf_tables_kernel<<<(unsigned)((nrows + 255) / 256), 256, 0, s>>>(d_row_map, nrows, ntiles, d_tr_tile, d_tr_pal, d_tr_attrs);
    tm_count_launch();
    return tm_check(cudaGetLastError(), "kf_tables_kernel");
}
