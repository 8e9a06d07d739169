// psyv.cu — descriptor pack/mirror kernels (sm_100a).
//
// Replaces ComputeTilePsyVisFeatures (reference main.pas:2997-3177) for the two
// producers on the FrameTiling path: query tiles from true-colour pixels
// (main.pas:4023) and tileset rows rendered from a palette in the four H/V
// orientations (DoPsyV main.pas:3883-3919), plus the stage-1 palette-index rows
// of PrepareGlobalFT (main.pas:3736-3780).
//
// The arithmetic is FP64 in the reference's operation order, compiled with
// -fmad=false, so the float32 rows written here are bit-identical to what the
// Pascal code hands to ANN.  One thread owns one (tile, colour plane): 64 doubles
// in registers, the 3-level Haar (WaveletGS main.pas:2805-2840) fully unrolled.
// HBM-bound by design: 256 B in, 768 B out per query tile.
#include "tm_common.cuh"
#include <math.h>
#include <cuda_fp16.h>

namespace {

struct PsyvTables {      // host-built, mirrors InitLuts main.pas:592-623
    double f[256];       // gGammaCorLut[g][c] (c/255 or pow)
    double fr[256];      // cRedMul * f
    double fg[256];      // cGreenMul * f
    double fb[256];      // cBlueMul * f
};

struct PsyvConsts {
    double cu, cv, luma_div, luma_rcp, haar, rcp255;
};


// One Haar level on the top-left N x N of an 8 x 8 (stride 8) block; same
// operation order as WaveletGS: rows first into tx, then columns.
template <int N>
__device__ __forceinline__ void haar_level(double (&a)[64], double f) {
    // in place, one row (then one column) of temporaries at a time: 64 + N live doubles
#pragma unroll
    for (int y = 0; y < N; y++) {
        double t[N];
#pragma unroll
        for (int x = 0; x < N / 2; x++) {
            double p = a[y * 8 + 2 * x], q = a[y * 8 + 2 * x + 1];
            t[x] = (p + q) * f;
            t[x + N / 2] = (p - q) * f;
        }
#pragma unroll
        for (int x = 0; x < N; x++) a[y * 8 + x] = t[x];
    }
#pragma unroll
    for (int x = 0; x < N; x++) {
        double t[N];
#pragma unroll
        for (int y = 0; y < N / 2; y++) {
            double p = a[(2 * y) * 8 + x], q = a[(2 * y + 1) * 8 + x];
            t[y] = (p + q) * f;
            t[y + N / 2] = (p - q) * f;
        }
#pragma unroll
        for (int y = 0; y < N; y++) a[y * 8 + x] = t[y];
    }
}

__device__ __forceinline__ double plane_value(int32_t col, int plane, const PsyvTables *T,
                                              const PsyvConsts &K) {
    int r = col & 0xff, g = (col >> 8) & 0xff, b = (col >> 16) & 0xff; // FromRGB main.pas:571-576
    // main.pas:2674 divides by cLumaDiv = 10000.  x / d is formed as q = RN(x * RN(1/d)), r = x - q*d (exact,
    // fma), RN(q + r * RN(1/d)): three FP64 operations instead of the ~15-instruction division sequence,
    // and bit-identical to the IEEE quotient for every value the three LUT sums can take (all 3 x 2^24
    // (r, g, b, gamma) combinations were checked exhaustively on the host against x / 10000.0).
    const double x = (T->fr[r] + T->fg[g]) + T->fb[b];
    const double q0 = x * K.luma_rcp;
    const double y = fma(fma(-q0, K.luma_div, x), K.luma_rcp, q0);
    if (plane == 0) return y;
    if (plane == 1) return (T->f[b] - y) * K.cu;                         // main.pas:2675
    return (T->f[r] - y) * K.cv;                                         // main.pas:2676
}

// DCT-II branch main.pas:3075-3174: z = sum_{y,x} cpn*lut (sequential), optional Q weight, * cUVRatio
__device__ __forceinline__ void dct_plane(const double (&a)[64], const double *lut, const double *qw,
                                          int plane, float *dst) {
    for (int vu = 0; vu < 64; vu++) {
        const double *l = lut + vu * 64;
        double z = 0.0;
#pragma unroll
        for (int i = 0; i < 64; i++) z += a[i] * l[i];
        if (qw) z *= qw[plane * 64 + vu];
        dst[vu] = (float)(z * lut[4096 + vu]);
    }
}

// Query tiles, two phases per block of 32 tiles (96 threads):
//   1. pixel phase: consecutive threads take consecutive pixels straight from global memory (coalesced), compute
//      (y, u, v) ONCE per pixel and park them as doubles in shared memory.  With GammaCor = -1 the colour LUT
//      gGammaCorLut[c] = c / 255.0 (main.pas:596-601) is evaluated arithmetically (exact 3-operation division,
//      checked against c / 255.0 for all 256 values when the tables are built), which removes the random-index
//      shared-memory reads whose bank conflicts bounded the one-phase kernel (11 LUT reads per pixel, ~5-way).
//   2. plane phase: thread = (tile, plane) pulls its 64 doubles into registers (16-byte reads, row stride 66
//      doubles = 4 banks apart -> conflict free), runs the Haar / DCT, and stages the float row in the same
//      shared memory (dead by then) for a coalesced 16-byte store.
constexpr int PQ_TILES = 32;
constexpr int PQ_THREADS = 96;   // DCT branch: thread = (tile, plane)
constexpr int PQ_THREADS_W = 192; // wavelet branch: thread = (tile, plane, row half), see haar_half below
__host__ __device__ constexpr int pq_threads(bool wavelets, bool split) { return wavelets && split ? PQ_THREADS_W : PQ_THREADS; }
constexpr int PQ_CPN_STRIDE = 66;  // doubles per (plane, tile) row
constexpr int PQ_OUT_STRIDE = 196; // words per staged output row (192 + 4)
constexpr size_t PQ_CPN_BYTES = (size_t)3 * PQ_TILES * PQ_CPN_STRIDE * 8;
constexpr size_t PQ_SMEM_LINEAR = PQ_CPN_BYTES;
constexpr size_t PQ_SMEM_LUT = PQ_CPN_BYTES + 256 * 8;
static_assert((size_t)PQ_TILES * PQ_OUT_STRIDE * 4 + (size_t)PQ_THREADS_W * 16 <= PQ_CPN_BYTES, "output staging + the halves' exchange area alias the component buffer");

// Optional second output of the query kernel: the tensor-core engine's centred fp16 operand and the two norms
// the certificate needs (what search_tc.cu's operand_kernel derives from the float rows), written from the
// staged rows so the 768 B / query are not read back from HBM.  q16 == nullptr: not requested.
struct PsyvOperandOut {
    const float *mean; // dataset column means [192]
    __half *q16;       // [nq][192]
    float *qn16;       // sum of squares of the rounded values
    float *qne;        // Euclidean norm of the centred float row
    float *qde;        // Euclidean norm of the fp16 rounding error of the centred row
    const int *dst;    // optional: input tile i goes to output row dst[i] (coherence order, qorder.cu)
};

template <bool LINEAR>
__device__ __forceinline__ double channel_value(int c, const double *ftab, double rcp255) {
    if (LINEAR) { // c / 255.0: q = RN(c * RN(1/255)), r = c - 255 q (exact), RN(q + r * RN(1/255))
        const double cd = (double)c;
        const double q0 = cd * rcp255;
        return fma(fma(-q0, 255.0, cd), rcp255, q0);
    }
    return ftab[c];
}

__device__ __forceinline__ void store_row_smem(float *dst, const double (&a)[64]) {
    float4 *d4 = reinterpret_cast<float4 *>(dst);
#pragma unroll
    for (int i = 0; i < 16; i++)
        d4[i] = make_float4((float)a[4 * i], (float)a[4 * i + 1], (float)a[4 * i + 2], (float)a[4 * i + 3]);
}

// The 3-level Haar of one 8 x 8 plane split over TWO threads by row halves (hf = 0: input rows 0-3, hf = 1: rows 4-7),
// 32 doubles each instead of 64: ~100 registers instead of 168, i.e. 18-24 warps per SM instead of 12 for a kernel whose
// pipes are all under half busy at 12.  Same operations on the same operands in the same order as
// haar_level (WaveletGS, main.pas:2805-2840) - every output is one (p + q) * f or (p - q) * f of the same p, q - so the
// rows stay bit-identical.  A row pass is local to a row; a column pass pairs rows (2y, 2y + 1), which one thread owns
// at level 8 (its rows 0-1 and 2-3) and at level 4 (the two low rows it has just produced); only the level-2 column
// pass pairs a row of one half with a row of the other: two values change hands through shared memory.
// After the call, with S = the thread's four 8-wide rows:  S0[0..3] = coefficient row hf, S0[4..7] = row 2 hf (cols 4-7),
// S1[0..3] = row 2 + hf, S1[4..7] = row 2 hf + 1 (cols 4-7), S2 = row 4 + 2 hf, S3 = row 5 + 2 hf.
__device__ __forceinline__ void haar_rows(double (&r)[8], int n, double f) {
    // n = 8 or 4 or 2 leading values of the row, in place
    double t[8];
#pragma unroll
    for (int x = 0; x < 4; x++)
        if (x < n / 2) {
            const double p = r[2 * x], q = r[2 * x + 1];
            t[x] = (p + q) * f;
            t[x + n / 2] = (p - q) * f;
        }
#pragma unroll
    for (int x = 0; x < 8; x++)
        if (x < n) r[x] = t[x];
}
__device__ __forceinline__ void haar_half(double (&S)[4][8], int hf, double f, double *xch_mine, const double *xch_other) {
    // level 8: rows, then the column pairs (S0, S1) and (S2, S3); low rows to S0, S1, high rows to S2, S3
#pragma unroll
    for (int y = 0; y < 4; y++) haar_rows(S[y], 8, f);
#pragma unroll
    for (int x = 0; x < 8; x++) {
        const double p0 = S[0][x], q0 = S[1][x], p1 = S[2][x], q1 = S[3][x];
        S[0][x] = (p0 + q0) * f;
        S[2][x] = (p0 - q0) * f;
        S[1][x] = (p1 + q1) * f;
        S[3][x] = (p1 - q1) * f;
    }
    // level 4 on the low rows' first four columns: rows, then the column pair (S0, S1)
    haar_rows(S[0], 4, f);
    haar_rows(S[1], 4, f);
#pragma unroll
    for (int x = 0; x < 4; x++) {
        const double p = S[0][x], q = S[1][x];
        S[0][x] = (p + q) * f;
        S[1][x] = (p - q) * f;
    }
    // level 2 on coefficient rows 0 (half 0's S0) and 1 (half 1's S0), columns 0-1: row pass, exchange, column pass
    haar_rows(S[0], 2, f);
    xch_mine[0] = S[0][0];
    xch_mine[1] = S[0][1];
    __syncthreads();
    const double o0 = xch_other[0], o1 = xch_other[1];
    if (hf == 0) { // p = row 0 (mine), q = row 1 (the other half's): the low row stays here
        S[0][0] = (S[0][0] + o0) * f;
        S[0][1] = (S[0][1] + o1) * f;
    } else {       // p = row 0 (the other half's), q = row 1 (mine): the high row stays here
        S[0][0] = (o0 - S[0][0]) * f;
        S[0][1] = (o1 - S[0][1]) * f;
    }
}

// SPLIT: the two-threads-per-plane wavelet path.  Used when the kernel also writes the tensor-core operand (the fused
// variant the FrameTiling path runs: 24 instead of 12 warps per SM hide the operand phase, descriptor phase of a C2 step
// 3.77 -> 3.41 ms); the plain descriptor call keeps one thread per plane (0.69 ms per 2 M tiles against 0.90 ms split:
// alone, the extra index arithmetic of twice the threads costs more than the occupancy buys).
template <bool WAVELETS, bool LINEAR, bool SPLIT>
__global__ void __launch_bounds__(pq_threads(WAVELETS, SPLIT)) psyv_queries_kernel(const int32_t *__restrict__ rgb, int64_t nq,
                                                                   const PsyvTables *__restrict__ tabs,
                                                                   PsyvConsts K, const double *__restrict__ lut,
                                                                   const double *__restrict__ qw,
                                                                   float *__restrict__ out, PsyvOperandOut op) {
    extern __shared__ __align__(16) uint8_t psm[];
    double *cpn = reinterpret_cast<double *>(psm);
    double *ftab = reinterpret_cast<double *>(psm + PQ_CPN_BYTES);
    float *sout = reinterpret_cast<float *>(psm);
    constexpr int NT = pq_threads(WAVELETS, SPLIT);
    double *xch = reinterpret_cast<double *>(psm + (size_t)PQ_TILES * PQ_OUT_STRIDE * 4); // behind the output staging
    __shared__ int64_t dst_row[PQ_TILES]; // output row of each of the block's tiles
    const int64_t q0 = (int64_t)blockIdx.x * PQ_TILES;
    const int ntile = (int)min((int64_t)PQ_TILES, nq - q0);
    if (!LINEAR) {
        for (int i = threadIdx.x; i < 256; i += blockDim.x) ftab[i] = tabs->f[i];
        __syncthreads();
    }
    {   // pixel phase (FromRGB main.pas:571-576, RGBToYUV main.pas:2662-2677); all of a thread's pixels are
        // requested before the first one is used (22 independent loads in flight per thread)
        const int32_t *src = rgb + q0 * 64;
        constexpr int PER_THREAD = (PQ_TILES * 64 + NT - 1) / NT;
        int32_t colv[PER_THREAD];
#pragma unroll
        for (int k = 0; k < PER_THREAD; k++) {
            const int e = threadIdx.x + k * NT;
            colv[k] = e < ntile * 64 ? __ldg(src + e) : 0;
        }
        // destination rows: requested together with the pixels (no dependent load), used only by the stores
        if ((int)threadIdx.x < PQ_TILES)
            dst_row[threadIdx.x] = (int)threadIdx.x < ntile ? (op.dst ? (int64_t)__ldg(op.dst + q0 + threadIdx.x) : q0 + threadIdx.x) : 0;
#pragma unroll
        for (int k = 0; k < PER_THREAD; k++) {
            const int e = threadIdx.x + k * NT;
            if (e < ntile * 64) {
                const int32_t col = colv[k];
                const double fr = channel_value<LINEAR>(col & 0xff, ftab, K.rcp255);
                const double fg = channel_value<LINEAR>((col >> 8) & 0xff, ftab, K.rcp255);
                const double fb = channel_value<LINEAR>((col >> 16) & 0xff, ftab, K.rcp255);
                const double x = (2126.0 * fr + 7152.0 * fg) + 722.0 * fb;
                const double qd = x * K.luma_rcp; // x / cLumaDiv, see plane_value
                const double y = fma(fma(-qd, K.luma_div, x), K.luma_rcp, qd);
                double *dst = cpn + (e >> 6) * PQ_CPN_STRIDE + (e & 63);
                dst[0] = y;
                dst[PQ_TILES * PQ_CPN_STRIDE] = (fb - y) * K.cu;
                dst[2 * PQ_TILES * PQ_CPN_STRIDE] = (fr - y) * K.cv;
            }
        }
    }
    __syncthreads();
    if (WAVELETS && SPLIT) {
        // lanes of a warp take consecutive tiles; warp = (plane, row half): 192 threads = 3 planes x 2 halves x 32 tiles
        const int tl = threadIdx.x & 31, plane = threadIdx.x >> 6, hf = (threadIdx.x >> 5) & 1;
        double S[4][8];
        if (tl < ntile) {
            const double2 *src = reinterpret_cast<const double2 *>(cpn + (plane * PQ_TILES + tl) * PQ_CPN_STRIDE + hf * 32);
#pragma unroll
            for (int i = 0; i < 16; i++) {
                const double2 v = src[i];
                S[i >> 2][2 * (i & 3)] = v.x;
                S[i >> 2][2 * (i & 3) + 1] = v.y;
            }
        } else {
#pragma unroll
            for (int y = 0; y < 4; y++)
#pragma unroll
                for (int x = 0; x < 8; x++) S[y][x] = 0.0;
        }
        __syncthreads(); // every component row is in registers: the buffer becomes the output staging + exchange area
        haar_half(S, hf, K.haar, xch + 2 * threadIdx.x, xch + 2 * (threadIdx.x ^ 32)); // (contains one __syncthreads)
        if (tl < ntile) {
            float *dst = sout + tl * PQ_OUT_STRIDE + plane * 64;
            auto st4 = [&](int row, int col, const double *v) {
                *reinterpret_cast<float4 *>(dst + row * 8 + col) = make_float4((float)v[0], (float)v[1], (float)v[2], (float)v[3]);
            };
            st4(hf, 0, &S[0][0]);          st4(2 * hf, 4, &S[0][4]);
            st4(2 + hf, 0, &S[1][0]);      st4(2 * hf + 1, 4, &S[1][4]);
            st4(4 + 2 * hf, 0, &S[2][0]);  st4(4 + 2 * hf, 4, &S[2][4]);
            st4(5 + 2 * hf, 0, &S[3][0]);  st4(5 + 2 * hf, 4, &S[3][4]);
        }
    } else {
        // lanes of a warp take consecutive tiles of ONE plane (96 threads = 3 planes x 32 tiles)
        const int plane = threadIdx.x / PQ_TILES, tl = threadIdx.x % PQ_TILES;
        double a[64];
        if (tl < ntile) {
            const double2 *src = reinterpret_cast<const double2 *>(cpn + (plane * PQ_TILES + tl) * PQ_CPN_STRIDE);
#pragma unroll
            for (int i = 0; i < 32; i++) {
                const double2 v = src[i];
                a[2 * i] = v.x;
                a[2 * i + 1] = v.y;
            }
        }
        __syncthreads(); // every component row is in registers: the buffer becomes the output staging
        if (tl < ntile) {
            float *dst = sout + tl * PQ_OUT_STRIDE + plane * 64;
            if (WAVELETS) {
                haar_level<8>(a, K.haar);
                haar_level<4>(a, K.haar);
                haar_level<2>(a, K.haar);
                store_row_smem(dst, a);
            } else {
                dct_plane(a, lut, qw, plane, dst);
            }
        }
    }
    __syncthreads();
    {   // coalesced store of the block's rows (48 float4 per row; rows are contiguous unless op.dst scatters them)
        for (int e = threadIdx.x; e < ntile * 48; e += NT) {
            const int t = e / 48, j = e - t * 48;
            reinterpret_cast<float4 *>(out + dst_row[t] * TM_DCT)[j] = *reinterpret_cast<const float4 *>(sout + t * PQ_OUT_STRIDE + j * 4);
        }
    }
    if (op.q16) { // 16 lanes per row, 3 float4 each; NT / 16 rows per pass (trip count uniform: shuffles need full warps)
        const int l16 = threadIdx.x & 15, grp = threadIdx.x >> 4;
        const float4 *mean4 = reinterpret_cast<const float4 *>(op.mean);
#pragma unroll 1
        for (int i = 0; i < (PQ_TILES + NT / 16 - 1) / (NT / 16); i++) {
            const int t = grp + (NT / 16) * i;
            float s16 = 0.f, sex = 0.f, sde = 0.f;
            if (t < ntile) {
                uint2 *drow = reinterpret_cast<uint2 *>(op.q16 + dst_row[t] * TM_DCT);
#pragma unroll
                for (int c = l16; c < 48; c += 16) {
                    const float4 x = *reinterpret_cast<const float4 *>(sout + t * PQ_OUT_STRIDE + c * 4);
                    const float4 m = __ldg(mean4 + c);
                    const float v0 = x.x - m.x, v1 = x.y - m.y, v2 = x.z - m.z, v3 = x.w - m.w;
                    const __half2 h01 = __floats2half2_rn(v0, v1), h23 = __floats2half2_rn(v2, v3);
                    const float2 f01 = __half22float2(h01), f23 = __half22float2(h23);
                    s16 += f01.x * f01.x; s16 += f01.y * f01.y; s16 += f23.x * f23.x; s16 += f23.y * f23.y;
                    sex += v0 * v0; sex += v1 * v1; sex += v2 * v2; sex += v3 * v3;
                    const float e0 = v0 - f01.x, e1 = v1 - f01.y, e2 = v2 - f23.x, e3 = v3 - f23.y; // exact
                    sde += e0 * e0; sde += e1 * e1; sde += e2 * e2; sde += e3 * e3;
                    uint2 pk;
                    pk.x = *reinterpret_cast<const unsigned int *>(&h01);
                    pk.y = *reinterpret_cast<const unsigned int *>(&h23);
                    drow[c] = pk;
                }
            }
#pragma unroll
            for (int o = 8; o > 0; o >>= 1) {
                s16 += __shfl_xor_sync(0xffffffffu, s16, o);
                sex += __shfl_xor_sync(0xffffffffu, sex, o);
                sde += __shfl_xor_sync(0xffffffffu, sde, o);
            }
            if (t < ntile && l16 == 0) {
                op.qn16[dst_row[t]] = s16;
                op.qne[dst_row[t]] = sqrtf(sex);
                op.qde[dst_row[t]] = sqrtf(sde);
            }
        }
    }
}

// Palette colours -> (y, u, v) once per call: a palette has pal_size colours, a dataset renders each of them
// thousands of times.  yuv is plane-major: [3][npal * pal_size].
__global__ void psyv_palette_yuv_kernel(const int32_t *__restrict__ pals, int ncol, const PsyvTables *__restrict__ tabs,
                                        PsyvConsts K, double *__restrict__ yuv) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= ncol) return;
    const int32_t col = pals[i];
    yuv[i] = plane_value(col, 0, tabs, K);
    yuv[ncol + i] = plane_value(col, 1, tabs, K);
    yuv[2 * ncol + i] = plane_value(col, 2, tabs, K);
}

// tileset rows: thread = (row, plane); row r = ((p*ntiles + t)*2 + vmir)*2 + hmir.  The colour transform is a
// gather from the (L1-resident) palette table; the float rows leave through padded shared memory so that the
// block's 32 x 768 B go out as coalesced 16-byte stores.
template <bool WAVELETS>
__global__ void __launch_bounds__(96) psyv_tileset_kernel(const uint8_t *__restrict__ palpix, int64_t ntiles,
                                                            const double *__restrict__ yuv, int npal,
                                                            int pal_size, const uint8_t *__restrict__ tflags,
                                                            PsyvConsts K,
                                                            const double *__restrict__ lut,
                                                            const double *__restrict__ qw,
                                                            const int64_t *__restrict__ row_map, int64_t nrows,
                                                            float *__restrict__ out) {
    __shared__ __align__(16) float sout[32 * PQ_OUT_STRIDE];
    const int plane = threadIdx.x / 32, rl = threadIdx.x & 31;
    const int64_t row0 = (int64_t)blockIdx.x * 32;
    const int64_t orow = row0 + rl; // output row
    const int nrow = (int)min((int64_t)32, nrows - row0);
    if (rl < nrow) {
        // source (pal, tile, vmir, hmir): every combination in order, or the keyframe's used subset
        const int64_t row = row_map ? row_map[orow] : orow;
        int hmir = (int)(row & 1), vmir = (int)((row >> 1) & 1);
        const int64_t t = (row >> 2) % ntiles;
        const int p = (int)((row >> 2) / ntiles);
        const int fl = tflags ? tflags[t] : 0;
        hmir ^= (fl & 1);          // hmir xor T^.HMirror main.pas:3912
        vmir ^= ((fl >> 1) & 1);
        const double *pal = yuv + (int64_t)plane * npal * pal_size + (int64_t)p * pal_size;
        // 64 palette indices as 4 x uint4
        const uint4 *src = reinterpret_cast<const uint4 *>(palpix + t * 64);
        uint32_t w[16];
#pragma unroll
        for (int i = 0; i < 4; i++) {
            uint4 v = __ldg(src + i);
            w[4 * i] = v.x; w[4 * i + 1] = v.y; w[4 * i + 2] = v.z; w[4 * i + 3] = v.w;
        }
        double a[64];
#pragma unroll
        for (int y = 0; y < 8; y++)
#pragma unroll
            for (int x = 0; x < 8; x++) {
                // output pixel (x,y) takes source pixel (xx,yy) main.pas:3041-3046
                int word_n = y * 2 + (x >> 2), word_v = (7 - y) * 2 + (x >> 2);
                int word_hn = y * 2 + ((7 - x) >> 2), word_hv = (7 - y) * 2 + ((7 - x) >> 2);
                uint32_t wn = vmir ? (hmir ? w[word_hv] : w[word_v]) : (hmir ? w[word_hn] : w[word_n]);
                int sh = hmir ? ((7 - x) & 3) * 8 : (x & 3) * 8;
                int pi = (wn >> sh) & 0xff;
                a[y * 8 + x] = __ldg(pal + pi);
            }
        float *dst = sout + rl * PQ_OUT_STRIDE + plane * 64;
        if (WAVELETS) {
            haar_level<8>(a, K.haar);
            haar_level<4>(a, K.haar);
            haar_level<2>(a, K.haar);
            store_row_smem(dst, a);
        } else {
            dct_plane(a, lut, qw, plane, dst);
        }
    }
    __syncthreads();
    {   // coalesced store of the block's rows (48 float4 per row)
        float4 *dst = reinterpret_cast<float4 *>(out + row0 * TM_DCT);
        for (int e = threadIdx.x; e < nrow * 48; e += 96) {
            const int t = e / 48, j = e - t * 48;
            dst[e] = *reinterpret_cast<const float4 *>(sout + t * PQ_OUT_STRIDE + j * 4);
        }
    }
}

// stage-1 rows (PrepareGlobalFT main.pas:3736-3780): order (none, H, H+V, V)
__global__ void global_rows_kernel(const uint8_t *__restrict__ palpix, int64_t ntiles,
                                   float *__restrict__ out) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; // one thread per output float4
    int64_t total = ntiles * 4 * 16;
    if (i >= total) return;
    int c4 = (int)(i & 15);
    int o = (int)((i >> 4) & 3);
    int64_t t = i >> 6;
    int hm = (o == 1 || o == 2), vm = (o >= 2);
    int y = c4 >> 1, x0 = (c4 & 1) * 4;
    int sy = vm ? 7 - y : y;
    const uint8_t *src = palpix + t * 64 + sy * 8;
    float v[4];
#pragma unroll
    for (int j = 0; j < 4; j++) {
        int x = x0 + j;
        v[j] = (float)src[hm ? 7 - x : x];
    }
    reinterpret_cast<float4 *>(out)[i] = make_float4(v[0], v[1], v[2], v[3]);
}

// ---- temporal smoothing (btnSmoothClick main.pas:1336-1370, DoTemporalSmoothing main.pas:4071-4119) -------------
// A tilemap position is a chain of items along the frame axis: step f compares the (possibly already replaced)
// item of frame f-1 with the original item of frame f, so frames are sequential and positions independent.  One
// warp owns one position and walks the frames; the descriptor of the carried item stays in shared memory, so
// each step renders at most one new Q-weighted DCT descriptor (lanes = 6 coefficients each, 64 sequential
// multiply-adds in the reference's order) and none when the two items are the same (tile, palette, mirrors):
// identical descriptors give cmp = 0 exactly.  The 192-term squared-difference sum is sequential like
// CompareEuclideanDCTPtr (main.pas:659-675).
constexpr int TS_WARPS = 8;
constexpr size_t TS_SMEM = sizeof(PsyvTables) + sizeof(double) * (4096 + 64 + 192 + TS_WARPS * 4 * 192);

struct TsItem {
    int tile, pal, attrs, src;
};

__device__ __forceinline__ void ts_descriptor(const TsItem &it, const uint8_t *__restrict__ palpix,
                                              const int32_t *__restrict__ kpal, int pal_size, const PsyvTables &T,
                                              const PsyvConsts &K, const double *lutT, const double *ratio,
                                              const double *qw, double *cpn, double *d, int lane, int *bad) {
    const int hm = it.attrs & 1, vm = (it.attrs >> 1) & 1;
    const int32_t *pal = kpal + (int64_t)it.pal * pal_size;
#pragma unroll
    for (int k = 0; k < 2; k++) {
        const int o = lane + 32 * k, x = o & 7, y = o >> 3;
        const int xx = hm ? 7 - x : x, yy = vm ? 7 - y : y; // main.pas:3041-3046
        int pi = palpix[(int64_t)it.tile * 64 + yy * 8 + xx];
        if (pi >= pal_size) { *bad = 1; pi = 0; }
        const int32_t col = __ldg(pal + pi);
        cpn[o] = plane_value(col, 0, &T, K);
        cpn[64 + o] = plane_value(col, 1, &T, K);
        cpn[128 + o] = plane_value(col, 2, &T, K);
    }
    __syncwarp();
    double z[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
    const int vu0 = lane, vu1 = lane + 32;
#pragma unroll 4
    for (int i = 0; i < 64; i++) { // main.pas:3083-3166, sequential in (y, x)
        const double l0 = lutT[i * 64 + vu0], l1 = lutT[i * 64 + vu1];
        const double c0 = cpn[i], c1 = cpn[64 + i], c2 = cpn[128 + i];
        z[0] += c0 * l0; z[1] += c0 * l1;
        z[2] += c1 * l0; z[3] += c1 * l1;
        z[4] += c2 * l0; z[5] += c2 * l1;
    }
#pragma unroll
    for (int j = 0; j < 6; j++) {
        const int c = lane + 32 * j;
        d[c] = (z[j] * qw[c]) * ratio[c & 63]; // z *= cDCTQuantization; pDCT^ := z * pRatio^ (main.pas:3168-3171)
    }
    __syncwarp();
}

__global__ void __launch_bounds__(TS_WARPS * 32) temporal_smoothing_kernel(
    const uint8_t *__restrict__ palpix, int64_t ntiles, const int32_t *__restrict__ pals, int npal, int pal_size,
    const int32_t *__restrict__ frame_kf, int64_t nframes, int64_t npos, const int32_t *__restrict__ tile,
    const int32_t *__restrict__ pal, const uint8_t *__restrict__ attrs, double strength,
    const PsyvTables *__restrict__ tabs, PsyvConsts K, const double *__restrict__ lut, const double *__restrict__ qwg,
    int32_t *__restrict__ src_frame, int32_t *__restrict__ out_tile, int32_t *__restrict__ out_pal,
    uint8_t *__restrict__ out_attrs, int *__restrict__ bad) {
    extern __shared__ __align__(16) uint8_t tsm[];
    PsyvTables &T = *reinterpret_cast<PsyvTables *>(tsm);
    double *lutT = reinterpret_cast<double *>(tsm + sizeof(PsyvTables)); // [i][vu]: lanes read consecutive words
    double *ratio = lutT + 4096;
    double *qw = ratio + 64;
    double *wbuf = qw + 192;
    for (int i = threadIdx.x; i < (int)(sizeof(PsyvTables) / 8); i += blockDim.x)
        reinterpret_cast<double *>(&T)[i] = reinterpret_cast<const double *>(tabs)[i];
    for (int i = threadIdx.x; i < 4096; i += blockDim.x) lutT[(i & 63) * 64 + (i >> 6)] = lut[i];
    for (int i = threadIdx.x; i < 64; i += blockDim.x) ratio[i] = lut[4096 + i];
    for (int i = threadIdx.x; i < 192; i += blockDim.x) qw[i] = qwg[i];
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    double *cpn = wbuf + warp * 4 * 192, *dP = cpn + 192, *dC = dP + 192, *sq = dC + 192;
    const double sqrt_factor = 1.0 / (8.0 * 8.0 * 3.0); // cSqrtFactor main.pas:4073
    int lbad = 0;
    for (int64_t p = (int64_t)blockIdx.x * TS_WARPS + warp; p < npos; p += (int64_t)gridDim.x * TS_WARPS) {
        TsItem prev{tile[p], pal[p], attrs[p], 0};
        if (prev.tile < 0 || prev.tile >= ntiles || prev.pal < 0 || prev.pal >= npal) { lbad = 1; prev.tile = 0; prev.pal = 0; }
        bool have_prev = false; // dP holds prev's descriptor
        int kf_prev = frame_kf[0];
        for (int64_t f = 1; f < nframes; f++) {
            const int64_t c = f * npos + p;
            TsItem cur{tile[c], pal[c], attrs[c], (int)f};
            if (cur.tile < 0 || cur.tile >= ntiles || cur.pal < 0 || cur.pal >= npal) { lbad = 1; cur.tile = 0; cur.pal = 0; }
            const int kf = frame_kf[f];
            if (kf == kf_prev) { // main.pas:4082-4083: only inside one keyframe
                double cmp = 0.0;
                const bool same = cur.tile == prev.tile && cur.pal == prev.pal && ((cur.attrs ^ prev.attrs) & 3) == 0;
                if (!same) {
                    const int32_t *kpal = pals + (int64_t)kf * npal * pal_size;
                    if (!have_prev) ts_descriptor(prev, palpix, kpal, pal_size, T, K, lutT, ratio, qw, cpn, dP, lane, &lbad);
                    ts_descriptor(cur, palpix, kpal, pal_size, T, K, lutT, ratio, qw, cpn, dC, lane, &lbad);
                    have_prev = true;
#pragma unroll
                    for (int j = 0; j < 6; j++) {
                        const double e = dC[lane + 32 * j] - dP[lane + 32 * j];
                        sq[lane + 32 * j] = e * e;
                    }
                    __syncwarp();
                    for (int i = 0; i < 192; i++) cmp += sq[i]; // every lane: same sequential sum, no broadcast needed
                    cmp = sqrt(cmp * sqrt_factor);
                    __syncwarp();
                }
                bool cur_desc_is_dC = !same;
                if (fabs(cmp) <= strength) {
                    if (cur.tile >= prev.tile) { cur = prev; cur_desc_is_dC = false; } // TMI^ := PrevTMI^
                    else prev = cur;                                                   // PrevTMI^ := TMI^
                    cur.attrs |= 4;
                } else {
                    cur.attrs &= ~4;
                }
                if (cur_desc_is_dC) { double *t = dP; dP = dC; dC = t; }
            } else {
                have_prev = false;
            }
            if (lane == 0) {
                const int64_t q = c - npos;
                src_frame[q] = prev.src; out_tile[q] = prev.tile; out_pal[q] = prev.pal; out_attrs[q] = (uint8_t)prev.attrs;
            }
            prev = cur;
            kf_prev = kf;
        }
        if (lane == 0 && nframes > 0) {
            const int64_t q = (nframes - 1) * npos + p;
            src_frame[q] = prev.src; out_tile[q] = prev.tile; out_pal[q] = prev.pal; out_attrs[q] = (uint8_t)prev.attrs;
        }
    }
    if (lbad) atomicOr(bad, 1);
}

struct PsyvDeviceState {
    PsyvTables *tabs[3] = {nullptr, nullptr, nullptr}; // gamma -1, 0, 1
    double *lut = nullptr;                            // 4096 DCT lut + 64 cUVRatio
    double *qw = nullptr;                             // 192 Q weights
    int device = -1;
    bool linear_ok = false; // the arithmetic c / 255.0 matched the table for all 256 values
} g_ps_dev[TM_MAX_DEV]; // one table set per device; callers hold that device's lock
#define g_ps (g_ps_dev[tm_dev()->device])

const int q_luma[64] = {16, 11, 10, 16, 24, 40, 51, 61, 12, 12, 14, 19, 26, 58, 60, 55,
                        14, 13, 16, 24, 40, 57, 69, 56, 14, 17, 22, 29, 51, 87, 80, 62,
                        18, 22, 37, 56, 68, 109, 103, 77, 24, 35, 55, 64, 81, 104, 113, 92,
                        49, 64, 78, 87, 103, 121, 120, 101, 72, 92, 95, 98, 112, 100, 103, 99};
const int q_chroma[64] = {17, 18, 24, 47, 99, 99, 99, 99, 18, 21, 26, 66, 99, 99, 99, 112,
                          24, 26, 56, 99, 99, 99, 112, 128, 47, 66, 99, 99, 99, 112, 128, 144,
                          99, 99, 99, 99, 112, 128, 144, 160, 99, 99, 99, 112, 128, 144, 160, 176,
                          99, 99, 112, 128, 144, 160, 176, 192, 99, 112, 128, 144, 160, 176, 192, 208};

int ensure_tables() {
    const int dev = tm_dev()->device;
    if (g_ps.device == dev && g_ps.lut) return 0;
    static const double gam[2] = {2.0, 0.6}; // main.pas:586
    {   // the kernels' table-free evaluation of c / 255.0 (channel_value), replayed on the host
        const double rcp = 1.0 / 255.0;
        bool ok = true;
        for (int c = 0; c < 256; c++) {
            const double cd = (double)c, q0 = cd * rcp;
            ok = ok && std::fma(std::fma(-q0, 255.0, cd), rcp, q0) == c / 255.0;
        }
        g_ps.linear_ok = ok;
    }
    for (int g = -1; g < 2; g++) {
        PsyvTables *h = new PsyvTables;
        for (int c = 0; c < 256; c++) {
            double f = g >= 0 ? pow(c / 255.0, gam[g]) : c / 255.0; // main.pas:596-601
            h->f[c] = f;
            h->fr[c] = 2126.0 * f;
            h->fg[c] = 7152.0 * f;
            h->fb[c] = 722.0 * f;
        }
        TM_CUDA(cudaMalloc(&g_ps.tabs[g + 1], sizeof(PsyvTables)));
        TM_CUDA(cudaMemcpy(g_ps.tabs[g + 1], h, sizeof(PsyvTables), cudaMemcpyHostToDevice));
        delete h;
    }
    double *lut = new double[4096 + 64];
    int i = 0;
    for (int v = 0; v < 8; v++)
        for (int u = 0; u < 8; u++)
            for (int y = 0; y < 8; y++)
                for (int x = 0; x < 8; x++)
                    lut[i++] = cos((x + 0.5) * u * M_PI / 16.0) * cos((y + 0.5) * v * M_PI / 16.0);
    for (int v = 0; v < 8; v++)
        for (int u = 0; u < 8; u++) {
            double r = (u == 0 ? sqrt(0.5) : 1.0) * (v == 0 ? sqrt(0.5) : 1.0);
            if (u == 0 && v == 0) r = 0.5;
            lut[4096 + v * 8 + u] = r;
        }
    TM_CUDA(cudaMalloc(&g_ps.lut, sizeof(double) * (4096 + 64)));
    TM_CUDA(cudaMemcpy(g_ps.lut, lut, sizeof(double) * (4096 + 64), cudaMemcpyHostToDevice));
    delete[] lut;
    double qw[192];
    for (int c = 0; c < 3; c++)
        for (int j = 0; j < 64; j++) qw[c * 64 + j] = sqrt(16.0) / sqrt((double)(c == 0 ? q_luma : q_chroma)[j]);
    TM_CUDA(cudaMalloc(&g_ps.qw, sizeof(qw)));
    TM_CUDA(cudaMemcpy(g_ps.qw, qw, sizeof(qw), cudaMemcpyHostToDevice));
    g_ps.device = dev;
    return 0;
}

PsyvConsts make_consts() {
    PsyvConsts K;
    const double luma = 2126.0 + 7152.0 + 722.0;
    K.cu = 0.5 / (1.0 - 722.0 / luma);
    K.cv = 0.5 / (1.0 - 2126.0 / luma);
    K.luma_div = luma;
    K.luma_rcp = 1.0 / luma;
    K.haar = 1.0 / sqrt(2.0);
    K.rcp255 = 1.0 / 255.0;
    return K;
}

} // namespace

int tm_launch_psyv_queries(const int32_t *d_rgb, int64_t nq, int flags, int gamma, float *d_out,
                           cudaStream_t s) {
    return tm_launch_psyv_queries_operand(d_rgb, nq, flags, gamma, d_out, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, s);
}

int tm_launch_psyv_queries_operand(const int32_t *d_rgb, int64_t nq, int flags, int gamma, float *d_out,
                                   const float *d_mean, void *d_q16, float *d_qn16, float *d_qne, float *d_qde,
                                   const int *d_dst, cudaStream_t s) {
    if (gamma < -1 || gamma > 1) { tm_set_error("gamma must be -1, 0 or 1"); return TM_EINVAL; }
    if (nq == 0) return 0;
    int r = ensure_tables();
    if (r) return r;
    PsyvConsts K = make_consts();
    unsigned grid = (unsigned)((nq + PQ_TILES - 1) / PQ_TILES);
    const double *qw = (flags & TM_PSYV_QWEIGHT) ? g_ps.qw : nullptr;
    static std::atomic<uint32_t> attr{0}; // per device (tm_first_on_device)
    if (tm_first_on_device(attr)) {
        TM_CUDA(cudaFuncSetAttribute(psyv_queries_kernel<true, true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)PQ_SMEM_LINEAR));
        TM_CUDA(cudaFuncSetAttribute(psyv_queries_kernel<true, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)PQ_SMEM_LINEAR));
        TM_CUDA(cudaFuncSetAttribute(psyv_queries_kernel<false, true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)PQ_SMEM_LINEAR));
        TM_CUDA(cudaFuncSetAttribute(psyv_queries_kernel<true, false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)PQ_SMEM_LUT));
        TM_CUDA(cudaFuncSetAttribute(psyv_queries_kernel<true, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)PQ_SMEM_LUT));
        TM_CUDA(cudaFuncSetAttribute(psyv_queries_kernel<false, false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)PQ_SMEM_LUT));
    }
    const bool linear = gamma < 0 && g_ps.linear_ok; // GammaCor = -1: c / 255.0 without the table
    const bool wav = (flags & TM_PSYV_WAVELETS) != 0;
    const bool split = wav && d_q16 != nullptr; // fused operand: two threads per plane (see the kernel)
    const PsyvTables *tabs = g_ps.tabs[gamma + 1];
    const PsyvOperandOut op{d_mean, static_cast<__half *>(d_q16), d_qn16, d_qne, d_qde, d_dst};
    if (wav && linear && split)
        psyv_queries_kernel<true, true, true><<<grid, PQ_THREADS_W, PQ_SMEM_LINEAR, s>>>(d_rgb, nq, tabs, K, g_ps.lut, qw, d_out, op);
    else if (wav && linear)
        psyv_queries_kernel<true, true, false><<<grid, PQ_THREADS, PQ_SMEM_LINEAR, s>>>(d_rgb, nq, tabs, K, g_ps.lut, qw, d_out, op);
    else if (wav && split)
        psyv_queries_kernel<true, false, true><<<grid, PQ_THREADS_W, PQ_SMEM_LUT, s>>>(d_rgb, nq, tabs, K, g_ps.lut, qw, d_out, op);
    else if (wav)
        psyv_queries_kernel<true, false, false><<<grid, PQ_THREADS, PQ_SMEM_LUT, s>>>(d_rgb, nq, tabs, K, g_ps.lut, qw, d_out, op);
    else if (linear)
        psyv_queries_kernel<false, true, false><<<grid, PQ_THREADS, PQ_SMEM_LINEAR, s>>>(d_rgb, nq, tabs, K, g_ps.lut, qw, d_out, op);
    else
        psyv_queries_kernel<false, false, false><<<grid, PQ_THREADS, PQ_SMEM_LUT, s>>>(d_rgb, nq, tabs, K, g_ps.lut, qw, d_out, op);
    tm_count_launch();
    return tm_check(cudaGetLastError(), "psyv_queries_kernel");
}

int tm_launch_psyv_tileset_rows(const uint8_t *d_palpix, int64_t ntiles, const int32_t *d_pals, int npal,
                                int pal_size, const uint8_t *d_tile_flags, int flags, int gamma,
                                const int64_t *d_row_map, int64_t nrows, float *d_out, cudaStream_t s) {
    if (gamma < -1 || gamma > 1) { tm_set_error("gamma must be -1, 0 or 1"); return TM_EINVAL; }
    if (nrows == 0) return 0;
    int r = ensure_tables();
    if (r) return r;
    PsyvConsts K = make_consts();
    unsigned grid = (unsigned)((nrows + 31) / 32);
    const double *qw = (flags & TM_PSYV_QWEIGHT) ? g_ps.qw : nullptr;
    const int ncol = npal * pal_size;
    double *yuv = nullptr;
    TM_CUDA(tm_dev_malloc(&yuv, sizeof(double) * 3 * (size_t)ncol));
    psyv_palette_yuv_kernel<<<(unsigned)((ncol + 127) / 128), 128, 0, s>>>(d_pals, ncol, g_ps.tabs[gamma + 1], K, yuv);
    if (flags & TM_PSYV_WAVELETS)
        psyv_tileset_kernel<true><<<grid, 96, 0, s>>>(d_palpix, ntiles, yuv, npal, pal_size, d_tile_flags, K, g_ps.lut, qw,
                                                       d_row_map, nrows, d_out);
    else
        psyv_tileset_kernel<false><<<grid, 96, 0, s>>>(d_palpix, ntiles, yuv, npal, pal_size, d_tile_flags, K, g_ps.lut, qw,
                                                        d_row_map, nrows, d_out);
    tm_dev_free(yuv); // stream-ordered: released after the kernel above
    tm_count_launch();
    tm_count_launch();
    return tm_check(cudaGetLastError(), "psyv_tileset_kernel");
}

int tm_launch_psyv_tileset(const uint8_t *d_palpix, int64_t ntiles, const int32_t *d_pals, int npal,
                           int pal_size, const uint8_t *d_tile_flags, int flags, int gamma,
                           float *d_out, cudaStream_t s) {
    return tm_launch_psyv_tileset_rows(d_palpix, ntiles, d_pals, npal, pal_size, d_tile_flags, flags, gamma, nullptr,
                                       ntiles * 4 * npal, d_out, s);
}

int tm_launch_global_rows(const uint8_t *d_palpix, int64_t ntiles, float *d_out, cudaStream_t s) {
    int64_t total = ntiles * 64;
    if (total == 0) return 0;
    global_rows_kernel<<<(unsigned)((total + 255) / 256), 256, 0, s>>>(d_palpix, ntiles, d_out);
    tm_count_launch();
    return tm_check(cudaGetLastError(), "global_rows_kernel");
}

int tm_launch_temporal_smoothing(const uint8_t *d_palpix, int64_t ntiles, const int32_t *d_pals, int npal, int pal_size,
                                 const int32_t *d_frame_kf, int64_t nframes, int64_t npos, const int32_t *d_tile,
                                 const int32_t *d_pal, const uint8_t *d_attrs, double strength, int32_t *d_src_frame,
                                 int32_t *d_out_tile, int32_t *d_out_pal, uint8_t *d_out_attrs, int *d_bad,
                                 cudaStream_t s) {
    if (nframes == 0 || npos == 0) return 0;
    int r = ensure_tables();
    if (r) return r;
    static std::atomic<uint32_t> attr{0}; // per device (tm_first_on_device)
    if (tm_first_on_device(attr)) {
        TM_CUDA(cudaFuncSetAttribute(temporal_smoothing_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TS_SMEM));
    }
    const int64_t blocks = (npos + TS_WARPS - 1) / TS_WARPS;
    const unsigned grid = (unsigned)(blocks < 148 * 8 ? blocks : 148 * 8);
    temporal_smoothing_kernel<<<grid, TS_WARPS * 32, TS_SMEM, s>>>(d_palpix, ntiles, d_pals, npal, pal_size, d_frame_kf,
                                                                  nframes, npos, d_tile, d_pal, d_attrs, strength,
                                                                  g_ps.tabs[0], make_consts(), g_ps.lut, g_ps.qw,
                                                                  d_src_frame, d_out_tile, d_out_pal, d_out_attrs, d_bad);
    tm_count_launch();
    return tm_check(cudaGetLastError(), "temporal_smoothing_kernel");
}
