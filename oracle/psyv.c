/* psyv.c — oracle restatement of the PsyV descriptor and tile mirror helpers.
 * TEST INFRASTRUCTURE ONLY (see tmoracle.h).  PARITY UNPINNED.
 * Follows /root/reference main.pas: RGBToYUV 2656-2679, WaveletGS 2805-2840,
 * ComputeTilePsyVisFeatures 2997-3177, InitLuts 592-623, mirrors 3179-3209,
 * PrepareTileMirrors 4049-4069, PrepareGlobalFT 3736-3780.
 */
#include "tmoracle.h"
#include <math.h>
#include <string.h>

/* main.pas:25-27,37 */
#define RED_MUL 2126.0
#define GREEN_MUL 7152.0
#define BLUE_MUL 722.0
#define LUMA_DIV (2126.0 + 7152.0 + 722.0)

static const double g_gamma[2] = {2.0, 0.6}; /* main.pas:586 */

/* JPEG-derived divisors behind cDCTQuantization (main.pas:62-98): weight = sqrt(16)/sqrt(div) */
static const int q_luma[64] = {
    16, 11, 10, 16, 24, 40, 51, 61,   12, 12, 14, 19, 26, 58, 60, 55,
    14, 13, 16, 24, 40, 57, 69, 56,   14, 17, 22, 29, 51, 87, 80, 62,
    18, 22, 37, 56, 68, 109, 103, 77, 24, 35, 55, 64, 81, 104, 113, 92,
    49, 64, 78, 87, 103, 121, 120, 101, 72, 92, 95, 98, 112, 100, 103, 99};
static const int q_chroma[64] = {
    17, 18, 24, 47, 99, 99, 99, 99,      18, 21, 26, 66, 99, 99, 99, 112,
    24, 26, 56, 99, 99, 99, 112, 128,    47, 66, 99, 99, 99, 112, 128, 144,
    99, 99, 99, 99, 112, 128, 144, 160,  99, 99, 99, 112, 128, 144, 160, 176,
    99, 99, 112, 128, 144, 160, 176, 192, 99, 112, 128, 144, 160, 176, 192, 208};

static double g_dct_lut[64 * 64];
static int g_lut_ready = 0;

static void init_luts(void) {
    /* main.pas:613-623: index order v,u,y,x */
    int i = 0;
    for (int v = 0; v < 8; v++)
        for (int u = 0; u < 8; u++)
            for (int y = 0; y < 8; y++)
                for (int x = 0; x < 8; x++)
                    g_dct_lut[i++] = cos((x + 0.5) * u * M_PI / 16.0) *
                                     cos((y + 0.5) * v * M_PI / 16.0);
    g_lut_ready = 1;
}

void tmo_rgb_to_yuv(int r, int g, int b, int gamma, double *y, double *u, double *v) {
    double fr, fg, fb;
    if (gamma >= 0) { /* main.pas:596-601 */
        fr = pow(r / 255.0, g_gamma[gamma]);
        fg = pow(g / 255.0, g_gamma[gamma]);
        fb = pow(b / 255.0, g_gamma[gamma]);
    } else {
        fr = r / 255.0; fg = g / 255.0; fb = b / 255.0;
    }
    double yy = (RED_MUL * fr + GREEN_MUL * fg + BLUE_MUL * fb) / LUMA_DIV;
    *u = (fb - yy) * (0.5 / (1.0 - BLUE_MUL / LUMA_DIV));
    *v = (fr - yy) * (0.5 / (1.0 - RED_MUL / LUMA_DIV));
    *y = yy;
}

void tmo_wavelet_gs(const double *in, double *out, int dx, int dy, int depth) {
    double tx[64], ty[64];
    memset(tx, 0, sizeof tx);
    memset(ty, 0, sizeof ty);
    const double f = 1.0 / sqrt(2.0);
    for (int y = 0; y < dy; y++) {
        int o = y * 8;
        for (int x = 0; x < dx / 2; x++) {
            tx[x + o] = (in[x * 2 + o] + in[x * 2 + 1 + o]) * f;
            tx[x + dx / 2 + o] = (in[x * 2 + o] - in[x * 2 + 1 + o]) * f;
        }
    }
    for (int x = 0; x < dx; x++)
        for (int y = 0; y < dy / 2; y++) {
            ty[x + y * 8] = (tx[x + y * 2 * 8] + tx[x + (y * 2 + 1) * 8]) * f;
            ty[x + (y + dy / 2) * 8] = (tx[x + y * 2 * 8] - tx[x + (y * 2 + 1) * 8]) * f;
        }
    for (int y = 0; y < dy; y++) memcpy(out + y * 8, ty + y * 8, dx * sizeof(double));
    if (depth > 0) tmo_wavelet_gs(out, out, dx / 2, dy / 2, depth - 1);
}

void tmo_psyv(const int32_t *rgb, const uint8_t *palpix, const int32_t *pal,
              int from_pal, int use_wavelets, int qweighting, int hmirror,
              int vmirror, int gamma, double *out) {
    double cpn[3][64];
    if (!g_lut_ready) init_luts();
    for (int y = 0; y < 8; y++)
        for (int x = 0; x < 8; x++) {
            int xx = hmirror ? 7 - x : x, yy = vmirror ? 7 - y : y;
            int32_t col = from_pal ? pal[palpix[yy * 8 + xx]] : rgb[yy * 8 + xx];
            /* FromRGB main.pas:571-576: r = low byte */
            tmo_rgb_to_yuv(col & 0xff, (col >> 8) & 0xff, (col >> 16) & 0xff, gamma,
                           &cpn[0][y * 8 + x], &cpn[1][y * 8 + x], &cpn[2][y * 8 + x]);
        }
    if (use_wavelets) {
        for (int c = 0; c < 3; c++) tmo_wavelet_gs(cpn[c], out + c * 64, 8, 8, 2);
        return;
    }
    for (int c = 0; c < 3; c++) {
        const double *lut = g_dct_lut;
        for (int v = 0; v < 8; v++)
            for (int u = 0; u < 8; u++) {
                double z = 0.0;
                for (int i = 0; i < 64; i++) z += cpn[c][i] * lut[i];
                lut += 64;
                if (qweighting)
                    z *= sqrt(16.0) / sqrt((double)(c == 0 ? q_luma : q_chroma)[v * 8 + u]);
                /* cUVRatio main.pas:3000-3009 */
                double ratio = (u == 0 ? sqrt(0.5) : 1.0) * (v == 0 ? sqrt(0.5) : 1.0);
                if (u == 0 && v == 0) ratio = 0.5;
                out[c * 64 + v * 8 + u] = z * ratio;
            }
    }
}

void tmo_psyv_queries_f32(const int32_t *rgb, int64_t nq, int use_wavelets,
                          int gamma, float *out) {
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < nq; i++) {
        double d[192];
        tmo_psyv(rgb + i * 64, 0, 0, 0, use_wavelets, 0, 0, 0, gamma, d);
        for (int j = 0; j < 192; j++) out[i * 192 + j] = (float)d[j];
    }
}

void tmo_psyv_tileset_f32(const uint8_t *palpix, int64_t ntiles, const int32_t *pal,
                          int use_wavelets, int gamma, float *out) {
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < ntiles; i++)
        for (int vm = 0; vm < 2; vm++)
            for (int hm = 0; hm < 2; hm++) {
                double d[192];
                tmo_psyv(0, palpix + i * 64, pal, 1, use_wavelets, 0, hm, vm, gamma, d);
                float *o = out + ((i * 4) + vm * 2 + hm) * 192;
                for (int j = 0; j < 192; j++) o[j] = (float)d[j];
            }
}

void tmo_hmirror(uint8_t *p) {
    for (int j = 0; j < 8; j++)
        for (int i = 0; i < 4; i++) {
            uint8_t v = p[j * 8 + i];
            p[j * 8 + i] = p[j * 8 + 7 - i];
            p[j * 8 + 7 - i] = v;
        }
}

void tmo_vmirror(uint8_t *p) {
    for (int j = 0; j < 4; j++)
        for (int i = 0; i < 8; i++) {
            uint8_t v = p[j * 8 + i];
            p[j * 8 + i] = p[(7 - j) * 8 + i];
            p[(7 - j) * 8 + i] = v;
        }
}

int tmo_prepare_tile_mirrors(uint8_t *p) {
    int best = -1, hf_best = 0, vf_best = 0;
    for (int vf = 0; vf < 2; vf++)
        for (int hf = 0; hf < 2; hf++) {
            int s = 0; /* GetTileZoneSum main.pas:4132-4140 */
            for (int j = vf * 4; j < vf * 4 + 4; j++)
                for (int i = hf * 4; i < hf * 4 + 4; i++) s += p[j * 8 + i];
            if (s > best) { best = s; hf_best = hf; vf_best = vf; }
        }
    if (hf_best) tmo_hmirror(p);
    if (vf_best) tmo_vmirror(p);
    return hf_best | (vf_best << 1);
}

void tmo_global_rows_f32(const uint8_t *palpix, int64_t ntiles, float *out,
                         int32_t *tr_tile, uint8_t *tr_attrs) {
    static const int hm[4] = {0, 1, 1, 0}, vm[4] = {0, 0, 1, 1};
    for (int64_t t = 0; t < ntiles; t++) {
        uint8_t tmp[64];
        memcpy(tmp, palpix + t * 64, 64);
        for (int o = 0; o < 4; o++) {
            /* sequence main.pas:3769-3776: none, H, then V (=HV), then H (=V) */
            if (o == 1) tmo_hmirror(tmp);
            if (o == 2) tmo_vmirror(tmp);
            if (o == 3) tmo_hmirror(tmp);
            for (int j = 0; j < 64; j++) out[(t * 4 + o) * 64 + j] = tmp[j];
            if (tr_tile) tr_tile[t * 4 + o] = (int32_t)t;
            if (tr_attrs) tr_attrs[t * 4 + o] = (uint8_t)(hm[o] | (vm[o] << 1));
        }
    }
}
