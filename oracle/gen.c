/* gen.c — deterministic synthetic inputs, built on the
 * reference's own LCG RandInt (kmodes.pas:82-86) so an FPC build could
 * regenerate identical data.  TEST INFRASTRUCTURE ONLY, PARITY UNPINNED (see tmoracle.h).
 */
#include "tmoracle.h"
#include <string.h>

uint32_t tmo_randint(uint32_t range, uint32_t *seed) {
    *seed = (uint32_t)(*seed * 0x08088405u) + 1u;
    return (uint32_t)(((uint64_t)*seed * (uint64_t)range) >> 32);
}

static int luma_i(int32_t c) {
    return 2126 * (c & 0xff) + 7152 * ((c >> 8) & 0xff) + 722 * ((c >> 16) & 0xff);
}

void tmo_gen_palette(uint32_t *seed, int pal_size, int32_t *pal) {
    for (int i = 0; i < pal_size; i++) {
        int r = (int)tmo_randint(256, seed), g = (int)tmo_randint(256, seed),
            b = (int)tmo_randint(256, seed);
        pal[i] = (b << 16) | (g << 8) | r; /* ToRGB main.pas:566-569 */
    }
    for (int i = 1; i < pal_size; i++) { /* insertion sort by luma */
        int32_t c = pal[i];
        int j = i;
        while (j > 0 && luma_i(pal[j - 1]) > luma_i(c)) { pal[j] = pal[j - 1]; j--; }
        pal[j] = c;
    }
}

/* per-tile: random base + linear gradient in a random direction + +-1 noise,
 * then canonicalised like PrepareTileMirrors (main.pas:4049-4069). */
void tmo_gen_tileset(uint32_t *seed, int64_t ntiles, int pal_size, uint8_t *palpix) {
    uint32_t s0 = *seed;
    for (int64_t t = 0; t < ntiles; t++) {
        uint32_t s = s0 + (uint32_t)t * 0x9E3779B9u;
        tmo_randint(2, &s);
        int base = (int)tmo_randint((uint32_t)pal_size, &s);
        int gx = (int)tmo_randint(7, &s) - 3, gy = (int)tmo_randint(7, &s) - 3;
        int div = 2 + (int)tmo_randint(6, &s);
        uint8_t *p = palpix + t * 64;
        for (int y = 0; y < 8; y++)
            for (int x = 0; x < 8; x++) {
                int v = base + (gx * (x - 4) + gy * (y - 4)) / div + (int)tmo_randint(3, &s) - 1;
                if (v < 0) v = 0;
                if (v > pal_size - 1) v = pal_size - 1;
                p[y * 8 + x] = (uint8_t)v;
            }
        tmo_prepare_tile_mirrors(p);
    }
    *seed = s0 + (uint32_t)ntiles * 0x9E3779B9u + 1u;
}

void tmo_gen_queries(uint32_t *seed, int64_t nq, const uint8_t *palpix, int64_t ntiles,
                     const int32_t *pal, int noise, int32_t *rgb) {
    uint32_t s0 = *seed;
#pragma omp parallel for schedule(static)
    for (int64_t qi = 0; qi < nq; qi++) {
        uint32_t s = s0 + (uint32_t)qi * 0x9E3779B9u;
        tmo_randint(2, &s);
        int32_t *o = rgb + qi * 64;
        if (tmo_randint(16, &s) == 0) { /* pure random RGB: no planted answer */
            for (int i = 0; i < 64; i++) o[i] = (int32_t)tmo_randint(1u << 24, &s);
            continue;
        }
        int64_t t = (int64_t)tmo_randint((uint32_t)ntiles, &s);
        int hm = (int)tmo_randint(2, &s), vm = (int)tmo_randint(2, &s);
        for (int y = 0; y < 8; y++)
            for (int x = 0; x < 8; x++) {
                int xx = hm ? 7 - x : x, yy = vm ? 7 - y : y;
                int32_t c = pal[palpix[t * 64 + yy * 8 + xx]];
                int ch[3];
                for (int k = 0; k < 3; k++) {
                    int v = ((c >> (8 * k)) & 0xff) + (int)tmo_randint((uint32_t)(2 * noise + 1), &s) - noise;
                    ch[k] = v < 0 ? 0 : (v > 255 ? 255 : v);
                }
                o[y * 8 + x] = (ch[2] << 16) | (ch[1] << 8) | ch[0];
            }
    }
    *seed = s0 + (uint32_t)nq * 0x9E3779B9u + 1u;
}
