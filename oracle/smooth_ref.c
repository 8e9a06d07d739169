/* smooth_ref.c — CPU restatement of the temporal smoothing pass (TEST INFRASTRUCTURE ONLY, PARITY
 * UNPINNED: the reference is FreePascal + a Win32 DLL and cannot be built here; see tmoracle.h).
 *
 * btnSmoothClick (reference main.pas:1336-1370): SmoothedTileMap := TileMap, then for every tilemap row Y
 * (parallel) and every frame i >= 1 (sequential) DoTemporalSmoothing(Frames[i], Frames[i-1], Y, strength).
 * DoTemporalSmoothing (main.pas:4071-4119): skip when the two frames belong to different keyframes;
 * otherwise, position by position, the Q-weighted DCT descriptors (FromPal, no wavelets, no LAB,
 * QWeighting, GammaCor = -1) of the two tilemap items are compared (CompareEuclideanDCT main.pas:659-680,
 * sequential sum of squared differences), cmp = sqrt(sum / 192); if cmp <= strength the item with the
 * higher tile index is replaced by the other one (whole record) and the current item is flagged Smoothed,
 * else the current item's flag is cleared.
 *
 * A position is a column of items along the frame axis; rows Y only batch positions, so the restatement
 * loops positions x frames.  attrs: bit0 HMirror, bit1 VMirror, bit2 Smoothed. */
#include "tmoracle.h"
#include <math.h>
#include <string.h>

int tmo_temporal_smoothing(const uint8_t *palpix, int64_t ntiles, const int32_t *pals, int nkf, int npal,
                           int pal_size, const int32_t *frame_kf, int64_t nframes, int64_t npos,
                           const int32_t *tile, const int32_t *pal, const uint8_t *attrs, double strength,
                           int32_t *src_frame, int32_t *out_tile, int32_t *out_pal, uint8_t *out_attrs) {
    const double sqrt_factor = 1.0 / (8.0 * 8.0 * 3.0); /* cSqrtFactor main.pas:4073 */
    for (int64_t f = 0; f < nframes; f++)
        if (frame_kf[f] < 0 || frame_kf[f] >= nkf) return -1;
    for (int64_t i = 0; i < nframes * npos; i++) { /* main.pas:1360-1363 */
        if (tile[i] < 0 || tile[i] >= ntiles || pal[i] < 0 || pal[i] >= npal) return -1;
        src_frame[i] = (int32_t)(i / npos);
        out_tile[i] = tile[i];
        out_pal[i] = pal[i];
        out_attrs[i] = attrs[i];
    }
    int bad = 0;
#pragma omp parallel for schedule(dynamic, 16)
    for (int64_t p = 0; p < npos; p++) {
        double dprev[192], dcur[192];
        for (int64_t f = 1; f < nframes; f++) {
            if (frame_kf[f] != frame_kf[f - 1]) continue; /* main.pas:4082-4083 */
            const int64_t c = f * npos + p, q = (f - 1) * npos + p;
            const int32_t *kpal = pals + (int64_t)frame_kf[f] * npal * pal_size;
            for (int i = 0; i < 64; i++)
                if (palpix[(int64_t)out_tile[q] * 64 + i] >= pal_size || palpix[(int64_t)out_tile[c] * 64 + i] >= pal_size) bad = 1;
            if (bad) break;
            tmo_psyv(0, palpix + (int64_t)out_tile[q] * 64, kpal + (int64_t)out_pal[q] * pal_size, 1, 0, 1,
                     out_attrs[q] & 1, (out_attrs[q] >> 1) & 1, -1, dprev);
            tmo_psyv(0, palpix + (int64_t)out_tile[c] * 64, kpal + (int64_t)out_pal[c] * pal_size, 1, 0, 1,
                     out_attrs[c] & 1, (out_attrs[c] >> 1) & 1, -1, dcur);
            double cmp = 0.0;
            for (int i = 0; i < 192; i++) cmp += (dcur[i] - dprev[i]) * (dcur[i] - dprev[i]);
            cmp = sqrt(cmp * sqrt_factor);
            if (fabs(cmp) <= strength) {
                if (out_tile[c] >= out_tile[q]) { /* TMI^ := PrevTMI^ */
                    out_tile[c] = out_tile[q]; out_pal[c] = out_pal[q]; out_attrs[c] = out_attrs[q]; src_frame[c] = src_frame[q];
                } else {                          /* PrevTMI^ := TMI^ */
                    out_tile[q] = out_tile[c]; out_pal[q] = out_pal[c]; out_attrs[q] = out_attrs[c]; src_frame[q] = src_frame[c];
                }
                out_attrs[c] |= 4;
            } else {
                out_attrs[c] &= (uint8_t)~4;
            }
        }
    }
    return bad ? -1 : 0;
}
