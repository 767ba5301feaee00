/* oracle/ois_oracle.c -- TEST INFRASTRUCTURE ONLY (never linked into the product).
 *
 * Plain-C restatement of the open-loop intra search of the ME process, row (k) of the hot-path scope table
 * (DESIGN.md section 0, section 3.8): OpenLoopIntraSearchLcu and what it calls
 * (reference Source/Lib/Codec/EbMotionEstimation.c:7419-8365, EbIntraPrediction.c:3718-3800, 5230-5430; leaves in
 * ASM_AVX2/EbIntraPrediction_Intrinsic_AVX2.c, ASM_SSE2/EbIntraPrediction_Intrinsic_SSE2.c). For every 32x32 / 16x16 /
 * 8x8 CU of an SB (raster-scan CU index 1..84; index == the ME pu_index of the same square) the reference predicts the CU
 * from neighbours taken out of the SOURCE picture (missing neighbours = 128, no substitution), measures the SAD of a few
 * or all of the 35 HEVC-style intra modes, and leaves a candidate list per CU that mode decision reads.
 *
 * Pinned by tests/test_ois_oracle_vs_ref.py against the compiled reference function (oracle/_ref/libsvtref.so,
 * ref_ois_picture in ref_driver.c). It is the checker of svtme_ois: tests/test_ois_emul.py (host emulation of the kernels'
 * arithmetic), tests/test_zzz_ois_parity_gpu.py (the kernels through the C-ABI), __graft_entry__.smoke().
 */
#include "me_oracle.h"
#include <stddef.h>
#include <string.h>

enum { OIS_PLANAR = 0, OIS_DC = 1, OIS_HOR = 10, OIS_VER = 26 };

/* neighbours of a CU in the reference's "regular" order (IntraReferenceSamplesOpenLoop_t::y_intra_reference_array after
 * UpdateNeighborSamplesArrayOpenLoop, EbIntraPrediction.c:5233-5348): r[0..2N-1] = left column bottom-up (r[2N-1] is the
 * sample left of row 0), r[2N] = top-left, r[2N+1..4N] = top row left to right. Unavailable samples are 128. */
static void ois_neighbours(const MeoPlane *pic, int x0, int y0, int n, uint8_t *r) {
    const int w = pic->width, h = pic->height;
    const uint8_t *src = pic->buf + (size_t)(pic->origin_y + y0) * pic->stride + pic->origin_x + x0;
    memset(r, 128, (size_t)4 * n + 1);
    if (x0 != 0) {
        int count = 2 * n;
        if (y0 + count > h) count -= y0 + count - h;
        for (int i = 0; i < count; i++) r[2 * n - 1 - i] = src[(ptrdiff_t)i * pic->stride - 1];
    }
    if (x0 != 0 && y0 != 0) r[2 * n] = src[-(ptrdiff_t)pic->stride - 1];
    if (y0 != 0) {
        int count = 2 * n;
        if (x0 + count > w) count -= x0 + count - w;
        memcpy(r + 2 * n + 1, src - pic->stride, (size_t)count);
    }
}

static int clip8(int v) { return v < 0 ? 0 : v > 255 ? 255 : v; }

/* One intra prediction of an n x n CU into pred (stride 64), HEVC semantics on UNFILTERED neighbours:
 * planar, DC and pure vertical / horizontal with their edge filters for n < 32 (the `skip` argument of the leaves is
 * EB_FALSE on this path), angular modes 2..34 with the 1/32-sample two-tap interpolation and the inverse-angle
 * projection of the side reference for the negative angles. */
static void ois_predict(const uint8_t *r, int n, int mode, uint8_t *pred) {
    static const int8_t ang[35] = { 0, 0, 32, 26, 21, 17, 13, 9, 5, 2, 0, -2, -5, -9, -13, -17, -21, -26, -32,
                                    -26, -21, -17, -13, -9, -5, -2, 0, 2, 5, 9, 13, 17, 21, 26, 32 };
    static const int16_t inv[35] = { 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, -4096, -1638, -910, -630, -482, -390, -315, -256,
                                     -315, -390, -482, -630, -910, -1638, -4096, 0, 0, 0, 0, 0, 0, 0, 0, 0 };
    const uint8_t *left = r + 2 * n - 1;      /* left[-y] = sample left of row y (y = -1 is the top-left) */
    const uint8_t *top = r + 2 * n + 1;       /* top[x], top[-1] = top-left */
    const int lg = n == 8 ? 3 : n == 16 ? 4 : 5;
#define LEFT(y) left[-(y)]
    if (mode == OIS_PLANAR) {
        for (int y = 0; y < n; y++)
            for (int x = 0; x < n; x++)
                pred[y * 64 + x] = (uint8_t)(((n - 1 - x) * LEFT(y) + (x + 1) * top[n] + (n - 1 - y) * top[x] + (y + 1) * LEFT(n) + n) >> (lg + 1));
        return;
    }
    if (mode == OIS_DC) {
        int sum = n;
        for (int i = 0; i < n; i++) sum += top[i] + LEFT(i);
        const int dc = sum >> (lg + 1);
        for (int y = 0; y < n; y++) for (int x = 0; x < n; x++) pred[y * 64 + x] = (uint8_t)dc;
        if (n < 32) {
            pred[0] = (uint8_t)((LEFT(0) + 2 * dc + top[0] + 2) >> 2);
            for (int x = 1; x < n; x++) pred[x] = (uint8_t)((top[x] + 3 * dc + 2) >> 2);
            for (int y = 1; y < n; y++) pred[y * 64] = (uint8_t)((LEFT(y) + 3 * dc + 2) >> 2);
        }
        return;
    }
    const int vertical = mode >= 18;
    const int a = ang[mode];
    /* main reference ref[-n .. 2n]: the top row (vertical modes) or the left column (horizontal modes), index 0 = top-left */
    uint8_t buf[3 * 32 + 1];
    uint8_t *ref = buf + 32;
    for (int i = 0; i <= 2 * n; i++) ref[i] = vertical ? top[i - 1] : (i == 0 ? top[-1] : LEFT(i - 1));
    if (a < 0) {
        const int last = (n * a) >> 5;
        for (int i = -1; i >= last; i--) {
            const int k = (i * inv[mode] + 128) >> 8;          /* 1-based index into the side reference */
            ref[i] = vertical ? LEFT(k - 1) : top[k - 1];
        }
    }
    for (int j = 0; j < n; j++) {                                /* j runs along the prediction direction's normal */
        const int pos = (j + 1) * a, idx = pos >> 5, f = pos & 31;
        for (int i = 0; i < n; i++) {
            const int v = f ? ((32 - f) * ref[i + idx + 1] + f * ref[i + idx + 2] + 16) >> 5 : ref[i + idx + 1];
            if (vertical) pred[j * 64 + i] = (uint8_t)v; else pred[i * 64 + j] = (uint8_t)v;
        }
    }
    if (a == 0 && n < 32) {                                      /* boundary smoothing of the pure vertical / horizontal modes */
        for (int i = 0; i < n; i++) {
            if (vertical) pred[i * 64] = (uint8_t)clip8(top[0] + ((LEFT(i) - top[-1]) >> 1));
            else          pred[i] = (uint8_t)clip8(LEFT(0) + ((top[i] - top[-1]) >> 1));
        }
    }
#undef LEFT
}

static uint32_t ois_sad(const MeoPlane *pic, int x0, int y0, int n, const uint8_t *pred) {
    const uint8_t *src = pic->buf + (size_t)(pic->origin_y + y0) * pic->stride + pic->origin_x + x0;
    uint32_t s = 0;
    for (int y = 0; y < n; y++)
        for (int x = 0; x < n; x++) {
            const int d = src[(size_t)y * pic->stride + x] - pred[y * 64 + x];
            s += (uint32_t)(d < 0 ? -d : d);
        }
    return s;
}

/* test hook: one prediction + its SAD for one CU */
uint32_t meo_ois_predict_cu(const MeoPlane *pic, int32_t x0, int32_t y0, int32_t n, int32_t mode, uint8_t *pred64) {
    uint8_t r[4 * 32 + 1];
    ois_neighbours(pic, x0, y0, n, r);
    ois_predict(r, n, mode, pred64);
    return ois_sad(pic, x0, y0, n, pred64);
}

/* ---- candidate lists ---------------------------------------------------------------------------------------------- */
static const uint8_t k_islice_modes[7] = { OIS_PLANAR, OIS_DC, OIS_HOR, OIS_VER, 2, 18, 34 };            /* iSliceModesArray[0..6], ME.c:84, 8132 */
static const uint8_t k_stage1_modes[9] = { OIS_HOR, OIS_VER, 2, 18, 34, 6, 14, 22, 30 };                   /* stage1ModesArray, ME.c:85 */
static const uint8_t k_points[5] = { 1, 3, 5, 7, 9 };                                                      /* numberOfOisModePoints, ME.c:68 */
static const int16_t k_point_th[3][6][4] = {                                                                 /* OisPointTh, ME.c:28-58 */
    { { -20, 50, 150, 200 }, { -20, 50, 150, 200 }, { -20, 50, 100, 150 }, { -20, 50, 200, 300 }, { -20, 50, 200, 300 }, { -20, 50, 200, 300 } },
    { { -150, 0, 150, 200 }, { -150, 0, 150, 200 }, { -125, 0, 100, 150 }, { -50, 50, 200, 300 }, { -50, 50, 200, 300 }, { -50, 50, 200, 300 } },
    { { -400, -300, -200, 0 }, { -400, -300, -200, 0 }, { -400, -300, -200, 0 }, { -400, -300, -200, 0 }, { -400, -300, -200, 0 }, { -400, -300, -200, 0 } } };

/* InjectIntraCandidatesBasedOnBestMode, ME.c:7525-7745: the list that follows each stage-1 winner, and the stage-1 slot its
 * SAD is copied from */
static const struct { uint8_t best, sad_slot, modes[9]; } k_inject[9] = {
    { 2, 2, { 2, OIS_DC, OIS_PLANAR, 3, 4, 5, 7, 8, 9 } },
    { OIS_HOR, 0, { OIS_HOR, OIS_DC, OIS_PLANAR, 9, 11, 8, 12, 7, 13 } },
    { 18, 3, { 18, OIS_DC, OIS_PLANAR, 17, 19, 16, 20, 15, 21 } },
    { OIS_VER, 1, { OIS_VER, OIS_DC, OIS_PLANAR, 25, 27, 24, 28, 23, 29 } },
    { 34, 4, { 34, OIS_DC, OIS_PLANAR, 33, 32, 29, 31, 27, 28 } },
    { 6, 5, { 6, OIS_DC, OIS_PLANAR, 7, 5, 4, 8, 3, 9 } },
    { 14, 6, { 14, OIS_DC, OIS_PLANAR, 13, 15, 12, 16, 11, 17 } },
    { 22, 7, { 22, OIS_DC, OIS_PLANAR, 21, 23, 20, 24, 19, 25 } },
    { 30, 8, { 30, OIS_DC, OIS_PLANAR, 29, 31, 28, 32, 27, 33 } } };

/* raster-scan CU index -> position and size inside the SB (RASTER_SCAN_CU_X / _Y / _SIZE, EbUtility.c) */
static void ois_cu_geometry(int cu, int *x, int *y, int *n) {
    if (cu < 5) { *n = 32; *x = ((cu - 1) & 1) * 32; *y = ((cu - 1) >> 1) * 32; }
    else if (cu < 21) { *n = 16; *x = ((cu - 5) & 3) * 16; *y = ((cu - 5) >> 2) * 16; }
    else { *n = 8; *x = ((cu - 21) & 7) * 8; *y = ((cu - 21) >> 3) * 8; }
}

/* OpenLoopIntraSearchLcu for every SB of a picture (ME.c:8047-8365, ENCODER_MODE_CLEANUP = 1).
 * me_sad: [num_sb][85] best ME distortion per square PU (me_results[sb][pu].distortionDirection[0].distortion), read on
 * the inter path only; may be NULL for intra pictures. out: [num_sb][85] (entry 0, the 64x64 CU, is never written). */
void meo_ois_picture(const MeoOisParams *prm, const MeoPlane *pic, const uint32_t *me_sad, MeoOisCu *out) {
    const int w_sb = (prm->width + 63) / 64, h_sb = (prm->height + 63) / 64;
    uint8_t r[4 * 32 + 1], pred[64 * 64];
    for (int sb = 0; sb < w_sb * h_sb; sb++) {
        const int ox = (sb % w_sb) * 64, oy = (sb / w_sb) * 64;
        uint32_t stage1[11];
        uint32_t best_mode = OIS_PLANAR;                       /* both persist from CU to CU inside an SB, like the reference's locals */
        memset(stage1, 0, sizeof(stage1));
        for (int cu = 1; cu < 85; cu++) {
            int cx, cy, n;
            ois_cu_geometry(cu, &cx, &cy, &n);
            MeoOisCu *o = &out[(size_t)sb * 85 + cu];
            for (int i = 0; i < 9; i++) o->cand[i].valid = 0;                                  /* InitValidDistortion */
            /* raster_scan_cu_validity: the CU lies completely inside the picture (EbSequenceControlSet.c:395-420) */
            if (ox + cx + n > prm->width || oy + cy + n > prm->height) continue;
            const int x0 = ox + cx, y0 = oy + cy;
            if (prm->slice_type == 0) {                                                          /* I_SLICE */
                ois_neighbours(pic, x0, y0, n, r);
                if (n == 32) {
                    ois_predict(r, n, OIS_PLANAR, pred);
                    o->cand[0].distortion = ois_sad(pic, x0, y0, n, pred);
                    o->cand[0].mode = OIS_PLANAR; o->cand[0].valid = 1;
                    continue;
                }
                uint32_t best = 32 * 32 * 255;
                for (int i = 0; i < 7; i++) {                                                    /* IntraOpenLoopSearchTheseModesOutputBest */
                    ois_predict(r, n, k_islice_modes[i], pred);
                    stage1[i] = ois_sad(pic, x0, y0, n, pred);
                    if (stage1[i] < best) { best = stage1[i]; best_mode = k_islice_modes[i]; }
                }
                int c = 0;                                                                       /* InjectIntraCandidatesBasedOnBestModeIslice */
                o->cand[0].valid = 1; o->cand[0].distortion = stage1[0];
                o->cand[c++].mode = OIS_PLANAR; o->cand[c++].mode = OIS_DC;
                if (best_mode == 2) { o->cand[c++].mode = 2; o->cand[c++].mode = 4; o->cand[c++].mode = 6; }
                else if (best_mode == OIS_HOR) { o->cand[c++].mode = OIS_HOR; o->cand[c++].mode = 6; o->cand[c++].mode = 14; }
                else if (best_mode == 18) { o->cand[c++].mode = 18; o->cand[c++].mode = 14; o->cand[c++].mode = 22; }
                else if (best_mode == OIS_VER) { o->cand[c++].mode = OIS_VER; o->cand[c++].mode = 22; o->cand[c++].mode = 30; }
                else if (best_mode != OIS_PLANAR && best_mode != OIS_DC) { o->cand[c++].mode = 34; o->cand[c++].mode = 32; o->cand[c++].mode = 30; }
                o->total = (uint8_t)c;
                continue;
            }
            if (prm->cu8x8_mode == 1 && n == 8) continue;
            const int is_4k = prm->input_resolution == 3;                                        /* INPUT_SIZE_4K_RANGE */
            const int th_set = is_4k ? ((prm->temporal_layer_index == 0 || prm->is_used_as_reference) ? 2 : 1) : 2;
            if (!prm->limit_ois_to_dc)
                ois_neighbours(pic, x0, y0, n, r);
            if (prm->temporal_layer_index == 0 && !is_4k) {
                /* all 35 modes; the best MAX_OPEN_LOOP_INTRA_CANDIDATES = 18 are kept (SortIntraModesOpenLoop: a later mode
                 * replaces the FIRST worst entry when it is strictly better), then sorted by SAD (SortOisCandidateOpenLoop:
                 * a selection sort by swaps -- its order among equal SADs is reproduced literally) */
                for (int i = 0; i < 18; i++) o->cand[i].valid = 0;
                for (int m = 0; m < 35; m++) {
                    ois_predict(r, n, m, pred);
                    const uint32_t sad = ois_sad(pic, x0, y0, n, pred);
                    if (m < 18) { o->cand[m].distortion = sad; o->cand[m].mode = (uint8_t)m; }
                    else {
                        int worst = 0;
                        for (int i = 1; i < 18; i++) if (o->cand[i].distortion > o->cand[worst].distortion) worst = i;
                        if (sad < o->cand[worst].distortion) { o->cand[worst].distortion = sad; o->cand[worst].mode = (uint8_t)m; }
                    }
                }
                for (int i = 0; i < 18; i++)
                    for (int j = i; j < 18; j++)
                        if (o->cand[i].distortion > o->cand[j].distortion) {               /* mode and SAD swap; the flags stay */
                            const uint32_t d = o->cand[i].distortion; const uint8_t m = o->cand[i].mode;
                            o->cand[i].distortion = o->cand[j].distortion; o->cand[i].mode = o->cand[j].mode;
                            o->cand[j].distortion = d; o->cand[j].mode = m;
                        }
                o->total = 18;
                continue;
            }
            if (prm->limit_ois_to_dc) {                                                          /* OpenLoopIntraDC */
                ois_neighbours(pic, x0, y0, n, r);
                ois_predict(r, n, OIS_DC, pred);
                o->cand[0].distortion = ois_sad(pic, x0, y0, n, pred);
                o->cand[0].mode = OIS_DC; o->cand[0].valid = 1; o->total = 1;
                continue;
            }
            /* inter pictures above layer 0 (or 4K): DC first, then a number of stage-1 directions that grows with how much
             * better intra DC looks than the ME result of the same square */
            const uint32_t me = me_sad[(size_t)sb * 85 + cu];
            ois_predict(r, n, OIS_DC, pred);
            stage1[0] = ois_sad(pic, x0, y0, n, pred);
            const int32_t diff = (int32_t)(me - stage1[0]) * 100;
            const int32_t dist = stage1[0] ? diff / (int32_t)stage1[0] : 0;
            const int16_t *th = k_point_th[th_set][prm->temporal_layer_index];
            int point = 4;                                                                       /* OIS_VERY_COMPLEX_MODE */
            if (stage1[0] == 0 || me == 0 || dist <= th[0]) point = 0;
            else if (dist <= th[1]) point = 1;
            else if (dist <= th[2]) point = 2;
            else if (dist <= th[3]) point = 3;
            if (point == 0) {
                o->cand[0].mode = OIS_DC; o->cand[0].distortion = stage1[0]; o->total = 1;      /* (valid stays false) */
                continue;
            }
            uint32_t best = 32 * 32 * 255;
            for (int i = 0; i < k_points[point]; i++) {
                ois_predict(r, n, k_stage1_modes[i], pred);
                stage1[i] = ois_sad(pic, x0, y0, n, pred);
                if (stage1[i] < best) { best = stage1[i]; best_mode = k_stage1_modes[i]; }
            }
            int e = 8;                                                                           /* default: the list of mode 30 */
            for (int i = 0; i < 9; i++) if (k_inject[i].best == best_mode) e = i;
            o->cand[0].distortion = stage1[k_inject[e].sad_slot];
            /* quirk: with ENCODER_MODE_CLEANUP every case of the reference's switch sets valid = (layer > 1) -- except the case
             * of the vertical winner, which kept its enc-mode test (ME.c:7604-7612): there it is TRUE above enc-mode 1 */
            o->cand[0].valid = (best_mode == OIS_VER && prm->enc_mode > 1) ? 1 : prm->temporal_layer_index > 1;
            for (int i = 0; i < 9; i++) o->cand[i].mode = k_inject[e].modes[i];
            o->total = point == 4 ? 9 : k_points[point];                                         /* intraSearchInMd[point][depth 1..3] */
        }
    }
}
