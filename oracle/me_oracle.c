/* oracle/me_oracle.c -- TEST INFRASTRUCTURE ONLY (checker; never shipped, never the thing measured).
 *
 * Plain-C restatement of SVT-AV1 (v0.4.0) open-loop motion estimation for one picture:
 *   pyramid    : EbPictureAnalysisProcess.c:100-125 (Decimation2D), :4866-4937; EbMcp.c:173-216
 *   per SB     : EbMotionEstimationProcess.c:481-555 (staging) -> EbMotionEstimation.c:6152-7162
 * as it behaves when dispatched with asm_type = ASM_AVX2 and built by gcc (the only runnable
 * configuration, EbEncHandle.c:2535-2539), including the arithmetic quirks of the SIMD leaves:
 *   Q1  left/top search-area clipping moves the origin but never shrinks the size
 *       (the second ternary re-tests the already-corrected origin)       ME.c:4378-4384, 6690-6696
 *   Q2  hme_mv_center_check evaluates position A at the zero-MV address   ME.c:5953-5959
 *   Q3  HME level 1 (32x16 block) keeps 16-bit partial sums and uses the min-subtraction trick:
 *       value = min(lo+hi, 65535 + min_group(lo))     ASM_AVX2/EbComputeSAD_Intrinsic_AVX2.c:2341-2363
 *   Q4  32x32 best-MV x offset is wrong by +-4 px under gcc and ties prefer positions 4..7
 *                                                     ASM_AVX2/EbComputeSAD_Intrinsic_AVX2.c:3845-3888
 *   Q5  half-pel candidates are ranked with an 8-bit *wrapping* "SSD"
 *                                                     ASM_SSE4_1/EbPictureOperators_Intrinsic_SSE4_1.c:534-630
 *   Q6  64x64 quarter-pel refinement looks at the top-left 32x32 only        ME.c:3395-3409
 *   Q7  the per-region HME centre arrays are initialised for list 0 only (the region counters are
 *       not reset before the initialisation loop of list 1), so a disabled HME level hands list 1
 *       the list-0 values                                                    ME.c:6323-6345
 * PARITY PIN: checked bit-exactly against oracle/_ref/libsvtref.so (the reference's own objects)
 * by tests/test_oracle_vs_ref.py and against tests/golden/ vectors generated from it.
 *
 * Supported geometry: width, height multiples of 8; every width except width % 64 in {40, 56}: for those two
 * edge-SB widths the reference dispatches a void function in hme_mv_center_check (EbComputeSAD.h:143-147) and its
 * result is undefined. Edge SBs 8, 16, 24, 32 and 48 wide run the leaves the reference picks for them (ME.c:4417-4496,
 * 4579-4615, 4711-4750): the C SadLoopKernel (C_DEFAULT/EbComputeSAD_C.c:72-120) for the block widths 2, 6, 12 of the
 * decimated levels, the AVX2 cases 4 / 8 / 16 / 24 / 48 otherwise, with two more leaf quirks:
 *   Q12 4-wide blocks, height % 4 == 0 (HME level 0 of 16-wide SBs, level 1 of 8-wide SBs): the 8-position groups of
 *       the AVX2 kernels pair reference row 2 with source row 0 and reference row 0 with source row 2 inside every
 *       group of four rows (the 128-bit halves of ss0 are swapped); the leftover (< 8 positions) code is right
 *                                                     ASM_AVX2/EbComputeSAD_Intrinsic_AVX2.c:1650-1652, 3973-3975
 *   Q13 24-wide blocks, height <= 16 (HME level 1 of 48-wide SBs): 16-bit sums with the min-subtraction trick of Q3,
 *       value = min(a + b, 65535 + min_group(a)), a = columns 0-7 and 16-23, b = columns 8-15
 *                                                     ASM_AVX2/EbComputeSAD_Intrinsic_AVX2.c:2123-2160
 * Search areas up to 127 x 127 (MIN(search_area, 127), ME.c:6644-6645).
 *   Q10 width-8 PUs taller than 8 (8x16, 8x32; 209-PU mode): the half-pel "SSD" kernel picked by
 *       Log2f(width) is the fixed 8x8 one, so only the top eight rows are ranked   ME.c:1909-1916,
 *                                                     ASM_SSE4_1/EbPictureOperators_Intrinsic_SSE4_1.c:534-569
 *   Q11 209-PU full-pel search: the update test of 32x16 PU #5 (internal order) reads the stale
 *       variable `sad` (= the bottom 64x32 sum) but stores its own sum             ME.c:363-367
 * 85-PU (square, FullPelSearch_LCU) and 209-PU (open_loop_me_fullpel_search_sblock) modes.
 */
#include "me_oracle.h"
#include <stdlib.h>
#include <string.h>

#define MAX_SAD_VALUE (128 * 128 * 255)           /* EbMotionEstimation.h:78 */
#define MIN(a, b) ((a) < (b) ? (a) : (b))
#define MAX(a, b) ((a) > (b) ? (a) : (b))

/* ------------------------------------------------------------------ planes / params */
int64_t meo_plane_bytes(int32_t width, int32_t height, int32_t pad) {
    return (int64_t)(width + 2 * pad) * (height + 2 * pad);
}

void meo_plane_init(MeoPlane *p, uint8_t *mem, int32_t width, int32_t height, int32_t pad) {
    p->buf = mem; p->stride = width + 2 * pad;
    p->origin_x = pad; p->origin_y = pad; p->width = width; p->height = height;
}

/* EbDefinitions.h:3320-3469, column M0 (EbMotionEstimationProcess.c:102) */
static const int16_t T_L0_TOTAL_W[5] = { 48, 64, 96, 96, 128 }, T_L0_W[5] = { 24, 32, 48, 48, 64 };
static const int16_t T_L0_TOTAL_H[5] = { 40, 48, 48, 48, 80 },  T_L0_H[5] = { 20, 24, 24, 24, 40 };
/* HME enables come from column enc_mode (=3, default): only 4K level 2 is off
 * (EbResourceCoordinationProcess.c:366-380, EbDefinitions.h:3416-3422) */
static const int8_t  T_EN_L2_M3[5] = { 1, 1, 1, 1, 0 };
/* EbDefinitions.h:2980-2996 */
static const int16_t T_L0_MULT[6][6] = {
    { 100 }, { 100, 100 }, { 100, 100, 100 }, { 200, 140, 100, 70 },
    { 350, 200, 100, 100, 100 }, { 525, 350, 200, 100, 100, 100 } };

void meo_default_params(MeoParams *p, int32_t width, int32_t height, int32_t num_lists,
                        int32_t temporal_layer_index, int32_t hierarchical_levels, int32_t is_ref) {
    /* derive_input_resolution, EbSequenceControlSet.c:332-346; thresholds EbDefinitions.h:1824-1827 */
    uint32_t size = (uint32_t)width * (uint32_t)height;
    int ir = size < 0xB71B0 ? 0 : size < 0x1AB3F0 ? 1 : size < 0x29F630 ? 2 : 3;
    uint32_t ratio = (uint32_t)width / (uint32_t)height;
    int ri = ir <= 0 ? 0 : (ir <= 1 && ratio < 2) ? 1 : (ir <= 1 && ratio > 3) ? 2 : (ir <= 2) ? 3 : 4;
    memset(p, 0, sizeof(*p));
    p->width = width; p->height = height; p->num_lists = num_lists;
    p->temporal_layer_index = temporal_layer_index; p->hierarchical_levels = hierarchical_levels;
    p->is_used_as_reference = is_ref;
    p->enable_hme_flag = 1; p->enable_hme_level0 = 1; p->enable_hme_level1 = 1;
    p->enable_hme_level2 = T_EN_L2_M3[ri];
    p->use_subpel = 1; p->num_pus = MEO_SQUARE_PU_COUNT; p->cu8x8_mode = 0;
    p->num_hme_regions_w = 2; p->num_hme_regions_h = 2;
    p->hme_l0_total_w = T_L0_TOTAL_W[ri]; p->hme_l0_total_h = T_L0_TOTAL_H[ri];
    for (int i = 0; i < 2; i++) {
        p->hme_l0_w[i] = T_L0_W[ri]; p->hme_l0_h[i] = T_L0_H[ri];
        p->hme_l1_w[i] = 16; p->hme_l1_h[i] = 16;
        p->hme_l2_w[i] = 8;  p->hme_l2_h[i] = 8;
    }
    p->search_area_w = 64; p->search_area_h = 64;
    p->hme_l0_mult_x = T_L0_MULT[hierarchical_levels][temporal_layer_index];
    p->hme_l0_mult_y = T_L0_MULT[hierarchical_levels][temporal_layer_index];
}

/* ------------------------------------------------------------------ pyramid */
/* generate_padding, EbMcp.c:173-216: rows first (left/right replicate), then whole padded rows
 * are copied up and down. */
static void pad_plane(MeoPlane *p) {
    uint8_t *row = p->buf + p->origin_x + (size_t)p->origin_y * p->stride;
    for (int y = 0; y < p->height; y++, row += p->stride) {
        memset(row - p->origin_x, row[0], p->origin_x);
        memset(row + p->width, row[p->width - 1], p->origin_x);
    }
    uint8_t *top = p->buf + (size_t)p->origin_y * p->stride;
    uint8_t *bot = p->buf + (size_t)(p->origin_y + p->height - 1) * p->stride;
    for (int i = 1; i <= p->origin_y; i++) {
        memcpy(top - (size_t)i * p->stride, top, p->stride);
        memcpy(bot + (size_t)i * p->stride, bot, p->stride);
    }
}

/* Decimation2D, EbPictureAnalysisProcess.c:100-125: plain point sampling, no filter. */
static void decimate(const MeoPlane *f, MeoPlane *d, int step) {
    const uint8_t *src = f->buf + f->origin_x + (size_t)f->origin_y * f->stride;
    uint8_t *dst = d->buf + d->origin_x + (size_t)d->origin_y * d->stride;
    for (int y = 0; y < f->height; y += step) {
        for (int x = 0; x < f->width; x += step) dst[x / step] = src[x];
        src += (size_t)f->stride * step; dst += d->stride;
    }
}

void meo_build_pyramid(const uint8_t *luma, int32_t src_stride, MeoPyramid *pyr) {
    MeoPlane *f = &pyr->full;
    for (int y = 0; y < f->height; y++)
        memcpy(f->buf + (size_t)(f->origin_y + y) * f->stride + f->origin_x, luma + (size_t)y * src_stride, f->width);
    pad_plane(f);
    decimate(f, &pyr->quarter, 2);   pad_plane(&pyr->quarter);
    decimate(f, &pyr->sixteenth, 4); pad_plane(&pyr->sixteenth);
}

/* ------------------------------------------------------------------ SAD primitives */
static inline uint8_t *px(const MeoPlane *p, int x, int y) {
    return p->buf + (size_t)(p->origin_y + y) * p->stride + p->origin_x + x;
}

/* NxMSadKernel (AVX2 Compute{8..64}xMSad, exact 32-bit):  EbComputeSAD_Intrinsic_AVX2.c:3217-3406 */
static uint32_t sad_wxh(const uint8_t *s, int ss, const uint8_t *r, int rs, int w, int h) {
    uint32_t acc = 0;
    for (int y = 0; y < h; y++, s += ss, r += rs)
        for (int x = 0; x < w; x++) acc += (uint32_t)abs((int)s[x] - (int)r[x]);
    return acc;
}

/* The four-step edge correction used by every search (ME.c:4378-4411, 4541-4573, 4674-4707,
 * 6690-6723), kept in the reference's statement order because of quirk Q1. */
static void clip_search_area(int org, int *sao, int *sa, int pad, int dim) {
    int o = *sao, a = *sa;
    o = ((org + o) < -pad) ? -pad - org : o;
    a = ((org + o) < -pad) ? a - (-pad - (org + o)) : a;            /* never true after the line above */
    o = ((org + o) > dim - 1) ? o - ((org + o) - (dim - 1)) : o;
    a = ((org + o + a) > dim) ? MAX(1, a - ((org + o + a) - dim)) : a;
    *sao = o; *sa = a;
}

/* Edge correction of a single candidate centre (ME.c:5935-5950, 5504-5518). */
static int clip_center(int org, int c, int pad, int dim) {
    c = ((org + c) < -pad) ? -pad - org : c;
    c = ((org + c) > dim - 1) ? c - ((org + c) - (dim - 1)) : c;
    return c;
}

/* SadLoopKernel_AVX2_INTRIN / _HmeL0_ (EbComputeSAD_Intrinsic_AVX2.c:1609-3182, 3940-4139):
 * exhaustive search, positions visited row by row, eight at a time; a group replaces the
 * best only when strictly smaller, and inside a group the lowest index wins. The C SadLoopKernel
 * (C_DEFAULT/EbComputeSAD_C.c:72-120: one position at a time, strict '<') is the same argmin.
 *   quirk = SL_Q3  reproduces the 32-wide, <=16-row branch (Q3), SL_Q13 the 24-wide one (Q13), SL_Q12 the swapped
 *   rows of the 4-wide kernels (Q12, full groups only); everything else used by the
 *   ME path is an exact 32-bit argmin (the 16-bit overflow branches recompute exactly). */
enum { SL_EXACT = 0, SL_Q3 = 1, SL_Q13 = 2, SL_Q12 = 3 };
static uint32_t sad4_rows_swapped(const uint8_t *s, int ss, const uint8_t *r, int rs, int h) {
    uint32_t acc = 0;
    for (int y = 0; y < h; y++) {
        const int ry = (y & 3) == 0 ? y + 2 : (y & 3) == 2 ? y - 2 : y;     /* reference rows 0 and 2 of every four trade places */
        for (int x = 0; x < 4; x++) acc += (uint32_t)abs((int)s[y * ss + x] - (int)r[ry * rs + x]);
    }
    return acc;
}
static void sad_loop(const uint8_t *src, int ss, const uint8_t *ref, int rs, int h, int w,
                     int ref_row_stride, int saw, int sah, int quirk,
                     uint64_t *best_sad, int *best_x, int *best_y) {
    uint32_t low = 0xffffff; int bx = 0, by = 0;
    for (int i = 0; i < sah; i++) {
        const uint8_t *rrow = ref + (size_t)i * ref_row_stride;
        for (int j = 0; j < saw; j += 8) {
            int n = MIN(8, saw - j);
            uint32_t v[8];
            if (quirk == SL_Q3 || quirk == SL_Q13) {
                uint32_t lo[8], hi[8], mlo = 65535;
                for (int k = 0; k < 8; k++) {
                    if (k >= n) lo[k] = hi[k] = 65535;     /* _mm_or_si128(s, s8) */
                    else if (quirk == SL_Q3) {
                        lo[k] = sad_wxh(src, ss, rrow + j + k, rs, 16, h);
                        hi[k] = sad_wxh(src + 16, ss, rrow + j + k + 16, rs, 16, h);
                    } else {
                        lo[k] = sad_wxh(src, ss, rrow + j + k, rs, 8, h) + sad_wxh(src + 16, ss, rrow + j + k + 16, rs, 8, h);
                        hi[k] = sad_wxh(src + 8, ss, rrow + j + k + 8, rs, 8, h);
                    }
                    mlo = MIN(mlo, lo[k]);
                }
                for (int k = 0; k < 8; k++) v[k] = MIN(lo[k] + hi[k], 65535 + mlo);
                n = 8;                                     /* masked lanes take part in minpos */
            } else if (quirk == SL_Q12 && n == 8) {
                for (int k = 0; k < 8; k++) v[k] = sad4_rows_swapped(src, ss, rrow + j + k, rs, h);
            } else {
                for (int k = 0; k < n; k++) v[k] = sad_wxh(src, ss, rrow + j + k, rs, w, h);
            }
            int kb = 0;
            for (int k = 1; k < n; k++) if (v[k] < v[kb]) kb = k;
            if (v[kb] < low) { low = v[kb]; bx = j + kb; by = i; }
        }
    }
    *best_sad = low; *best_x = bx; *best_y = by;
}
/* which leaf quirk applies to a block of w x h (h = rows actually summed) */
static int sad_loop_quirk(int w, int h) {
    if (w == 32 && h <= 16) return SL_Q3;
    if (w == 24 && h <= 16) return SL_Q13;
    if (w == 4 && (h & 3) == 0) return SL_Q12;
    return SL_EXACT;
}

/* ------------------------------------------------------------------ per-SB state */
typedef struct SbCtx {
    const MeoParams *prm;
    const MeoPyramid *cur, *ref[2];
    int ox, oy, sb_w, sb_h;                 /* SB origin / clipped size */
    const uint8_t *src; int src_stride;     /* sb_src_ptr: padded current picture at the SB */
    uint8_t  quarter_sb[32 * 32];           /* quarter_sb_buffer (persists across SBs like the reference's) */
    uint8_t  sixteenth_sb[16 * 16];         /* sixteenth_sb_buffer: every 2nd row, stride 16 */
    int      sa_x[2], sa_y[2], sa_w[2], sa_h[2];
    uint32_t best_sad[2][MEO_MAX_PU_COUNT], best_mv[2][MEO_MAX_PU_COUNT];   /* internal order, like p_sb_best_sad */
    uint32_t best_ssd[2][MEO_MAX_PU_COUNT];
    uint8_t  sub_dir[MEO_MAX_PU_COUNT];  /* psub_pel_direction*, shared by both lists like the reference */
    uint32_t bipred_sad[MEO_MAX_PU_COUNT];
    int      hx[3][2][2], hy[3][2][2];      /* per-region HME centres of level 0/1/2 (live across lists: Q7) */
    uint64_t hs[3][2][2];
} SbCtx;

static inline int16_t mvx(uint32_t mv) { return (int16_t)(mv & 0xFFFF); }
static inline int16_t mvy(uint32_t mv) { return (int16_t)(mv >> 16); }
static inline uint32_t mkmv(int x, int y) { return ((uint32_t)(uint16_t)y << 16) | (uint16_t)x; }

/* 64 x (sb_h/2) SAD on even rows, doubled: the sub-sampled SB SAD used for centre decisions. */
static uint64_t sb_sub_sad(const SbCtx *c, const MeoPlane *rp, int dx, int dy) {
    return (uint64_t)sad_wxh(c->src, c->src_stride * 2, px(rp, c->ox + dx, c->oy + dy), rp->stride * 2,
                             c->sb_w, c->sb_h >> 1) << 1;
}

/* hme_mv_center_check, ME.c:5882-6145 */
static void hme_center_check(const SbCtx *c, int list, int *xsc, int *ysc) {
    const MeoPlane *rp = &c->ref[list]->full;
    const int tw = c->prm->hme_l0_total_w, th = c->prm->hme_l0_total_h, pad = 63;
    uint64_t zero = sb_sub_sad(c, rp, 0, 0);
    uint64_t a = zero;                                              /* Q2 */
    uint64_t b = sb_sub_sad(c, rp, clip_center(c->ox, tw, pad, rp->width), clip_center(c->oy, 0, pad, rp->height));
    uint64_t cc = sb_sub_sad(c, rp, clip_center(c->ox, 0, pad, rp->width), clip_center(c->oy, -th, pad, rp->height));
    uint64_t d = sb_sub_sad(c, rp, clip_center(c->ox, 0, pad, rp->width), clip_center(c->oy, th, pad, rp->height));
    uint64_t direct = 0xFFFFFFFFFFFFFull >> 8;                      /* compared after << COST_PRECISION */
    int dxv = 0, dyv = 0;
    if (list == 1) {
        dxv = 0 - (mvx(c->best_mv[0][0]) >> 2);
        dyv = 0 - (mvy(c->best_mv[0][0]) >> 2);
        dxv = (int16_t)dxv; dyv = (int16_t)dyv;
        direct = sb_sub_sad(c, rp, clip_center(c->ox, dxv, pad, rp->width), clip_center(c->oy, dyv, pad, rp->height));
    }
    uint64_t best = MIN(zero, MIN(a, MIN(b, MIN(cc, MIN(d, direct)))));
    if (best == zero)        { *xsc = 0;   *ysc = 0; }
    else if (best == a)      { *xsc = -tw; *ysc = 0; }
    else if (best == b)      { *xsc = tw;  *ysc = 0; }
    else if (best == cc)     { *xsc = 0;   *ysc = -th; }
    else if (best == direct) { *xsc = dxv; *ysc = dyv; }
    else                     { *xsc = 0;   *ysc = th; }
}

/* HmeLevel0, ME.c:4306-4503 */
static void hme_level0(SbCtx *c, int list, int rw, int rh, int xc, int yc,
                       uint64_t *sad, int *xo, int *yo) {
    const MeoParams *p = c->prm; const MeoPlane *rp = &c->ref[list]->sixteenth;
    int ox = c->ox >> 2, oy = c->oy >> 2, w = c->sb_w >> 2, h = c->sb_h >> 2;
    int saw = (int16_t)((p->hme_l0_w[rw] * p->hme_l0_mult_x) / 100);
    int sah = (int16_t)((p->hme_l0_h[rh] * p->hme_l0_mult_y) / 100);
    int xd = xc, yd = yc;
    for (int i = rw; i > 0; i--) xd += (int16_t)((p->hme_l0_w[i - 1] * p->hme_l0_mult_x) / 100);
    for (int i = rh; i > 0; i--) yd += (int16_t)((p->hme_l0_h[i - 1] * p->hme_l0_mult_y) / 100);
    int sax = -(int16_t)(((p->hme_l0_total_w * p->hme_l0_mult_x) / 100) >> 1) + xd;
    int say = -(int16_t)(((p->hme_l0_total_h * p->hme_l0_mult_y) / 100) >> 1) + yd;
    clip_search_area(ox, &sax, &saw, rp->origin_x - 1, rp->width);
    clip_search_area(oy, &say, &sah, rp->origin_y - 1, rp->height);
    int bx, by;
    sad_loop(c->sixteenth_sb, 16, px(rp, ox + sax, oy + say), rp->stride * 2, h >> 1, w,
             rp->stride, saw, sah, sad_loop_quirk(w, h >> 1), sad, &bx, &by);
    *sad *= 2;
    *xo = (int16_t)((bx + sax) * 4); *yo = (int16_t)((by + say) * 4);
}

static int round_sa_width(int w) { return (w < 8) ? 8 : (w & 7) ? w + (w - ((w >> 3) << 3)) : w; }

/* HmeLevel1, ME.c:4505-4625 */
static void hme_level1(SbCtx *c, int list, int rw, int rh, int xc, int yc,
                       uint64_t *sad, int *xo, int *yo) {
    const MeoParams *p = c->prm; const MeoPlane *rp = &c->ref[list]->quarter;
    int ox = c->ox >> 1, oy = c->oy >> 1, w = c->sb_w >> 1, h = c->sb_h >> 1;
    int saw = round_sa_width(p->hme_l1_w[rw]), sah = p->hme_l1_h[rh];
    int sax = -(saw >> 1) + xc, say = -(sah >> 1) + yc;
    clip_search_area(ox, &sax, &saw, rp->origin_x - 1, rp->width);
    clip_search_area(oy, &say, &sah, rp->origin_y - 1, rp->height);
    int bx, by;
    sad_loop(c->quarter_sb, 32 * 2, px(rp, ox + sax, oy + say), rp->stride * 2, h >> 1, w,
             rp->stride, saw, sah, sad_loop_quirk(w, h >> 1), sad, &bx, &by);
    *sad *= 2;
    *xo = (int16_t)((bx + sax) * 2); *yo = (int16_t)((by + say) * 2);
}

/* HmeLevel2, ME.c:4627-4758 */
static void hme_level2(SbCtx *c, int list, int rw, int rh, int xc, int yc,
                       uint64_t *sad, int *xo, int *yo) {
    const MeoParams *p = c->prm; const MeoPlane *rp = &c->ref[list]->full;
    int saw = round_sa_width(p->hme_l2_w[rw]), sah = p->hme_l2_h[rh];
    int sax = -(saw >> 1) + xc, say = -(sah >> 1) + yc;
    clip_search_area(c->ox, &sax, &saw, 63, rp->width);
    clip_search_area(c->oy, &say, &sah, 63, rp->height);
    int bx, by;
    sad_loop(c->src, c->src_stride * 2, px(rp, c->ox + sax, c->oy + say), rp->stride * 2, c->sb_h >> 1, c->sb_w,
             rp->stride, saw, sah, 0, sad, &bx, &by);
    *sad *= 2;
    *xo = (int16_t)(bx + sax); *yo = (int16_t)(by + say);
}

/* The HME part of MotionEstimateLcu, ME.c:6300-6642: returns the full-pel search centre. */
static void find_search_center(SbCtx *c, int list, int *xsc, int *ysc) {
    const MeoParams *p = c->prm;
    *xsc = 0; *ysc = 0;
    if (!(p->temporal_layer_index > 0 || list == 0)) return;
    hme_center_check(c, list, xsc, ysc);
    if (!(p->enable_hme_flag && c->sb_h == 64)) return;

    const int nw = p->num_hme_regions_w, nh = p->num_hme_regions_h;
    int (*x0)[2] = c->hx[0], (*y0)[2] = c->hy[0], (*x1)[2] = c->hx[1], (*y1)[2] = c->hy[1], (*x2)[2] = c->hx[2], (*y2)[2] = c->hy[2];
    uint64_t (*s0)[2] = c->hs[0], (*s1)[2] = c->hs[1], (*s2)[2] = c->hs[2];
    if (list == 0)                                                  /* Q7 */
        for (int h = 0; h < nh; h++) for (int w = 0; w < nw; w++) {
            x0[w][h] = x1[w][h] = x2[w][h] = *xsc; y0[w][h] = y1[w][h] = y2[w][h] = *ysc;
            s0[w][h] = s1[w][h] = s2[w][h] = 0;
        }
    if (p->enable_hme_level0)
        for (int h = 0; h < nh; h++) for (int w = 0; w < nw; w++)
            hme_level0(c, list, w, h, (int16_t)*xsc >> 2, (int16_t)*ysc >> 2, &s0[w][h], &x0[w][h], &y0[w][h]);
    if (p->enable_hme_level1)
        for (int h = 0; h < nh; h++) for (int w = 0; w < nw; w++)
            hme_level1(c, list, w, h, (int16_t)x0[w][h] >> 1, (int16_t)y0[w][h] >> 1, &s1[w][h], &x1[w][h], &y1[w][h]);
    if (p->enable_hme_level2)
        for (int h = 0; h < nh; h++) for (int w = 0; w < nw; w++)
            hme_level2(c, list, w, h, x1[w][h], y1[w][h], &s2[w][h], &x2[w][h], &y2[w][h]);

    int xh = 0, yh = 0; uint64_t sh;
#define PICK_MIN(X, Y, S)                                                             \
    do { xh = X[0][0]; yh = Y[0][0]; sh = S[0][0];                                    \
         for (int h = 0; h < nh; h++) for (int w = (h == 0 ? 1 : 0); w < nw; w++)     \
             if (S[w][h] < sh) { xh = X[w][h]; yh = Y[w][h]; sh = S[w][h]; } } while (0)
    if (p->enable_hme_level0 && !p->enable_hme_level1 && !p->enable_hme_level2) PICK_MIN(x0, y0, s0);
    if (p->enable_hme_level1 && !p->enable_hme_level2) PICK_MIN(x1, y1, s1);
    if (p->enable_hme_level2) {
        PICK_MIN(x2, y2, s2);
        const int tot = nw * nh;
        if (p->same_ref_poc && list == 1 && tot > 1) {          /* ME.c:6606-6631 */
            for (int q = 0; q < tot - 1; q++) for (int n = q + 1; n < tot; n++) {
                int qa = q / nw, qb = q % nw, na = n / nw, nb = n % nw;
                if (s2[qa][qb] > s2[na][nb]) {
                    int tx = x2[qa][qb], ty = y2[qa][qb]; uint64_t ts = s2[qa][qb];
                    x2[qa][qb] = x2[na][nb]; y2[qa][qb] = y2[na][nb]; s2[qa][qb] = s2[na][nb];
                    x2[na][nb] = tx; y2[na][nb] = ty; s2[na][nb] = ts;
                }
            }
            xh = x2[0][1]; yh = y2[0][1];
        }
    }
#undef PICK_MIN
    (void)sh;
    *xsc = xh; *ysc = yh;
}

/* ------------------------------------------------------------------ integer full-pel search */
/* EbMotionEstimation.h:89-128 : raster index -> internal (z-order) index */
static const uint8_t TAB16[16] = { 0, 1, 4, 5, 2, 3, 6, 7, 8, 9, 12, 13, 10, 11, 14, 15 };
static const uint8_t TAB8[64] = {
    0, 1, 4, 5, 16, 17, 20, 21, 2, 3, 6, 7, 18, 19, 22, 23, 8, 9, 12, 13, 24, 25, 28, 29, 10, 11, 14, 15, 26, 27, 30, 31,
    32, 33, 36, 37, 48, 49, 52, 53, 34, 35, 38, 39, 50, 51, 54, 55, 40, 41, 44, 45, 56, 57, 60, 61, 42, 43, 46, 47, 58, 59, 62, 63 };

/* Sub-sampled SADs of one search position: 8x8 on rows 0,2,4,6 (un-doubled), then sums.
 * s8 / s16 are in internal order; s32[4] raster; returns 64x64. */
static uint32_t position_sads(const SbCtx *c, const uint8_t *r, int rs, uint32_t s8[64], uint32_t s16[16], uint32_t s32[4]) {
    uint32_t raw8[64];
    for (int by = 0; by < 8; by++) for (int bx = 0; bx < 8; bx++)
        raw8[by * 8 + bx] = sad_wxh(c->src + (size_t)by * 8 * c->src_stride + bx * 8, c->src_stride * 2,
                                    r + (size_t)by * 8 * rs + bx * 8, rs * 2, 8, 4);
    for (int i = 0; i < 64; i++) s8[TAB8[i]] = raw8[i];
    uint32_t raw16[16];
    for (int by = 0; by < 4; by++) for (int bx = 0; bx < 4; bx++) {
        int b = (by * 2) * 8 + bx * 2;
        raw16[by * 4 + bx] = raw8[b] + raw8[b + 1] + raw8[b + 8] + raw8[b + 9];
    }
    for (int i = 0; i < 16; i++) s16[TAB16[i]] = raw16[i];
    for (int by = 0; by < 2; by++) for (int bx = 0; bx < 2; bx++) {
        int b = (by * 2) * 4 + bx * 2;
        s32[by * 2 + bx] = raw16[b] + raw16[b + 1] + raw16[b + 4] + raw16[b + 5];
    }
    return s32[0] + s32[1] + s32[2] + s32[3];
}

/* FullPelSearch_LCU, ME.c:1504-1551, with the 8-position kernels
 * (EbComputeSAD_Intrinsic_AVX2.c:3414-3932) and the single-position SSE2 ones
 * (EbMeSadCalculation_Intrinsic_SSE2.c:10-127) for the width & 7 leftovers. */
static void full_pel_search(SbCtx *c, int list) {
    const MeoPlane *rp = &c->ref[list]->full;
    const int sax = c->sa_x[list], say = c->sa_y[list], saw = c->sa_w[list], sah = c->sa_h[list];
    uint32_t *bs = c->best_sad[list], *bm = c->best_mv[list];
    for (int i = 0; i < 21; i++) bs[i] = MAX_SAD_VALUE;          /* InitializeBuffer_32bits(.., 21, 1, ..) = 85 words */
    for (int i = 21; i < MEO_SQUARE_PU_COUNT; i++) bs[i] = MAX_SAD_VALUE;
    uint32_t *b64 = bs, *b32 = bs + 1, *b16 = bs + 5, *b8 = bs + 21;
    uint32_t *m64 = bm, *m32 = bm + 1, *m16 = bm + 5, *m8 = bm + 21;
    const int mult8 = saw & ~7;
    for (int ys = 0; ys < sah; ys++) {
        const uint8_t *rrow = px(rp, c->ox + sax, c->oy + say + ys);
        const int ymv = (int16_t)((say + ys) * 4);
        for (int xs = 0; xs < mult8; xs += 8) {
            uint32_t s8[8][64], s16[8][16], s32[8][4], s64[8];
            for (int k = 0; k < 8; k++) s64[k] = position_sads(c, rrow + xs + k, rp->stride, s8[k], s16[k], s32[k]);
            const int xmv = (int16_t)((sax + xs) * 4);
            for (int b = 0; b < 64; b++) {                       /* minpos, then signed "best > new" */
                int kb = 0; for (int k = 1; k < 8; k++) if (s8[k][b] < s8[kb][b]) kb = k;
                if ((s8[kb][b] << 1) < b8[b]) { b8[b] = s8[kb][b] << 1; m8[b] = mkmv(xmv + kb * 4, ymv); }
            }
            for (int b = 0; b < 16; b++) {
                int kb = 0; for (int k = 1; k < 8; k++) if (s16[k][b] < s16[kb][b]) kb = k;
                if ((s16[kb][b] << 1) < b16[b]) { b16[b] = s16[kb][b] << 1; m16[b] = mkmv(xmv + kb * 4, ymv); }
            }
            {   /* 64x64: eight sequential strict compares */
                uint32_t best = b64[0]; int kb = 0;
                for (int k = 0; k < 8; k++) if ((s64[k] << 1) < best) { best = s64[k] << 1; kb = k; }
                if (b64[0] != best) { b64[0] = best; m64[0] = mkmv(xmv + kb * 4, ymv); }
            }
            for (int b = 0; b < 4; b++) {                        /* 32x32: quirk Q4 */
                int lo = 0, hi = 4;
                for (int k = 1; k < 4; k++) if (s32[k][b] < s32[lo][b]) lo = k;
                for (int k = 5; k < 8; k++) if (s32[k][b] < s32[hi][b]) hi = k;
                uint32_t v; int idx;
                if (s32[hi][b] > s32[lo][b]) { v = s32[lo][b]; idx = lo + 4; }     /* low half wins: offset +4 */
                else                         { v = s32[hi][b]; idx = hi - 4; }     /* high half (or tie): offset -4 */
                if ((v << 1) < b32[b]) { b32[b] = v << 1; m32[b] = mkmv(xmv + idx * 4, ymv); }
            }
        }
        for (int xs = mult8; xs < saw; xs++) {
            uint32_t s8[64], s16[16], s32[4], s64;
            s64 = position_sads(c, rrow + xs, rp->stride, s8, s16, s32);
            const uint32_t mv = mkmv((int16_t)((sax + xs) * 4), ymv);
            for (int b = 0; b < 64; b++) if ((s8[b] << 1) < b8[b]) { b8[b] = s8[b] << 1; m8[b] = mv; }
            for (int b = 0; b < 16; b++) if ((s16[b] << 1) < b16[b]) { b16[b] = s16[b] << 1; m16[b] = mv; }
            for (int b = 0; b < 4; b++) if ((s32[b] << 1) < b32[b]) { b32[b] = s32[b] << 1; m32[b] = mv; }
            if ((s64 << 1) < b64[0]) { b64[0] = s64 << 1; m64[0] = mv; }
        }
    }
}

/* ------------------------------------------------------------------ 209-PU mode: PU classes */
/* EbMotionEstimation.h:123-175 (raster index inside a class -> internal index inside the class) */
static const uint8_t TAB_ID[4] = { 0, 1, 2, 3 };
static const uint8_t TAB32x16[8] = { 0, 2, 1, 3, 4, 6, 5, 7 };
static const uint8_t TAB16x32[8] = { 0, 1, 2, 3, 4, 5, 6, 7 };
static const uint8_t TAB32x8[16] = { 0, 4, 1, 5, 2, 6, 3, 7, 8, 12, 9, 13, 10, 14, 11, 15 };
static const uint8_t TAB8x32[16] = { 0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15 };
static const uint8_t TAB16x8[32] = { 0, 2, 8, 10, 1, 3, 9, 11, 4, 6, 12, 14, 5, 7, 13, 15,
                                     16, 18, 24, 26, 17, 19, 25, 27, 20, 22, 28, 30, 21, 23, 29, 31 };
static const uint8_t TAB8x16[32] = { 0, 1, 2, 3, 8, 9, 10, 11, 4, 5, 6, 7, 12, 13, 14, 15,
                                     16, 17, 18, 19, 24, 25, 26, 27, 20, 21, 22, 23, 28, 29, 30, 31 };
/* pu_index ranges (EbMotionEstimation.h:176-262, EbMotionEstimationContext.h:45-261): raster inside each class */
typedef struct PuClass { int base, w, h, count; const uint8_t *tab; } PuClass;
static const PuClass PU_CLASS[14] = {
    { 0, 64, 64, 1, TAB_ID },   { 1, 32, 32, 4, TAB_ID },     { 5, 16, 16, 16, TAB16 },     { 21, 8, 8, 64, TAB8 },
    { 85, 64, 32, 2, TAB_ID },  { 87, 32, 16, 8, TAB32x16 },  { 95, 16, 8, 32, TAB16x8 },
    { 127, 32, 64, 2, TAB_ID }, { 129, 16, 32, 8, TAB16x32 }, { 137, 8, 16, 32, TAB8x16 },
    { 169, 32, 8, 16, TAB32x8 }, { 185, 8, 32, 16, TAB8x32 }, { 201, 64, 16, 4, TAB_ID },   { 205, 16, 64, 4, TAB_ID } };

/* open_loop_me_fullpel_search_sblock, ME.c:1556-1595: one position at a time, raster order, every update a plain
 * strict '<' (ExtSadCalculation_8x8_16x16 / _32x32_64x64 SSE4.1 kernels, ASM_SSE4_1/EbComputeSAD_Intrinsic_SSE4_1.c:21-164,
 * and the C ExtSadCalculation, ME.c:266-1053, for the 124 non-square PUs built from the 8x8 / 16x16 / 32x32 SADs). */
static void full_pel_search_ext(SbCtx *c, int list) {
    const MeoPlane *rp = &c->ref[list]->full;
    const int sax = c->sa_x[list], say = c->sa_y[list], saw = c->sa_w[list], sah = c->sa_h[list];
    uint32_t *bs = c->best_sad[list], *bm = c->best_mv[list];
    for (int i = 0; i < MEO_MAX_PU_COUNT; i++) bs[i] = MAX_SAD_VALUE;   /* InitializeBuffer_32bits(.., 52, 1, ..) = 209 words */
#define UPD(base, k, v) do { if ((v) < bs[(base) + (k)]) { bs[(base) + (k)] = (v); bm[(base) + (k)] = mv; } } while (0)
    for (int ys = 0; ys < sah; ys++) {
        const uint8_t *rrow = px(rp, c->ox + sax, c->oy + say + ys);
        for (int xs = 0; xs < saw; xs++) {
            uint32_t s8[64], s16[16], s32[4], s64;
            s64 = position_sads(c, rrow + xs, rp->stride, s8, s16, s32) << 1;
            for (int b = 0; b < 64; b++) s8[b] <<= 1;
            for (int b = 0; b < 16; b++) s16[b] <<= 1;
            for (int b = 0; b < 4; b++) s32[b] <<= 1;
            const uint32_t mv = mkmv((int16_t)((sax + xs) * 4), (int16_t)((say + ys) * 4));
            for (int b = 0; b < 64; b++) UPD(21, b, s8[b]);
            for (int b = 0; b < 16; b++) UPD(5, b, s16[b]);
            for (int b = 0; b < 4; b++) UPD(1, b, s32[b]);
            UPD(0, 0, s64);
            uint32_t s16x8[32], s8x16[32], s32x16[8], s16x32[8], sad;
            /* 64x32 */
            sad = s32[0] + s32[1]; UPD(85, 0, sad);
            sad = s32[2] + s32[3]; UPD(85, 1, sad);
            /* 32x16: entry 5 tests the stale `sad` (quirk Q11) */
            for (int k = 0; k < 8; k++) {
                s32x16[k] = s16[2 * k] + s16[2 * k + 1];
                if (k == 5) { if (sad < bs[87 + 5]) { bs[87 + 5] = s32x16[5]; bm[87 + 5] = mv; } }
                else UPD(87, k, s32x16[k]);
            }
            /* 64x16 */
            UPD(201, 0, s32x16[0] + s32x16[2]); UPD(201, 1, s32x16[1] + s32x16[3]);
            UPD(201, 2, s32x16[4] + s32x16[6]); UPD(201, 3, s32x16[5] + s32x16[7]);
            /* 16x8 */
            for (int k = 0; k < 32; k++) { s16x8[k] = s8[2 * k] + s8[2 * k + 1]; UPD(95, k, s16x8[k]); }
            /* 32x64 */
            UPD(127, 0, s32[0] + s32[2]); UPD(127, 1, s32[1] + s32[3]);
            /* 16x32, 16x64 */
            for (int m = 0; m < 4; m++) for (int i = 0; i < 2; i++) {
                s16x32[2 * m + i] = s16[4 * m + i] + s16[4 * m + i + 2]; UPD(129, 2 * m + i, s16x32[2 * m + i]);
            }
            for (int k = 0; k < 4; k++) UPD(205, k, s16x32[k] + s16x32[k + 4]);
            /* 8x16 */
            for (int m = 0; m < 16; m++) for (int i = 0; i < 2; i++) {
                s8x16[2 * m + i] = s8[4 * m + i] + s8[4 * m + i + 2]; UPD(137, 2 * m + i, s8x16[2 * m + i]);
            }
            /* 32x8 */
            for (int m = 0; m < 8; m++) for (int i = 0; i < 2; i++) UPD(169, 2 * m + i, s16x8[4 * m + i] + s16x8[4 * m + i + 2]);
            /* 8x32 */
            for (int m = 0; m < 4; m++) for (int i = 0; i < 4; i++) UPD(185, 4 * m + i, s8x16[8 * m + i] + s8x16[8 * m + i + 4]);
        }
    }
#undef UPD
}

/* ------------------------------------------------------------------ sub-pel samples */
/* AVC-style 4-tap {-2,18,18,-2}, +16 >> 5, clipped to u8
 * (ASM_SSSE3/EbAvcStyleMcp_Intrinsic_SSSE3.c:200-440, fracPos = 2). All three half-pel planes are
 * pure functions of the padded reference picture at absolute coordinates:
 *   A(x,y)  integer sample
 *   B(x,y)  half-pel between columns x and x+1            (pos_b)
 *   H(x,y)  half-pel between rows y and y+1               (pos_h)
 *   J(x,y)  vertical filter of the clipped B samples       (pos_j, ME.c:1780-1792) */
static inline int clip8(int v) { return v < 0 ? 0 : v > 255 ? 255 : v; }
static inline int sA(const MeoPlane *p, int x, int y) { return *px(p, x, y); }
static inline int sB(const MeoPlane *p, int x, int y) {
    const uint8_t *r = px(p, x, y);
    return clip8((-2 * r[-1] + 18 * r[0] + 18 * r[1] - 2 * r[2] + 16) >> 5);
}
static inline int sH(const MeoPlane *p, int x, int y) {
    const uint8_t *r = px(p, x, y); int s = p->stride;
    return clip8((-2 * r[-s] + 18 * r[0] + 18 * r[s] - 2 * r[2 * s] + 16) >> 5);
}
static inline int sJ(const MeoPlane *p, int x, int y) {
    return clip8((-2 * sB(p, x, y - 1) + 18 * sB(p, x, y) + 18 * sB(p, x, y + 1) - 2 * sB(p, x, y + 2) + 16) >> 5);
}
enum { S_A = 0, S_B = 1, S_H = 2, S_J = 3 };
static inline int samp(const MeoPlane *p, int type, int x, int y) {
    switch (type) { case S_A: return sA(p, x, y); case S_B: return sB(p, x, y); case S_H: return sH(p, x, y); default: return sJ(p, x, y); }
}
/* Materialise a w x h block of one sample type at (x,y) into dst (stride 64). */
static void fetch_block(const MeoPlane *p, int type, int x, int y, int w, int h, uint8_t *dst) {
    for (int r = 0; r < h; r++) for (int q = 0; q < w; q++) dst[r * 64 + q] = (uint8_t)samp(p, type, x + q, y + r);
}

/* SpatialFullDistortionKernel*_SSSE3_INTRIN: |int8(a-b)| squared (quirk Q5) */
static uint32_t wrap_ssd(const uint8_t *s, int ss, const uint8_t *r, int rs, int w, int h) {
    uint32_t acc = 0;
    for (int y = 0; y < h; y++, s += ss, r += rs)
        for (int x = 0; x < w; x++) {
            int d = (int8_t)(uint8_t)(s[x] - r[x]);
            uint32_t e = (uint32_t)(uint8_t)(d < 0 ? -d : d);      /* -128 stays 128 */
            acc += e * e;
        }
    return acc;
}

/* direction codes, EbMotionEstimation.h:62-69 */
enum { D_TL = 0, D_T = 1, D_TR = 2, D_R = 3, D_BR = 4, D_B = 5, D_BL = 6, D_L = 7 };

/* PU_HalfPelRefinement, ME.c:1842-2240 (fractionalSearchMethod = SSD_SEARCH) */
static void half_pel_pu(SbCtx *c, int list, int px0, int py0, int w, int h, int idx) {
    const MeoPlane *rp = &c->ref[list]->full;
    uint32_t *bsad = &c->best_sad[list][idx], *bmv = &c->best_mv[list][idx], *bssd = &c->best_ssd[list][idx];
    const int xm = mvx(*bmv), ym = mvy(*bmv);
    const int X = c->ox + px0 + (xm >> 2), Y = c->oy + py0 + (ym >> 2);
    const uint8_t *src = c->src + (size_t)py0 * c->src_stride + px0;
    uint8_t blk[64 * 64];
    const int hs = (w == 8) ? 8 : h;                                  /* quirk Q10: the 8-wide kernel ranks eight rows */
    *bssd = wrap_ssd(src, c->src_stride, px(rp, X, Y), rp->stride, w, hs);
    /* order L R T B TL TR BR BL */
    static const int8_t ty[8] = { S_B, S_B, S_H, S_H, S_J, S_J, S_J, S_J };
    static const int8_t dx[8] = { -1, 0, 0, 0, -1, 0, 0, -1 };
    static const int8_t dy[8] = { 0, 0, -1, 0, -1, -1, 0, 0 };
    static const int8_t mx[8] = { -2, 2, 0, 0, -2, 2, 2, -2 };
    static const int8_t my[8] = { 0, 0, -2, 2, -2, -2, 2, 2 };
    uint64_t dist[8];
    for (int k = 0; k < 8; k++) {
        fetch_block(rp, ty[k], X + dx[k], Y + dy[k], w, h, blk);
        dist[k] = wrap_ssd(src, c->src_stride, blk, 64, w, hs);
        if (dist[k] < *bssd) {
            *bsad = sad_wxh(src, c->src_stride, blk, 64, w, h);
            *bmv = mkmv((int16_t)(xm + mx[k]), (int16_t)(ym + my[k]));
            *bssd = (uint32_t)dist[k];
        }
    }
    uint64_t best = dist[0];
    for (int k = 1; k < 8; k++) best = MIN(best, dist[k]);
    /* tie order L R T B TL TR BL BR */
    if      (best == dist[0]) c->sub_dir[idx] = D_L;
    else if (best == dist[1]) c->sub_dir[idx] = D_R;
    else if (best == dist[2]) c->sub_dir[idx] = D_T;
    else if (best == dist[3]) c->sub_dir[idx] = D_B;
    else if (best == dist[4]) c->sub_dir[idx] = D_TL;
    else if (best == dist[5]) c->sub_dir[idx] = D_TR;
    else if (best == dist[7]) c->sub_dir[idx] = D_BL;
    else                      c->sub_dir[idx] = D_BR;
}

/* SetQuarterPelRefinementInputsOnTheFly, ME.c:3246-3331, in sample-function form.
 * Entry = {type1, dx1, dy1, type2, dx2, dy2} relative to (X', Y') = integer part of (mv + 2) >> 2. */
static const int8_t QTAB[4][8][6] = {
    { /* EB_QUARTER_IN_FULL */
        { S_B, -1, 0, S_A, 0, 0 }, { S_A, 0, 0, S_B, 0, 0 }, { S_H, 0, -1, S_A, 0, 0 }, { S_A, 0, 0, S_H, 0, 0 },
        { S_B, -1, 0, S_H, 0, -1 }, { S_H, 0, -1, S_B, 0, 0 }, { S_H, 0, 0, S_B, 0, 0 }, { S_B, -1, 0, S_H, 0, 0 } },
    { /* EB_QUARTER_IN_HALF_HORIZONTAL */
        { S_A, -1, 0, S_B, -1, 0 }, { S_B, -1, 0, S_A, 0, 0 }, { S_J, -1, -1, S_B, -1, 0 }, { S_B, -1, 0, S_J, -1, 0 },
        { S_H, -1, -1, S_B, -1, 0 }, { S_B, -1, 0, S_H, 0, -1 }, { S_B, -1, 0, S_H, 0, 0 }, { S_H, -1, 0, S_B, -1, 0 } },
    { /* EB_QUARTER_IN_HALF_VERTICAL */
        { S_J, -1, -1, S_H, 0, -1 }, { S_H, 0, -1, S_J, 0, -1 }, { S_A, 0, -1, S_H, 0, -1 }, { S_H, 0, -1, S_A, 0, 0 },
        { S_B, -1, -1, S_H, 0, -1 }, { S_H, 0, -1, S_B, 0, -1 }, { S_H, 0, -1, S_B, 0, 0 }, { S_B, -1, 0, S_H, 0, -1 } },
    { /* EB_QUARTER_IN_HALF_DIAGONAL */
        { S_H, -1, -1, S_J, -1, -1 }, { S_J, -1, -1, S_H, 0, -1 }, { S_B, -1, -1, S_J, -1, -1 }, { S_J, -1, -1, S_B, -1, 0 },
        { S_H, -1, -1, S_B, -1, -1 }, { S_B, -1, -1, S_H, 0, -1 }, { S_B, -1, 0, S_H, 0, -1 }, { S_H, -1, -1, S_B, -1, 0 } },
};

/* PU_QuarterPelRefinementOnTheFly, ME.c:2824-3239 */
static void quarter_pel_pu(SbCtx *c, int list, int px0, int py0, int w, int h, int idx) {
    const MeoPlane *rp = &c->ref[list]->full;
    uint32_t *bsad = &c->best_sad[list][idx], *bmv = &c->best_mv[list][idx], *bssd = &c->best_ssd[list][idx];
    const int xm = mvx(*bmv), ym = mvy(*bmv), dir = c->sub_dir[idx];
    const int X = c->ox + px0 + ((xm + 2) >> 2), Y = c->oy + py0 + ((ym + 2) >> 2);
    const int method = (ym & 2) + ((xm & 2) >> 1);
    const uint8_t *src = c->src + (size_t)py0 * c->src_stride + px0;     /* sb_buffer == same samples */
    int valid[8];  /* order L R T B TL TR BR BL */
#define IN3(a, b, d) (dir == (a) || dir == (b) || dir == (d))
    if (method) {
        valid[4] = IN3(D_R, D_BR, D_B);   valid[2] = IN3(D_BR, D_B, D_BL);  valid[5] = IN3(D_B, D_BL, D_L);
        valid[1] = IN3(D_BL, D_L, D_TL);  valid[6] = IN3(D_L, D_TL, D_T);   valid[3] = IN3(D_TL, D_T, D_TR);
        valid[7] = IN3(D_T, D_TR, D_R);   valid[0] = IN3(D_TR, D_R, D_BR);
    } else {
        valid[4] = IN3(D_L, D_TL, D_T);   valid[2] = IN3(D_TL, D_T, D_TR);  valid[5] = IN3(D_T, D_TR, D_R);
        valid[1] = IN3(D_TR, D_R, D_BR);  valid[6] = IN3(D_R, D_BR, D_B);   valid[3] = IN3(D_BR, D_B, D_BL);
        valid[7] = IN3(D_B, D_BL, D_L);   valid[0] = IN3(D_BL, D_L, D_TL);
    }
#undef IN3
    static const int8_t mx[8] = { -1, 1, 0, 0, -1, 1, 1, -1 };
    static const int8_t my[8] = { 0, 0, -1, 1, -1, -1, 1, 1 };
    uint8_t b1[64 * 64], b2[64 * 64];
    for (int k = 0; k < 8; k++) {
        if (!valid[k]) continue;
        const int8_t *q = QTAB[method][k];
        fetch_block(rp, q[0], X + q[1], Y + q[2], w, h, b1);
        fetch_block(rp, q[3], X + q[4], Y + q[5], w, h, b2);
        uint32_t ssd = 0, sad = 0;                     /* CombinedAveragingSSD (true SSD) / CombinedAveraging*SAD */
        for (int r = 0; r < h; r++) for (int t = 0; t < w; t++) {
            int a = (b1[r * 64 + t] + b2[r * 64 + t] + 1) >> 1, d = (int)src[(size_t)r * c->src_stride + t] - a;
            ssd += (uint32_t)(d * d); sad += (uint32_t)abs(d);
        }
        if (ssd < *bssd) { *bsad = sad; *bmv = mkmv((int16_t)(xm + mx[k]), (int16_t)(ym + my[k])); *bssd = ssd; }
    }
}

/* HalfPelSearch_LCU ME.c:2246-2415 + QuarterPelSearch_LCU ME.c:3337-3585, square PUs */
static void subpel_search(SbCtx *c, int list) {
    half_pel_pu(c, list, 0, 0, 64, 64, 0);
    for (int i = 0; i < 4; i++) half_pel_pu(c, list, (i & 1) << 5, (i >> 1) << 5, 32, 32, 1 + i);
    for (int i = 0; i < 16; i++) half_pel_pu(c, list, (i & 3) << 4, (i >> 2) << 4, 16, 16, 5 + TAB16[i]);
    if (c->prm->cu8x8_mode != 1)
        for (int i = 0; i < 64; i++) half_pel_pu(c, list, (i & 7) << 3, (i >> 3) << 3, 8, 8, 21 + TAB8[i]);
    quarter_pel_pu(c, list, 0, 0, 32, 32, 0);                                       /* quirk Q6 */
    for (int i = 0; i < 4; i++) quarter_pel_pu(c, list, (i & 1) << 5, (i >> 1) << 5, 32, 32, 1 + i);
    for (int i = 0; i < 16; i++) quarter_pel_pu(c, list, (i & 3) << 4, (i >> 2) << 4, 16, 16, 5 + TAB16[i]);
    if (c->prm->cu8x8_mode != 1)
        for (int i = 0; i < 64; i++) quarter_pel_pu(c, list, (i & 7) << 3, (i >> 3) << 3, 8, 8, 21 + TAB8[i]);
    if (c->prm->num_pus != MEO_MAX_PU_COUNT) return;
    /* the ten non-square classes, ME.c:2417-2790 and 3582-4110 */
    for (int pass = 0; pass < 2; pass++)
        for (int k = 4; k < 14; k++) {
            const PuClass *pc = &PU_CLASS[k];
            const int ncols = 64 / pc->w;
            for (int i = 0; i < pc->count; i++) {
                const int x = (i % ncols) * pc->w, y = (i / ncols) * pc->h, idx = pc->base + pc->tab[i];
                if (pass == 0) half_pel_pu(c, list, x, y, pc->w, pc->h, idx);
                else           quarter_pel_pu(c, list, x, y, pc->w, pc->h, idx);
            }
        }
}

/* ------------------------------------------------------------------ bi-prediction + packing */
/* Prediction block of one list at quarter-pel MV: SelectBuffer / QuarterPelCompensation,
 * ME.c:4762-4920 (frac = fx + 4*fy). */
static void uni_pred_block(const MeoPlane *rp, int X, int Y, int frac, int w, int h, uint8_t *dst) {
    /* {type1,dx1,dy1,type2,dx2,dy2}; type2 < 0 => single buffer */
    static const int8_t T[16][6] = {
        { S_A, 0, 0, -1, 0, 0 },  { S_A, 0, 0, S_B, 0, 0 }, { S_B, 0, 0, -1, 0, 0 },  { S_B, 0, 0, S_A, 1, 0 },
        { S_A, 0, 0, S_H, 0, 0 }, { S_B, 0, 0, S_H, 0, 0 }, { S_B, 0, 0, S_J, 0, 0 }, { S_B, 0, 0, S_H, 1, 0 },
        { S_H, 0, 0, -1, 0, 0 },  { S_H, 0, 0, S_J, 0, 0 }, { S_J, 0, 0, -1, 0, 0 },  { S_J, 0, 0, S_H, 1, 0 },
        { S_H, 0, 0, S_A, 0, 1 }, { S_H, 0, 0, S_B, 0, 1 }, { S_J, 0, 0, S_B, 0, 1 }, { S_H, 1, 0, S_B, 0, 1 } };
    const int8_t *t = T[frac];
    fetch_block(rp, t[0], X + t[1], Y + t[2], w, h, dst);
    if (t[3] >= 0) {
        uint8_t b2[64 * 64];
        fetch_block(rp, t[3], X + t[4], Y + t[5], w, h, b2);
        for (int r = 0; r < h; r++) for (int q = 0; q < w; q++)
            dst[r * 64 + q] = (uint8_t)((dst[r * 64 + q] + b2[r * 64 + q] + 1) >> 1);
    }
}

/* BiPredictionCompensation + BiPredAverging, ME.c:4933-5254 (full SAD of src vs avg(L0,L1)) */
static uint32_t bipred_sad(const SbCtx *c, int px0, int py0, int w, int h, uint32_t mv0, uint32_t mv1) {
    uint8_t p0[64 * 64], p1[64 * 64];
    const uint32_t mv[2] = { mv0, mv1 }; uint8_t *pb[2] = { p0, p1 };
    for (int l = 0; l < 2; l++) {
        int x = mvx(mv[l]), y = mvy(mv[l]);
        uni_pred_block(&c->ref[l]->full, c->ox + px0 + (x >> 2), c->oy + py0 + (y >> 2),
                       (x & 3) + ((y & 3) << 2), w, h, pb[l]);
    }
    const uint8_t *src = c->src + (size_t)py0 * c->src_stride + px0;
    uint32_t sad = 0;
    for (int r = 0; r < h; r++) for (int q = 0; q < w; q++)
        sad += (uint32_t)abs((int)src[(size_t)r * c->src_stride + q] - ((p0[r * 64 + q] + p1[r * 64 + q] + 1) >> 1));
    return sad;
}

/* pu_index (raster inside each size class) -> geometry and internal index */
static void pu_geom(int pu, int *x, int *y, int *w, int *h, int *nidx) {
    int k = 13;
    while (pu < PU_CLASS[k].base) k--;
    const PuClass *pc = &PU_CLASS[k];
    const int i = pu - pc->base, ncols = 64 / pc->w;
    *x = (i % ncols) * pc->w; *y = (i / ncols) * pc->h; *w = pc->w; *h = pc->h; *nidx = pc->base + pc->tab[i];
}

/* The tail of MotionEstimateLcu, ME.c:6973-7146 */
static void pack_results(SbCtx *c, MeoPuResult *out) {
    const MeoParams *p = c->prm;
    const int ext = p->num_pus == MEO_MAX_PU_COUNT;
    for (int pu = 0; pu < p->num_pus; pu++) {
        int x, y, w, h, n; pu_geom(pu, &x, &y, &w, &h, &n);
        MeoPuResult *o = &out[pu];
        memset(o, 0, sizeof(*o));
        int total = p->num_lists;
        if (p->num_lists == 2 && (p->cu8x8_mode == 0 || pu < 21 || ext)) {
            c->bipred_sad[n] = bipred_sad(c, x, y, w, h, c->best_mv[0][n], c->best_mv[1][n]);
            total = 3;
        }
        o->total_cand = (uint8_t)total;
        o->xMvL0 = mvx(c->best_mv[0][n]); o->yMvL0 = mvy(c->best_mv[0][n]);
        if (p->num_lists == 2) { o->xMvL1 = mvx(c->best_mv[1][n]); o->yMvL1 = mvy(c->best_mv[1][n]); }
        const uint32_t a = c->best_sad[0][n], b = c->best_sad[1][n], cc = c->bipred_sad[n];
#define SET(k, d, dirv) do { o->cand[k].distortion = (d); o->cand[k].direction = (dirv); } while (0)
        if (total == 3) {                                         /* Sort3Elements, ME.c:5434-5463 */
            if (a <= b && a <= cc) {
                if (b <= cc) { SET(0, a, 0); SET(1, b, 1); SET(2, cc, 2); } else { SET(0, a, 0); SET(1, cc, 2); SET(2, b, 1); }
            } else if (b <= a && b <= cc) {
                if (a <= cc) { SET(0, b, 1); SET(1, a, 0); SET(2, cc, 2); } else { SET(0, b, 1); SET(1, cc, 2); SET(2, a, 0); }
            } else if (a <= b) { SET(0, cc, 2); SET(1, a, 0); SET(2, b, 1); }
            else               { SET(0, cc, 2); SET(1, b, 1); SET(2, a, 0); }
        } else if (total == 2) {
            if (a <= b) { SET(0, a, 0); SET(1, b, 1); } else { SET(0, b, 1); SET(1, a, 0); }
        } else SET(0, a, 0);
#undef SET
    }
}

/* ------------------------------------------------------------------ picture entry */
int meo_motion_estimate_picture(const MeoParams *prm, const MeoPyramid *cur,
                                const MeoPyramid *ref0, const MeoPyramid *ref1,
                                MeoPuResult *results, MeoSbDebug *debug) {
    if (prm->num_pus != MEO_SQUARE_PU_COUNT && prm->num_pus != MEO_MAX_PU_COUNT) return -2;
    if ((prm->width & 7) || (prm->height & 7) || (prm->width & 63) == 40 || (prm->width & 63) == 56) return -3;
    SbCtx *c = (SbCtx *)calloc(1, sizeof(SbCtx));
    if (!c) return -1;
    c->prm = prm; c->cur = cur; c->ref[0] = ref0; c->ref[1] = ref1 ? ref1 : ref0;
    const int w_sb = (prm->width + 63) / 64, h_sb = (prm->height + 63) / 64;
    for (int ys = 0; ys < h_sb; ys++) for (int xs = 0; xs < w_sb; xs++) {
        const int sb = xs + ys * w_sb;
        c->ox = xs * 64; c->oy = ys * 64;
        c->sb_w = MIN(64, prm->width - c->ox); c->sb_h = MIN(64, prm->height - c->oy);
        c->src = px(&cur->full, c->ox, c->oy); c->src_stride = cur->full.stride;
        /* staging, EbMotionEstimationProcess.c:519-544 */
        if (prm->enable_hme_level1)
            for (int r = 0; r < (c->sb_h >> 1); r++)
                memcpy(&c->quarter_sb[r * 32], px(&cur->quarter, c->ox >> 1, (c->oy >> 1) + r), c->sb_w >> 1);
        if (prm->enable_hme_level0)
            for (int r = 0, k = 0; r < (c->sb_h >> 2); r += 2, k++)
                memcpy(&c->sixteenth_sb[k * 16], px(&cur->sixteenth, c->ox >> 2, (c->oy >> 2) + r), c->sb_w >> 2);

        for (int list = 0; list < prm->num_lists; list++) {
            int xsc, ysc;
            find_search_center(c, list, &xsc, &ysc);
            int saw = MIN(prm->search_area_w, 127), sah = MIN(prm->search_area_h, 127);
            if ((xsc != 0 || ysc != 0) && prm->is_used_as_reference) {       /* CheckZeroZeroCenter, ME.c:5466-5552 */
                const MeoPlane *rp = &c->ref[list]->full;
                uint64_t zero = sb_sub_sad(c, rp, 0, 0);
                xsc = clip_center(c->ox, xsc, 63, rp->width);
                ysc = clip_center(c->oy, ysc, 63, rp->height);
                uint64_t hme = sb_sub_sad(c, rp, xsc, ysc);
                if (MIN(zero, hme) == zero) { xsc = 0; ysc = 0; }
            }
            int sax = xsc - (saw >> 1), say = ysc - (sah >> 1);
            clip_search_area(c->ox, &sax, &saw, 63, prm->width);
            clip_search_area(c->oy, &say, &sah, 63, prm->height);
            c->sa_x[list] = sax; c->sa_y[list] = say; c->sa_w[list] = saw; c->sa_h[list] = sah;
            if (prm->num_pus == MEO_MAX_PU_COUNT) full_pel_search_ext(c, list); else full_pel_search(c, list);
            if (prm->use_subpel == 1) subpel_search(c, list);
            if (debug) {
                MeoSbDebug *d = &debug[sb];
                if (list == 0) memset(d, 0, sizeof(*d));
                d->search_center_x[list] = (int16_t)xsc; d->search_center_y[list] = (int16_t)ysc;
                d->sa_origin_x[list] = (int16_t)sax; d->sa_origin_y[list] = (int16_t)say;
                d->sa_width[list] = (int16_t)saw; d->sa_height[list] = (int16_t)sah;
                memcpy(d->best_sad[list], c->best_sad[list], sizeof(d->best_sad[list]));   /* square PUs only */
                memcpy(d->best_mv[list], c->best_mv[list], sizeof(d->best_mv[list]));
            }
        }
        pack_results(c, &results[(size_t)sb * prm->num_pus]);
    }
    free(c);
    return 0;
}

/* UnPack2D (EbPictureOperators.c:642-662) -> EB_ENC_msbUnPack2D (C_DEFAULT/EbPackUnPack_C.c:127-151): the split of a
 * 16-bit-per-sample 10-bit input plane done in CopyInputBuffer (EbEncHandle.c:2954-2963). ME reads the msb plane. */
void meo_unpack16(const uint16_t *in, int32_t in_stride, uint8_t *msb, int32_t msb_stride, uint8_t *rem, int32_t rem_stride,
                  int32_t width, int32_t height) {
    for (int y = 0; y < height; y++)
        for (int x = 0; x < width; x++) {
            const uint16_t v = in[(size_t)y * in_stride + x];
            msb[(size_t)y * msb_stride + x] = (uint8_t)(v >> 2);
            /* quirk Q9 (ASM_SSE2/EbPackUnPack_Intrinsic_SSE2.c:139,180): the width == 32 and width == 64 special cases
             * forget the << 6 on their last 16 samples (second row of each row pair for 32, every row for 64) */
            const int q9 = (width == 64 && x >= 48) || (width == 32 && (y & 1) && x >= 16);
            if (rem) rem[(size_t)y * rem_stride + x] = q9 ? (uint8_t)(v & 3) : (uint8_t)(v << 6);
        }
}


/* ---- ComputeDecimatedZzSad (EbMotionEstimationProcess.c:219-349) ---------------------------------------------------
 * For every COMPLETE 64x64 SB: the previous picture's SB is decimated on the fly (Decimation2D with step 4: sample
 * (4x, 4y), EbPictureAnalysisProcess.c:100-125) and compared with the same SB of the current sixteenth picture by the plain
 * 16x16 SAD (NxMSadKernel_funcPtrArray[asm][2] = Compute16xMSad_AVX2_INTRIN: every row, no quirk). Thresholds:
 * BEA_CLASS_*_DEC_TH = 256, 512, 1024, 2048 -> zz cost 0, 3, 10, 20, 30 (EbDefinitions.h:2770-2790); the non-moving index
 * uses (sb_w / 4) * (sb_h / 4) * {2, 4, 8} -> 0, 10, 20, 30. Incomplete SBs: cost 255, SAD = ~0 -> index 30. */
void meo_zz_sad_picture(int32_t width, int32_t height, const MeoPyramid *cur, const MeoPyramid *prev,
                        uint8_t *zz_cost, uint8_t *non_moving_index) {
    const int w_sb = (width + 63) / 64, h_sb = (height + 63) / 64;
    for (int ys = 0; ys < h_sb; ys++)
        for (int xs = 0; xs < w_sb; xs++) {
            const int sb = xs + ys * w_sb, ox = xs * 64, oy = ys * 64;
            const int sb_w = width - ox < 64 ? width - ox : 64, sb_h = height - oy < 64 ? height - oy : 64;
            uint32_t sad = 0;
            if (sb_w == 64 && sb_h == 64) {
                const MeoPlane *c = &cur->sixteenth, *p = &prev->full;
                for (int y = 0; y < 16; y++)
                    for (int x = 0; x < 16; x++) {
                        const int a = c->buf[(size_t)(c->origin_y + (oy >> 2) + y) * c->stride + c->origin_x + (ox >> 2) + x];
                        const int b = p->buf[(size_t)(p->origin_y + oy + 4 * y) * p->stride + p->origin_x + ox + 4 * x];
                        sad += (uint32_t)(a > b ? a - b : b - a);
                    }
                zz_cost[sb] = sad < 256 ? 0 : sad < 512 ? 3 : sad < 1024 ? 10 : sad < 2048 ? 20 : 30;
            } else {
                zz_cost[sb] = 255;
                sad = ~0u;
            }
            const uint32_t area = (uint32_t)((sb_w >> 2) * (sb_h >> 2));
            non_moving_index[sb] = sad < area * 2 ? 0 : sad < area * 4 ? 10 : sad < area * 8 ? 20 : 30;
        }
}
