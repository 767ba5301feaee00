/* oracle/me_oracle.h -- TEST INFRASTRUCTURE ONLY (checker, never shipped, never measured
 * except as bench.py's cpu_baseline / --impl reference leg).
 *
 * Flat, picture-level description of the SVT-AV1 open-loop motion-estimation hot path
 * (MotionEstimationKernel -> MotionEstimateLcu, reference
 * Source/Lib/Codec/EbMotionEstimationProcess.c:478-556 and
 * Source/Lib/Codec/EbMotionEstimation.c:6152-7162) shared by
 *   - oracle/me_oracle.c   : our plain-C restatement        (symbols meo_*)
 *   - oracle/ref_driver.c  : driver around the real reference objects (symbols ref_*)
 * Both expose the same entry points so tests can diff them field by field.
 */
#ifndef ME_ORACLE_H
#define ME_ORACLE_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MEO_SQUARE_PU_COUNT 85   /* EbMotionEstimationLcuResults.h:21 */
#define MEO_MAX_PU_COUNT    209  /* EbMotionEstimationLcuResults.h:17 */
#define MEO_PAD_FULL        68   /* BLOCK_SIZE_64 + ME_FILTER_TAP, EbEncHandle.c:1006 */
#define MEO_PAD_QUARTER     32   /* sb_sz >> 1, EbEncHandle.c:1015 */
#define MEO_PAD_SIXTEENTH   16   /* sb_sz >> 2, EbEncHandle.c:1025 */

/* One padded 8-bit luma plane. buf points at the first byte of the allocation
 * (top-left of the padding); pixel (x,y) of the picture is
 * buf[(origin_y + y) * stride + origin_x + x]. */
typedef struct MeoPlane {
    uint8_t *buf;
    int32_t  stride;
    int32_t  origin_x, origin_y;
    int32_t  width, height;
} MeoPlane;

/* Full / quarter (1/2 x 1/2) / sixteenth (1/4 x 1/4) planes of one source picture
 * (EbPaReferenceObject_t, EbReferenceObject.c:219-258). */
typedef struct MeoPyramid {
    MeoPlane full, quarter, sixteenth;
} MeoPyramid;

/* Everything MotionEstimateLcu reads from PCS / SCS / MeContext. */
typedef struct MeoParams {
    int32_t width, height;            /* scs->luma_width / luma_height (multiples of 8) */
    int32_t num_lists;                /* 1 = P (list 0 only), 2 = B     ME.c:6271 */
    int32_t temporal_layer_index;     /* pcs->temporal_layer_index */
    int32_t hierarchical_levels;      /* pcs->hierarchical_levels */
    int32_t is_used_as_reference;     /* pcs->is_used_as_reference_flag */
    int32_t same_ref_poc;             /* ref_pic_poc_array[0] == ref_pic_poc_array[1]  ME.c:6606 */
    int32_t enable_hme_flag, enable_hme_level0, enable_hme_level1, enable_hme_level2;
    int32_t use_subpel;               /* pcs->use_subpel_flag */
    int32_t num_pus;                  /* 85 (PIC_SQ_*) or 209 (PIC_ALL_*)  EbPictureDecisionProcess.c:378 */
    int32_t cu8x8_mode;               /* CU_8x8_MODE_0 = 0 */
    /* search-area parameters (set_me_hme_params_oq, EbMotionEstimationProcess.c:94-156) */
    int32_t num_hme_regions_w, num_hme_regions_h;
    int32_t hme_l0_total_w, hme_l0_total_h;
    int32_t hme_l0_w[2], hme_l0_h[2];
    int32_t hme_l1_w[2], hme_l1_h[2];
    int32_t hme_l2_w[2], hme_l2_h[2];
    int32_t search_area_w, search_area_h;
    int32_t hme_l0_mult_x, hme_l0_mult_y;  /* HME_LEVEL_0_SEARCH_AREA_MULTIPLIER_X/Y  EbDefinitions.h:2980-2996 */
} MeoParams;

/* Binary image of MeCuResults_t (EbMotionEstimationLcuResults.h:62-75): 40 bytes. */
typedef struct MeoPuResult {
    int16_t  xMvL0, yMvL0, xMvL1, yMvL1;
    struct { uint32_t distortion; uint32_t direction; } cand[3];
    uint8_t  total_cand;
    uint8_t  pad[7];
} MeoPuResult;

/* Per-SB, per-list intermediate state, for stage-level parity. Indexing of sad/mv is the
 * reference's internal p_sb_best_* order (64x64, 32x32 x4, 16x16 x16 z-order, 8x8 x64 z-order). */
typedef struct MeoSbDebug {
    int16_t  search_center_x[2], search_center_y[2];   /* after HME + zero-centre check */
    int16_t  sa_origin_x[2], sa_origin_y[2];           /* clipped search-area origin */
    int16_t  sa_width[2], sa_height[2];                /* clipped search-area size   */
    uint32_t best_sad[2][MEO_SQUARE_PU_COUNT];
    uint32_t best_mv[2][MEO_SQUARE_PU_COUNT];
} MeoSbDebug;

/* ---- plane / pyramid helpers ------------------------------------------------------- */
/* Bytes needed for a padded plane of (width x height) with 'pad' on every side. */
int64_t meo_plane_bytes(int32_t width, int32_t height, int32_t pad);
/* Describe a plane living at 'mem'. */
void    meo_plane_init(MeoPlane *p, uint8_t *mem, int32_t width, int32_t height, int32_t pad);

/* Default ME/HME parameters for a resolution: M0 column of the tables
 * (hmeMeLevel = 0 for every legal preset, EbMotionEstimationProcess.c:102). */
void meo_default_params(MeoParams *p, int32_t width, int32_t height, int32_t num_lists,
                        int32_t temporal_layer_index, int32_t hierarchical_levels,
                        int32_t is_used_as_reference);

/* ---- the path: two implementations, same signatures ------------------------------- */
/* Build the padded pyramid of one picture from its unpadded luma (src_stride bytes/row):
 * copy + generate_padding + Decimation2D(2) + pad + Decimation2D(4) + pad
 * (EbPictureAnalysisProcess.c:4866-4937, EbMcp.c:173-216). Planes must be pre-described. */
void meo_build_pyramid(const uint8_t *luma, int32_t src_stride, MeoPyramid *pyr);
/* 10-bit input split: msb = sample >> 2, rem = (uint8_t)(sample << 6); strides in samples / bytes */
void meo_unpack16(const uint16_t *in, int32_t in_stride, uint8_t *msb, int32_t msb_stride, uint8_t *rem, int32_t rem_stride,
                  int32_t width, int32_t height);
void ref_build_pyramid(const uint8_t *luma, int32_t src_stride, MeoPyramid *pyr);

/* Motion-estimate every SB of one picture. results: [num_sb][params->num_pus];
 * debug (may be NULL): [num_sb]. Returns 0 on success. */
int meo_motion_estimate_picture(const MeoParams *params, const MeoPyramid *cur,
                                const MeoPyramid *ref0, const MeoPyramid *ref1,
                                MeoPuResult *results, MeoSbDebug *debug);
int ref_motion_estimate_picture(const MeoParams *params, const MeoPyramid *cur,
                                const MeoPyramid *ref0, const MeoPyramid *ref1,
                                MeoPuResult *results, MeoSbDebug *debug);

/* Zero-motion SADs of one picture against the previous picture in display order (ComputeDecimatedZzSad,
 * EbMotionEstimationProcess.c:219-349): per SB the 16x16 SAD between the current sixteenth picture and the previous FULL
 * picture sampled every fourth pixel, classified into zz_cost[num_sb] and non_moving_index[num_sb] (both belong to the
 * PREVIOUS picture's control set in the reference). Incomplete SBs get INVALID_ZZ_COST / class 3. */
void meo_zz_sad_picture(int32_t width, int32_t height, const MeoPyramid *cur, const MeoPyramid *prev,
                        uint8_t *zz_cost, uint8_t *non_moving_index);
int  ref_zz_sad_picture(int32_t width, int32_t height, const MeoPyramid *cur, const MeoPyramid *prev,
                        uint8_t *zz_cost, uint8_t *non_moving_index);

/* ---- open-loop intra search of the ME process (ois_oracle.c; checker of svtme_ois) -------------
 * OpenLoopIntraSearchLcu (EbMotionEstimation.c:8047-8365) for every SB. Per CU (raster-scan CU index 1..84 = the ME
 * pu_index of the same square) the candidate array sorted_ois_candidate[cu][] and total_intra_luma_mode[cu]; both start
 * from zero here and in the reference driver, so whole arrays compare. */
typedef struct MeoOisCand { uint32_t distortion; uint8_t mode, valid, pad[2]; } MeoOisCand;
typedef struct MeoOisCu { MeoOisCand cand[18]; uint8_t total, pad[3]; } MeoOisCu;
typedef struct MeoOisParams {
    int32_t width, height;
    int32_t slice_type;            /* 0 = I, 1 = P, 2 = B */
    int32_t temporal_layer_index, is_used_as_reference;
    int32_t cu8x8_mode;            /* 1 = no 8x8 CUs */
    int32_t input_resolution;      /* 0..3, 3 = INPUT_SIZE_4K_RANGE */
    int32_t limit_ois_to_dc;       /* enc-mode >= 6 only; kept for completeness */
    int32_t enc_mode;              /* read in ONE place: the valid flag of a stage-1 winner 'vertical' (quirk, see ois_oracle.c) */
} MeoOisParams;
void meo_ois_picture(const MeoOisParams *prm, const MeoPlane *pic, const uint32_t *me_sad, MeoOisCu *out);
int  ref_ois_picture(const MeoOisParams *prm, const MeoPlane *pic, const uint32_t *me_sad, MeoOisCu *out);
/* one prediction (n x n, mode 0..34) of the CU at (x0, y0) into pred64 (stride 64) + its SAD against the source */
uint32_t meo_ois_predict_cu(const MeoPlane *pic, int32_t x0, int32_t y0, int32_t n, int32_t mode, uint8_t *pred64);
uint32_t ref_ois_predict_cu(const MeoPlane *pic, int32_t x0, int32_t y0, int32_t n, int32_t mode, uint8_t *pred64);

#ifdef __cplusplus
}
#endif
#endif /* ME_ORACLE_H */
