/* oracle/ref_driver.c -- TEST INFRASTRUCTURE ONLY.
 *
 * Picture-level driver around the UNMODIFIED reference sources, compiled where they lie
 * under /root/reference (see oracle/Makefile). It fills the minimum of
 * PictureParentControlSet_t / SequenceControlSet_t / MeContext_t that
 * MotionEstimateLcu (EbMotionEstimation.c:6152) reads, stages each SB exactly like
 * MotionEstimationKernel does (EbMotionEstimationProcess.c:481-555) and converts
 * me_results[][] into the flat MeoPuResult image declared in me_oracle.h.
 *
 * EbMotionEstimation.c is #included (unity build) so its static functions are reachable;
 * no reference source is copied into this repository.
 */
/* The reference's MotionEstimateLcu is compiled under another name so that the wrapper at the end of this file can
 * record what the REAL encoder (oracle/enc_driver.c) feeds it and what it returns; with no dump open it is a plain call. */
#define MotionEstimateLcu MotionEstimateLcu_reference
#include "EbMotionEstimation.c"          /* resolved through -I$(REF)/Lib/Codec */
#undef MotionEstimateLcu
EbErrorType MotionEstimateLcu(PictureParentControlSet_t *, uint32_t, uint32_t, uint32_t, MeContext_t *, EbPictureBufferDesc_t *);
#include "EbMotionEstimationProcess.h"
#include "EbPictureAnalysisProcess.h"
#include "EbMcp.h"
#include "me_oracle.h"
#include <stdlib.h>
#include <string.h>

extern EbMemoryMapEntry *memoryMap;
extern uint32_t         *memoryMapIndex;
extern uint64_t         *totalLibMemory;

extern void *set_me_hme_params_oq(MeContext_t *, PictureParentControlSet_t *, SequenceControlSet_t *, EbInputResolution);
extern void Decimation2D(uint8_t *, uint32_t, uint32_t, uint32_t, uint8_t *, uint32_t, uint32_t);

static void ensure_memory_map(void) {
    static EbMemoryMapEntry *map; static uint32_t idx; static uint64_t total;
    if (!map) map = (EbMemoryMapEntry *)calloc(4096, sizeof(EbMemoryMapEntry));
    idx = 0;                    /* we never free through the map; just keep it from overflowing */
    memoryMap = map; memoryMapIndex = &idx; totalLibMemory = &total;
}

static void plane_to_desc(const MeoPlane *p, EbPictureBufferDesc_t *d) {
    memset(d, 0, sizeof(*d));
    d->bufferY = p->buf; d->strideY = (uint16_t)p->stride;
    d->origin_x = (uint16_t)p->origin_x; d->origin_y = (uint16_t)p->origin_y;
    d->width = (uint16_t)p->width; d->height = (uint16_t)p->height;
    d->maxWidth = d->width; d->maxHeight = d->height;
}

typedef struct RefPa {
    EbObjectWrapper_t wrapper;
    EbPaReferenceObject_t *obj;
    EbPictureBufferDesc_t full, quarter, sixteenth;
} RefPa;

static void pa_init(RefPa *pa, const MeoPyramid *pyr) {
    pa->obj = (EbPaReferenceObject_t *)calloc(1, sizeof(EbPaReferenceObject_t));
    plane_to_desc(&pyr->full, &pa->full);
    plane_to_desc(&pyr->quarter, &pa->quarter);
    plane_to_desc(&pyr->sixteenth, &pa->sixteenth);
    pa->obj->inputPaddedPicturePtr = &pa->full;
    pa->obj->quarterDecimatedPicturePtr = &pa->quarter;
    pa->obj->sixteenthDecimatedPicturePtr = &pa->sixteenth;
    memset(&pa->wrapper, 0, sizeof(pa->wrapper));
    pa->wrapper.objectPtr = pa->obj;
}

/* EbPictureAnalysisProcess.c:4866-4937 driven with the reference's own leaf functions. */
void ref_build_pyramid(const uint8_t *luma, int32_t src_stride, MeoPyramid *pyr) {
    MeoPlane *f = &pyr->full, *q = &pyr->quarter, *s = &pyr->sixteenth;
    for (int y = 0; y < f->height; y++)
        memcpy(f->buf + (size_t)(f->origin_y + y) * f->stride + f->origin_x, luma + (size_t)y * src_stride, f->width);
    generate_padding(f->buf, f->stride, f->width, f->height, f->origin_x, f->origin_y);
    Decimation2D(f->buf + f->origin_x + f->origin_y * f->stride, f->stride, f->width, f->height,
                 q->buf + q->origin_x + q->origin_x * q->stride, q->stride, 2);
    generate_padding(q->buf, q->stride, q->width, q->height, q->origin_x, q->origin_y);
    Decimation2D(f->buf + f->origin_x + f->origin_y * f->stride, f->stride, f->width, f->height,
                 s->buf + s->origin_x + s->origin_x * s->stride, s->stride, 4);
    generate_padding(s->buf, s->stride, s->width, s->height, s->origin_x, s->origin_y);
}

/* the reference's own 10-bit input split, AVX2 dispatch row (UnPack2D picks the SSE2 kernel for width % 4 == 0, even height) */
void ref_unpack16(uint16_t *in, int32_t in_stride, uint8_t *msb, int32_t msb_stride, uint8_t *rem, int32_t rem_stride,
                  int32_t width, int32_t height) {
    UnPack2D(in, (uint32_t)in_stride, msb, (uint32_t)msb_stride, rem, (uint32_t)rem_stride, (uint32_t)width, (uint32_t)height, ASM_AVX2);
}

int ref_motion_estimate_picture(const MeoParams *prm, const MeoPyramid *cur,
                                const MeoPyramid *ref0, const MeoPyramid *ref1,
                                MeoPuResult *results, MeoSbDebug *debug) {
    ensure_memory_map();

    SequenceControlSet_t *scs = (SequenceControlSet_t *)calloc(1, sizeof(SequenceControlSet_t));
    EncodeContext_t *enc = (EncodeContext_t *)calloc(1, sizeof(EncodeContext_t));
    PictureParentControlSet_t *pcs = (PictureParentControlSet_t *)calloc(1, sizeof(PictureParentControlSet_t));
    EbObjectWrapper_t scs_wrapper; memset(&scs_wrapper, 0, sizeof(scs_wrapper));
    enc->asm_type = ASM_AVX2;                     /* the only runnable path, EbEncHandle.c:2535-2539 */
    scs->encode_context_ptr = enc;
    scs->luma_width = (uint16_t)prm->width; scs->luma_height = (uint16_t)prm->height;
    scs->sb_sz = 64;
    scs->static_config.rate_control_mode = 0;
    scs->static_config.use_default_me_hme = 1;
    derive_input_resolution(scs, (uint32_t)prm->width * (uint32_t)prm->height);
    scs_wrapper.objectPtr = scs;

    RefPa pa_cur, pa_ref[2];
    pa_init(&pa_cur, cur);
    pa_init(&pa_ref[0], ref0);
    pa_init(&pa_ref[1], ref1 ? ref1 : ref0);

    const uint32_t w_sb = (prm->width + 63) / 64, h_sb = (prm->height + 63) / 64, n_sb = w_sb * h_sb;
    pcs->sequence_control_set_wrapper_ptr = &scs_wrapper;
    pcs->max_number_of_pus_per_sb = (uint8_t)prm->num_pus;
    pcs->pic_depth_mode = (prm->num_pus == MEO_MAX_PU_COUNT) ? PIC_ALL_DEPTH_MODE : PIC_SQ_DEPTH_MODE;
    pcs->enable_hme_flag = (EbBool)prm->enable_hme_flag;
    pcs->enable_hme_level0_flag = (EbBool)prm->enable_hme_level0;
    pcs->enable_hme_level1_flag = (EbBool)prm->enable_hme_level1;
    pcs->enable_hme_level2_flag = (EbBool)prm->enable_hme_level2;
    pcs->slice_type = (prm->num_lists == 1) ? P_SLICE : B_SLICE;
    pcs->temporal_layer_index = (uint8_t)prm->temporal_layer_index;
    pcs->hierarchical_levels = (uint8_t)prm->hierarchical_levels;
    pcs->is_used_as_reference_flag = (EbBool)prm->is_used_as_reference;
    pcs->use_subpel_flag = (uint8_t)prm->use_subpel;
    pcs->cu8x8_mode = (uint8_t)prm->cu8x8_mode;
    pcs->enc_mode = ENC_M3;
    pcs->ref_pa_pic_ptr_array[0] = &pa_ref[0].wrapper;
    pcs->ref_pa_pic_ptr_array[1] = &pa_ref[1].wrapper;
    pcs->ref_pic_poc_array[0] = 100;
    pcs->ref_pic_poc_array[1] = prm->same_ref_poc ? 100 : 200;
    pcs->enhanced_picture_ptr = &pa_cur.full;
    pcs->me_results = (MeCuResults_t **)calloc(n_sb, sizeof(MeCuResults_t *));
    for (uint32_t i = 0; i < n_sb; i++)
        pcs->me_results[i] = (MeCuResults_t *)calloc(MEO_MAX_PU_COUNT, sizeof(MeCuResults_t));

    static MeContext_t *ctx = NULL;              /* built once; reference allocations are never freed */
    if (!ctx && MeContextCtor(&ctx) != EB_ErrorNone) return -1;
    /* search-area parameters: take them from MeoParams (they default to the same tables) */
    set_me_hme_params_oq(ctx, pcs, scs, scs->input_resolution);
    ctx->number_hme_search_region_in_width = (uint16_t)prm->num_hme_regions_w;
    ctx->number_hme_search_region_in_height = (uint16_t)prm->num_hme_regions_h;
    ctx->hme_level0_total_search_area_width = (uint16_t)prm->hme_l0_total_w;
    ctx->hme_level0_total_search_area_height = (uint16_t)prm->hme_l0_total_h;
    for (int i = 0; i < 2; i++) {
        ctx->hme_level0_search_area_in_width_array[i] = (uint16_t)prm->hme_l0_w[i];
        ctx->hme_level0_search_area_in_height_array[i] = (uint16_t)prm->hme_l0_h[i];
        ctx->hme_level1_search_area_in_width_array[i] = (uint16_t)prm->hme_l1_w[i];
        ctx->hme_level1_search_area_in_height_array[i] = (uint16_t)prm->hme_l1_h[i];
        ctx->hme_level2_search_area_in_width_array[i] = (uint16_t)prm->hme_l2_w[i];
        ctx->hme_level2_search_area_in_height_array[i] = (uint16_t)prm->hme_l2_h[i];
    }
    ctx->search_area_width = (uint8_t)prm->search_area_w;
    ctx->search_area_height = (uint8_t)prm->search_area_h;
    ctx->lambda = 0;

    EbPictureBufferDesc_t *in = &pa_cur.full, *qd = &pa_cur.quarter, *sd = &pa_cur.sixteenth;
    for (uint32_t ys = 0; ys < h_sb; ys++) {
        for (uint32_t xs = 0; xs < w_sb; xs++) {
            uint32_t sb_index = xs + ys * w_sb, ox = xs * 64, oy = ys * 64;
            uint32_t sb_w = (prm->width - ox) < 64 ? prm->width - ox : 64;
            uint32_t sb_h = (prm->height - oy) < 64 ? prm->height - oy : 64;
            uint32_t bi = (in->origin_y + oy) * in->strideY + in->origin_x + ox, r;
            ctx->hme_search_type = HME_RECTANGULAR;
            for (r = 0; r < 64; r++)
                memcpy(&ctx->sb_buffer[r * 64], &in->bufferY[bi + r * in->strideY], 64);
            ctx->sb_src_ptr = &in->bufferY[bi];
            ctx->sb_src_stride = in->strideY;
            if (pcs->enable_hme_level1_flag) {
                bi = (qd->origin_y + (oy >> 1)) * qd->strideY + qd->origin_x + (ox >> 1);
                for (r = 0; r < (sb_h >> 1); r++)
                    memcpy(&ctx->quarter_sb_buffer[r * ctx->quarter_sb_buffer_stride], &qd->bufferY[bi + r * qd->strideY], sb_w >> 1);
            }
            if (pcs->enable_hme_level0_flag) {
                bi = (sd->origin_y + (oy >> 2)) * sd->strideY + sd->origin_x + (ox >> 2);
                uint8_t *fp = &sd->bufferY[bi], *lp = ctx->sixteenth_sb_buffer;
                for (r = 0; r < (sb_h >> 2); r += 2) {
                    memcpy(lp, fp, sb_w >> 2);
                    lp += 16; fp += sd->strideY << 1;
                }
            }
            MotionEstimateLcu(pcs, sb_index, ox, oy, ctx, in);

            for (int pu = 0; pu < prm->num_pus; pu++) {
                const MeCuResults_t *m = &pcs->me_results[sb_index][pu];
                MeoPuResult *o = &results[(size_t)sb_index * prm->num_pus + pu];
                memset(o, 0, sizeof(*o));
                o->xMvL0 = m->xMvL0; o->yMvL0 = m->yMvL0; o->xMvL1 = m->xMvL1; o->yMvL1 = m->yMvL1;
                o->total_cand = m->totalMeCandidateIndex;
                for (int c = 0; c < m->totalMeCandidateIndex && c < 3; c++) {
                    o->cand[c].distortion = m->distortionDirection[c].distortion;
                    o->cand[c].direction = m->distortionDirection[c].direction;
                }
            }
            if (debug) {
                MeoSbDebug *d = &debug[sb_index];
                memset(d, 0, sizeof(*d));
                for (int l = 0; l < prm->num_lists; l++) {
                    d->sa_origin_x[l] = ctx->x_search_area_origin[l][0];
                    d->sa_origin_y[l] = ctx->y_search_area_origin[l][0];
                    for (int i = 0; i < MEO_SQUARE_PU_COUNT; i++) {
                        d->best_sad[l][i] = ctx->p_sb_best_sad[l][0][i];
                        d->best_mv[l][i] = ctx->p_sb_best_mv[l][0][i];
                    }
                }
            }
        }
    }
    for (uint32_t i = 0; i < n_sb; i++) free(pcs->me_results[i]);
    free(pcs->me_results);
    /* MeContext buffers are owned by the (leaky by design) EB_MALLOC map of the reference */
    free(pa_cur.obj); free(pa_ref[0].obj); free(pa_ref[1].obj);
    free(pcs); free(enc); free(scs);
    return 0;
}

/* ComputeDecimatedZzSad of the reference on one picture pair: the minimum of SCS / PCS / context it reads. */
extern EbErrorType ComputeDecimatedZzSad(MotionEstimationContext_t *, SequenceControlSet_t *, PictureParentControlSet_t *,
                                         EbPictureBufferDesc_t *, uint32_t, uint32_t, uint32_t, uint32_t);
int ref_zz_sad_picture(int32_t width, int32_t height, const MeoPyramid *cur, const MeoPyramid *prev,
                       uint8_t *zz_cost, uint8_t *non_moving_index) {
    ensure_memory_map();
    const uint32_t w_sb = (width + 63) / 64, h_sb = (height + 63) / 64, n_sb = w_sb * h_sb;
    SequenceControlSet_t *scs = (SequenceControlSet_t *)calloc(1, sizeof(SequenceControlSet_t));
    EncodeContext_t *enc = (EncodeContext_t *)calloc(1, sizeof(EncodeContext_t));
    PictureParentControlSet_t *pcs = (PictureParentControlSet_t *)calloc(1, sizeof(PictureParentControlSet_t));
    PictureParentControlSet_t *prev_pcs = (PictureParentControlSet_t *)calloc(1, sizeof(PictureParentControlSet_t));
    MotionEstimationContext_t *mctx = (MotionEstimationContext_t *)calloc(1, sizeof(MotionEstimationContext_t));
    EbObjectWrapper_t prev_wrapper; memset(&prev_wrapper, 0, sizeof(prev_wrapper));
    enc->asm_type = ASM_AVX2;
    scs->encode_context_ptr = enc;
    scs->luma_width = (uint16_t)width; scs->luma_height = (uint16_t)height; scs->sb_sz = 64;
    scs->picture_width_in_sb = (uint8_t)w_sb; scs->picture_height_in_sb = (uint8_t)h_sb;
    scs->sb_params_array = (SbParams_t *)calloc(n_sb, sizeof(SbParams_t));
    for (uint32_t i = 0; i < n_sb; i++) {                      /* sb_params_init, EbSequenceControlSet.c:395-420 */
        SbParams_t *sp = &scs->sb_params_array[i];
        sp->origin_x = (uint16_t)((i % w_sb) * 64); sp->origin_y = (uint16_t)((i / w_sb) * 64);
        sp->width = (uint8_t)(width - sp->origin_x < 64 ? width - sp->origin_x : 64);
        sp->height = (uint8_t)(height - sp->origin_y < 64 ? height - sp->origin_y : 64);
        sp->is_complete_sb = (uint8_t)(sp->width == 64 && sp->height == 64);
    }
    static MeContext_t *ctx = NULL;
    if (!ctx && MeContextCtor(&ctx) != EB_ErrorNone) return -1;
    mctx->me_context_ptr = ctx;
    EbPictureBufferDesc_t prev_full, cur16;
    plane_to_desc(&prev->full, &prev_full);
    plane_to_desc(&cur->sixteenth, &cur16);
    prev_pcs->enhanced_picture_ptr = &prev_full;
    prev_pcs->zz_cost_array = zz_cost;
    prev_pcs->non_moving_index_array = non_moving_index;
    prev_wrapper.objectPtr = prev_pcs;
    pcs->previous_picture_control_set_wrapper_ptr = &prev_wrapper;
    const EbErrorType rc = ComputeDecimatedZzSad(mctx, scs, pcs, &cur16, 0, w_sb, 0, h_sb);
    free(scs->sb_params_array); free(mctx); free(prev_pcs); free(pcs); free(enc); free(scs);
    return rc == EB_ErrorNone ? 0 : -1;
}

/* ---- OpenLoopIntraSearchLcu of the reference on one picture: the minimum of SCS / PCS / context it reads ------------- */
#include "EbIntraPrediction.h"
extern EbErrorType OpenLoopIntraSearchLcu(PictureParentControlSet_t *, uint32_t, MotionEstimationContext_t *, EbPictureBufferDesc_t *, EbAsm);
static MotionEstimationContext_t *ois_context(void) {
    static MotionEstimationContext_t *mctx = NULL;
    if (!mctx) {
        ensure_memory_map();
        mctx = (MotionEstimationContext_t *)calloc(1, sizeof(MotionEstimationContext_t));
        if (MeContextCtor(&mctx->me_context_ptr) != EB_ErrorNone) return NULL;
        if (IntraOpenLoopReferenceSamplesCtor(&mctx->intra_ref_ptr) != EB_ErrorNone) return NULL;
    }
    return mctx;
}
uint32_t ref_ois_predict_cu(const MeoPlane *pic, int32_t x0, int32_t y0, int32_t n, int32_t mode, uint8_t *pred64) {
    ensure_memory_map();
    MotionEstimationContext_t *mctx = ois_context();
    EbPictureBufferDesc_t in;
    plane_to_desc(pic, &in);
    UpdateNeighborSamplesArrayOpenLoop(mctx->intra_ref_ptr, &in, in.strideY, (uint32_t)x0, (uint32_t)y0, (uint32_t)n);
    memset(mctx->me_context_ptr->sb_buffer, 0, 64 * 64);
    IntraPredictionOpenLoop((uint32_t)n, mctx, (uint32_t)mode, ASM_AVX2);
    memcpy(pred64, mctx->me_context_ptr->sb_buffer, 64 * 64);
    return (uint32_t)NxMSadKernel_funcPtrArray[ASM_AVX2][n >> 3](
        &in.bufferY[(in.origin_y + y0) * in.strideY + in.origin_x + x0], in.strideY, mctx->me_context_ptr->sb_buffer, 64, (uint32_t)n, (uint32_t)n);
}
int ref_ois_picture(const MeoOisParams *prm, const MeoPlane *pic, const uint32_t *me_sad, MeoOisCu *out) {
    ensure_memory_map();
    MotionEstimationContext_t *mctx = ois_context();
    if (!mctx) return -1;
    const uint32_t w_sb = (prm->width + 63) / 64, h_sb = (prm->height + 63) / 64, n_sb = w_sb * h_sb;
    SequenceControlSet_t *scs = (SequenceControlSet_t *)calloc(1, sizeof(SequenceControlSet_t));
    EncodeContext_t *enc = (EncodeContext_t *)calloc(1, sizeof(EncodeContext_t));
    PictureParentControlSet_t *pcs = (PictureParentControlSet_t *)calloc(1, sizeof(PictureParentControlSet_t));
    EbObjectWrapper_t scs_wrapper; memset(&scs_wrapper, 0, sizeof(scs_wrapper));
    enc->asm_type = ASM_AVX2;
    scs->encode_context_ptr = enc;
    scs->luma_width = (uint16_t)prm->width; scs->luma_height = (uint16_t)prm->height; scs->sb_sz = 64;
    scs->input_resolution = (EbInputResolution)prm->input_resolution;
    scs->picture_width_in_sb = (uint8_t)w_sb; scs->picture_height_in_sb = (uint8_t)h_sb;
    scs->sb_params_array = (SbParams_t *)calloc(n_sb, sizeof(SbParams_t));
    for (uint32_t i = 0; i < n_sb; i++) {                      /* sb_params_init, EbSequenceControlSet.c:395-440 */
        SbParams_t *sp = &scs->sb_params_array[i];
        sp->origin_x = (uint16_t)((i % w_sb) * 64); sp->origin_y = (uint16_t)((i / w_sb) * 64);
        sp->width = (uint8_t)(prm->width - sp->origin_x < 64 ? prm->width - sp->origin_x : 64);
        sp->height = (uint8_t)(prm->height - sp->origin_y < 64 ? prm->height - sp->origin_y : 64);
        sp->is_complete_sb = (uint8_t)(sp->width == 64 && sp->height == 64);
        for (uint32_t cu = 0; cu < CU_MAX_COUNT; cu++)
            sp->raster_scan_cu_validity[cu] = (EbBool)(sp->origin_x + RASTER_SCAN_CU_X[cu] + RASTER_SCAN_CU_SIZE[cu] <= (uint32_t)prm->width &&
                                                       sp->origin_y + RASTER_SCAN_CU_Y[cu] + RASTER_SCAN_CU_SIZE[cu] <= (uint32_t)prm->height);
    }
    scs_wrapper.objectPtr = scs;
    pcs->sequence_control_set_wrapper_ptr = &scs_wrapper;
    pcs->slice_type = prm->slice_type == 0 ? I_SLICE : prm->slice_type == 1 ? P_SLICE : B_SLICE;
    pcs->temporal_layer_index = (uint8_t)prm->temporal_layer_index;
    pcs->is_used_as_reference_flag = (EbBool)prm->is_used_as_reference;
    pcs->cu8x8_mode = (uint8_t)prm->cu8x8_mode;
    pcs->limit_ois_to_dc_mode_flag = (EbBool)prm->limit_ois_to_dc;
    pcs->enc_mode = (EbEncMode)prm->enc_mode;
    pcs->ois_cu32_cu16_results = (OisCu32Cu16Results_t **)calloc(n_sb, sizeof(void *));
    pcs->ois_cu8_results = (OisCu8Results_t **)calloc(n_sb, sizeof(void *));
    pcs->me_results = (MeCuResults_t **)calloc(n_sb, sizeof(MeCuResults_t *));
    OisCandidate_t *c32 = (OisCandidate_t *)calloc((size_t)n_sb * 21 * 18, sizeof(OisCandidate_t));
    OisCandidate_t *c8 = (OisCandidate_t *)calloc((size_t)n_sb * 64 * 18, sizeof(OisCandidate_t));
    for (uint32_t i = 0; i < n_sb; i++) {
        pcs->ois_cu32_cu16_results[i] = (OisCu32Cu16Results_t *)calloc(1, sizeof(OisCu32Cu16Results_t));
        pcs->ois_cu8_results[i] = (OisCu8Results_t *)calloc(1, sizeof(OisCu8Results_t));
        for (int cu = 0; cu < 21; cu++) pcs->ois_cu32_cu16_results[i]->sorted_ois_candidate[cu] = c32 + ((size_t)i * 21 + cu) * 18;
        for (int cu = 0; cu < 64; cu++) pcs->ois_cu8_results[i]->sorted_ois_candidate[cu] = c8 + ((size_t)i * 64 + cu) * 18;
        pcs->me_results[i] = (MeCuResults_t *)calloc(85, sizeof(MeCuResults_t));
        if (me_sad) for (int pu = 0; pu < 85; pu++) pcs->me_results[i][pu].distortionDirection[0].distortion = me_sad[(size_t)i * 85 + pu];
    }
    EbPictureBufferDesc_t in;
    plane_to_desc(pic, &in);
    for (uint32_t i = 0; i < n_sb; i++) OpenLoopIntraSearchLcu(pcs, i, mctx, &in, ASM_AVX2);
    for (uint32_t i = 0; i < n_sb; i++) {
        for (int cu = 1; cu < 85; cu++) {
            const OisCandidate_t *src = cu < 21 ? pcs->ois_cu32_cu16_results[i]->sorted_ois_candidate[cu] : pcs->ois_cu8_results[i]->sorted_ois_candidate[cu - 21];
            MeoOisCu *o = &out[(size_t)i * 85 + cu];
            for (int k = 0; k < 18; k++) {
                o->cand[k].distortion = src[k].distortion; o->cand[k].mode = (uint8_t)src[k].intra_mode; o->cand[k].valid = (uint8_t)src[k].valid_distortion;
            }
            o->total = cu < 21 ? pcs->ois_cu32_cu16_results[i]->total_intra_luma_mode[cu] : pcs->ois_cu8_results[i]->total_intra_luma_mode[cu - 21];
        }
        free(pcs->ois_cu32_cu16_results[i]); free(pcs->ois_cu8_results[i]); free(pcs->me_results[i]);
    }
    free(c32); free(c8); free(pcs->ois_cu32_cu16_results); free(pcs->ois_cu8_results); free(pcs->me_results);
    free(scs->sb_params_array); free(pcs); free(enc); free(scs);
    return 0;
}

/* Reference values of the search-area parameters (set_me_hme_params_oq + the L0 multiplier tables),
 * used to pin meo_default_params(). */
void ref_default_params(MeoParams *p, int32_t width, int32_t height, int32_t num_lists,
                        int32_t temporal_layer_index, int32_t hierarchical_levels, int32_t is_ref) {
    ensure_memory_map();
    SequenceControlSet_t *scs = (SequenceControlSet_t *)calloc(1, sizeof(SequenceControlSet_t));
    PictureParentControlSet_t *pcs = (PictureParentControlSet_t *)calloc(1, sizeof(PictureParentControlSet_t));
    MeContext_t *ctx = (MeContext_t *)calloc(1, sizeof(MeContext_t));
    scs->luma_width = (uint16_t)width; scs->luma_height = (uint16_t)height;
    derive_input_resolution(scs, (uint32_t)width * (uint32_t)height);
    pcs->enc_mode = ENC_M3;
    set_me_hme_params_oq(ctx, pcs, scs, scs->input_resolution);
    memset(p, 0, sizeof(*p));
    p->width = width; p->height = height; p->num_lists = num_lists;
    p->temporal_layer_index = temporal_layer_index; p->hierarchical_levels = hierarchical_levels;
    p->is_used_as_reference = is_ref;
    p->enable_hme_flag = 1; p->enable_hme_level0 = 1; p->enable_hme_level1 = 1;
    {   /* signal_derivation_pre_analysis_oq, EbResourceCoordinationProcess.c:366-380 (hme_me_level = enc_mode = 3) */
        uint32_t ratio = scs->luma_width / scs->luma_height; uint8_t ir = scs->input_resolution;
        uint8_t ri = ir <= INPUT_SIZE_576p_RANGE_OR_LOWER ? 0 : (ir <= INPUT_SIZE_1080i_RANGE && ratio < 2) ? 1 :
                     (ir <= INPUT_SIZE_1080i_RANGE && ratio > 3) ? 2 : (ir <= INPUT_SIZE_1080p_RANGE) ? 3 : 4;
        p->enable_hme_level0 = EnableHmeLevel0Flag[ri][ENC_M3];
        p->enable_hme_level1 = EnableHmeLevel1Flag[ri][ENC_M3];
        p->enable_hme_level2 = EnableHmeLevel2Flag[ri][ENC_M3];
    }
    p->use_subpel = 1; p->num_pus = MEO_SQUARE_PU_COUNT; p->cu8x8_mode = 0;
    p->num_hme_regions_w = ctx->number_hme_search_region_in_width;
    p->num_hme_regions_h = ctx->number_hme_search_region_in_height;
    p->hme_l0_total_w = ctx->hme_level0_total_search_area_width;
    p->hme_l0_total_h = ctx->hme_level0_total_search_area_height;
    for (int i = 0; i < 2; i++) {
        p->hme_l0_w[i] = ctx->hme_level0_search_area_in_width_array[i];
        p->hme_l0_h[i] = ctx->hme_level0_search_area_in_height_array[i];
        p->hme_l1_w[i] = ctx->hme_level1_search_area_in_width_array[i];
        p->hme_l1_h[i] = ctx->hme_level1_search_area_in_height_array[i];
        p->hme_l2_w[i] = ctx->hme_level2_search_area_in_width_array[i];
        p->hme_l2_h[i] = ctx->hme_level2_search_area_in_height_array[i];
    }
    p->search_area_w = ctx->search_area_width; p->search_area_h = ctx->search_area_height;
    p->hme_l0_mult_x = HME_LEVEL_0_SEARCH_AREA_MULTIPLIER_X[hierarchical_levels][temporal_layer_index];
    p->hme_l0_mult_y = HME_LEVEL_0_SEARCH_AREA_MULTIPLIER_Y[hierarchical_levels][temporal_layer_index];
    free(ctx); free(pcs); free(scs);
}


/* ---- ME dump / ME hook of real encodes (golden vectors of BASELINE configs[0]; encoder-in-the-loop test) -------------
 * Dump: record per MotionEstimateLcu call what the encoder's own processes set up for it (slice type, layers, HME flags,
 * the MeContext search parameters, reference POCs), the source luma once per picture, and the MeCuResults_t rows it
 * produced. Layout: a stream of int32 records, see tests/golden/make_encoder_golden.py.
 * Hook: when one is installed (oracle/enc_driver.c --cuda-me) the reference's MotionEstimateLcu is NOT run; the hook
 * receives the same header, the luma planes of the source and of the two PA reference pictures, and fills
 * me_results[sb_index] itself. Everything else the encoder does stays the reference's. */
#include <pthread.h>
#include <stdio.h>
typedef int (*ref_me_hook_fn)(void *user, const int32_t *hdr, const uint8_t *cur, int cur_stride, const uint8_t *ref0, int ref0_stride,
                              const uint8_t *ref1, int ref1_stride, void *row_out);
static FILE *g_dump;
static pthread_mutex_t g_dump_mu = PTHREAD_MUTEX_INITIALIZER;
static uint64_t g_seen[4096]; static int g_nseen;
static ref_me_hook_fn g_hook; static void *g_hook_user;

int ref_me_dump_open(const char *path) { g_dump = fopen(path, "wb"); g_nseen = 0; return g_dump ? 0 : -1; }
void ref_me_dump_close(void) { if (g_dump) fclose(g_dump); g_dump = NULL; }
void ref_me_hook_set(ref_me_hook_fn fn, void *user) { g_hook = fn; g_hook_user = user; }

static const uint8_t *luma_origin(const EbPictureBufferDesc_t *p) { return p->bufferY + (size_t)p->origin_y * p->strideY + p->origin_x; }

/* ME-stage time: nanoseconds spent inside MotionEstimateLcu summed over the encoder's ME threads, and the number of calls
 * (profiles/encoder_timing.py: stock vs hooked encoder) */
#include <time.h>
static long long g_me_ns, g_me_calls;
static int g_me_timing;
void ref_me_timing_enable(int on) { g_me_timing = on; __atomic_store_n(&g_me_ns, 0, __ATOMIC_RELAXED); __atomic_store_n(&g_me_calls, 0, __ATOMIC_RELAXED); }
void ref_me_timing_get(long long *ns, long long *calls) { *ns = __atomic_load_n(&g_me_ns, __ATOMIC_RELAXED); *calls = __atomic_load_n(&g_me_calls, __ATOMIC_RELAXED); }
static EbErrorType motion_estimate_lcu_impl(PictureParentControlSet_t *pcs, uint32_t sb_index, uint32_t sb_origin_x, uint32_t sb_origin_y,
                                            MeContext_t *ctx, EbPictureBufferDesc_t *input_ptr);
EbErrorType MotionEstimateLcu(PictureParentControlSet_t *pcs, uint32_t sb_index, uint32_t sb_origin_x, uint32_t sb_origin_y,
                              MeContext_t *ctx, EbPictureBufferDesc_t *input_ptr) {
    if (!g_me_timing) return motion_estimate_lcu_impl(pcs, sb_index, sb_origin_x, sb_origin_y, ctx, input_ptr);
    struct timespec a, b;
    clock_gettime(CLOCK_MONOTONIC, &a);
    const EbErrorType rc = motion_estimate_lcu_impl(pcs, sb_index, sb_origin_x, sb_origin_y, ctx, input_ptr);
    clock_gettime(CLOCK_MONOTONIC, &b);
    __atomic_fetch_add(&g_me_ns, (b.tv_sec - a.tv_sec) * 1000000000LL + (b.tv_nsec - a.tv_nsec), __ATOMIC_RELAXED);
    __atomic_fetch_add(&g_me_calls, 1, __ATOMIC_RELAXED);
    return rc;
}
static EbErrorType motion_estimate_lcu_impl(PictureParentControlSet_t *pcs, uint32_t sb_index, uint32_t sb_origin_x, uint32_t sb_origin_y,
                                            MeContext_t *ctx, EbPictureBufferDesc_t *input_ptr) {
    if (!g_hook && !g_dump) return MotionEstimateLcu_reference(pcs, sb_index, sb_origin_x, sb_origin_y, ctx, input_ptr);
    SequenceControlSet_t *scs = (SequenceControlSet_t *)pcs->sequence_control_set_wrapper_ptr->objectPtr;
    int32_t h[48]; memset(h, 0, sizeof(h));
    h[0] = 0x4D454447;                                    /* 'MEDG' */
    h[1] = (int32_t)pcs->picture_number; h[2] = (int32_t)sb_index; h[3] = (int32_t)sb_origin_x; h[4] = (int32_t)sb_origin_y;
    h[5] = scs->luma_width; h[6] = scs->luma_height;
    h[7] = pcs->slice_type == P_SLICE ? 1 : pcs->slice_type == B_SLICE ? 2 : 0;
    h[8] = pcs->temporal_layer_index; h[9] = pcs->hierarchical_levels; h[10] = pcs->is_used_as_reference_flag;
    h[11] = (int32_t)pcs->ref_pic_poc_array[0]; h[12] = (int32_t)pcs->ref_pic_poc_array[1];
    h[13] = pcs->enable_hme_flag; h[14] = pcs->enable_hme_level0_flag; h[15] = pcs->enable_hme_level1_flag; h[16] = pcs->enable_hme_level2_flag;
    h[17] = pcs->use_subpel_flag; h[18] = pcs->max_number_of_pus_per_sb; h[19] = pcs->cu8x8_mode;
    h[20] = ctx->number_hme_search_region_in_width; h[21] = ctx->number_hme_search_region_in_height;
    h[22] = ctx->hme_level0_total_search_area_width; h[23] = ctx->hme_level0_total_search_area_height;
    for (int i = 0; i < 2; i++) {
        h[24 + i] = ctx->hme_level0_search_area_in_width_array[i]; h[26 + i] = ctx->hme_level0_search_area_in_height_array[i];
        h[28 + i] = ctx->hme_level1_search_area_in_width_array[i]; h[30 + i] = ctx->hme_level1_search_area_in_height_array[i];
        h[32 + i] = ctx->hme_level2_search_area_in_width_array[i]; h[34 + i] = ctx->hme_level2_search_area_in_height_array[i];
    }
    h[36] = ctx->search_area_width; h[37] = ctx->search_area_height;
    h[38] = pcs->enc_mode; h[39] = pcs->pic_depth_mode;
    h[41] = (int32_t)sizeof(MeCuResults_t);
    EbErrorType rc = EB_ErrorNone;
    if (g_hook) {
        const EbPictureBufferDesc_t *r0 = ((EbPaReferenceObject_t *)pcs->ref_pa_pic_ptr_array[0]->objectPtr)->inputPaddedPicturePtr;
        const EbPictureBufferDesc_t *r1 = h[7] == 2 ? ((EbPaReferenceObject_t *)pcs->ref_pa_pic_ptr_array[1]->objectPtr)->inputPaddedPicturePtr : r0;
        if (g_hook(g_hook_user, h, luma_origin(input_ptr), input_ptr->strideY, luma_origin(r0), r0->strideY, luma_origin(r1), r1->strideY,
                   pcs->me_results[sb_index]) != 0) {
            fprintf(stderr, "ref_driver: the ME hook failed (picture %d, SB %d)\n", h[1], h[2]);
            abort();                                      /* no silent fallback to the CPU path */
        }
        if (scs->static_config.rate_control_mode) {       /* EbMotionEstimation.c:7150-7157 */
            pcs->rc_me_distortion[sb_index] = 0;
            for (int i = 0; i < 16; i++) pcs->rc_me_distortion[sb_index] += pcs->me_results[sb_index][5 + i].distortionDirection[0].distortion;
        }
    } else {
        rc = MotionEstimateLcu_reference(pcs, sb_index, sb_origin_x, sb_origin_y, ctx, input_ptr);
    }
    if (!g_dump) return rc;
    pthread_mutex_lock(&g_dump_mu);
    int first = 1;
    for (int i = 0; i < g_nseen; i++) if (g_seen[i] == pcs->picture_number) first = 0;
    if (first && g_nseen < 4096) g_seen[g_nseen++] = pcs->picture_number;
    h[40] = first;
    fwrite(h, sizeof(h), 1, g_dump);
    if (first)                                            /* source luma of this picture (interior of the padded input plane) */
        for (uint32_t y = 0; y < scs->luma_height; y++)
            fwrite(luma_origin(input_ptr) + (size_t)y * input_ptr->strideY, 1, scs->luma_width, g_dump);
    fwrite(pcs->me_results[sb_index], sizeof(MeCuResults_t), pcs->max_number_of_pus_per_sb, g_dump);
    pthread_mutex_unlock(&g_dump_mu);
    return rc;
}
