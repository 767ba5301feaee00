/* include/svtme_lcu.h -- SB-granular facade of libsvtme_cuda.so.
 *
 * The reference encoder calls its ME operator once per 64x64 SB, from several ME threads at once
 *   EbErrorType MotionEstimateLcu(PictureParentControlSet_t*, uint32_t sb_index, uint32_t sb_origin_x,
 *                                 uint32_t sb_origin_y, MeContext_t*, EbPictureBufferDesc_t *input_ptr)
 *   Source/Lib/Codec/EbMotionEstimation.c:6152, called from MotionEstimationKernel,
 *   Source/Lib/Codec/EbMotionEstimationProcess.c:546-552 (one call per SB of a segment, 60 segments per picture).
 * This facade keeps that call shape -- one call per SB, any thread, any order -- on top of the picture-level
 * entry points of svtme.h: the first SB call of a picture uploads whichever of its three pictures (source, list-0
 * and list-1 reference) are not yet resident in HBM, motion-estimates the WHOLE picture in one launch sequence and
 * keeps the rows; every call, that one included, then copies out the row of its SB, i.e. what MotionEstimateLcu
 * leaves in PictureParentControlSet_t::me_results[sb_index] (EbMotionEstimation.c:7047-7145). The rows of a picture
 * are dropped once each of its SBs has been fetched.
 *
 * A hook that swaps this in for MotionEstimateLcu needs nothing else from the encoder: INTEGRATION.md section 6 shows
 * it, and tests/test_encoder_in_loop_gpu.py runs the unmodified reference encoder with exactly that hook and checks
 * that the bitstream does not change.
 *
 * Pictures are identified by picture_number (PictureParentControlSet_t::picture_number, the POC of the stream):
 * the luma of a picture_number must never change while the session lives, which is what the reference's PA
 * reference pool guarantees. Errors mirror svtme.h. There is no CPU fallback.
 */
#ifndef SVTME_LCU_H
#define SVTME_LCU_H
#include "svtme.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct SvtMeLcuSession SvtMeLcuSession;    /* opaque; thread-safe */

/* One 8-bit luma plane in HOST memory: pixel (0,0) of the picture (not of its padding), 'stride' bytes per row.
 * = EbPictureBufferDesc_t::bufferY + origin_y * strideY + origin_x of the input / PA reference picture. */
typedef struct SvtMeLcuPicture {
    uint64_t       picture_number;
    const uint8_t *luma;
    int32_t        stride;
    int32_t        reserved;
} SvtMeLcuPicture;

/* num_slots >= 3: pictures kept resident in HBM (least recently used is replaced). A hierarchical mini-GOP of 8
 * touches 9 pictures, so 16 avoids every re-upload. */
int32_t svtme_lcu_session_create(SvtMeLcuSession **session, int32_t device, int32_t width, int32_t height, int32_t num_slots);
void    svtme_lcu_session_destroy(SvtMeLcuSession *session);
const char *svtme_lcu_last_error(const SvtMeLcuSession *session);

/* The MotionEstimateLcu-shaped call. 'params' must be the same for every SB of a picture (it is per picture in the
 * reference too); ref1 is NULL for P pictures (params->num_lists == 1). row_out receives params->num_pus results. */
int32_t svtme_lcu_motion_estimate(SvtMeLcuSession *session, const SvtMeParams *params, const SvtMeLcuPicture *cur,
                                  const SvtMeLcuPicture *ref0, const SvtMeLcuPicture *ref1, int32_t sb_index,
                                  SvtMePuResult *row_out);

/* The same for one ME SEGMENT, the unit the reference's ME threads work in (MotionEstimationKernel takes one segment of
 * me_segments_column_count x me_segments_row_count per input message, EbMotionEstimationProcess.c:432-439, and runs the SB
 * loop :478-556 over it): SB columns [sb_x0, sb_x1) x SB rows [sb_y0, sb_y1). rows_out[i] = where the row of the i-th SB
 * of the rectangle (raster order) goes, normally PictureParentControlSet_t::me_results[sb_index]; rc_me_distortion (may be
 * NULL) receives rc_me_distortion[sb_index] for the same SBs (EbMotionEstimation.c:7150-7157). The first segment of a
 * picture to arrive submits the whole picture to the GPU, every segment then copies its rows: this is the picture-level
 * drop-in of integration/svt_cuda_me.patch. */
int32_t svtme_lcu_motion_estimate_segment(SvtMeLcuSession *session, const SvtMeParams *params, const SvtMeLcuPicture *cur,
                                          const SvtMeLcuPicture *ref0, const SvtMeLcuPicture *ref1, int32_t sb_x0, int32_t sb_x1,
                                          int32_t sb_y0, int32_t sb_y1, SvtMePuResult *const *rows_out, uint32_t *rc_me_distortion);

/* Counters for tests and logs: pictures motion-estimated, picture uploads, SB rows served. */
void    svtme_lcu_session_stats(const SvtMeLcuSession *session, int64_t *pictures, int64_t *uploads, int64_t *rows);

#ifdef __cplusplus
}
#endif
#endif /* SVTME_LCU_H */
