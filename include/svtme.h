/* include/svtme.h -- C-ABI of libsvtme_cuda.so: SVT-AV1 open-loop motion estimation on B200.
 *
 * Drop-in boundary. Each entry point replaces a piece of the reference's CPU path; the host
 * encoder (plain C) keeps its own structures and calls these with plain pointers and sizes:
 *
 *   svtme_picture_upload        replaces PadPictureToMultipleOfLcuDimensions + DecimateInputPicture
 *                               Source/Lib/Codec/EbPictureAnalysisProcess.c:4866-4937 (generate_padding
 *                               EbMcp.c:173-216, Decimation2D :100-125): the padded full / quarter /
 *                               sixteenth planes of EbPaReferenceObject_t live in HBM instead of host RAM.
 *   svtme_motion_estimate       replaces the SB loop of MotionEstimationKernel,
 *                               Source/Lib/Codec/EbMotionEstimationProcess.c:478-556, i.e. one
 *                               MotionEstimateLcu (Source/Lib/Codec/EbMotionEstimation.c:6152) per SB,
 *                               for whole pictures (all ME segments at once) and for a batch of pictures.
 *   SvtMePuResult               is the byte image of MeCuResults_t
 *                               (Source/Lib/Codec/EbMotionEstimationLcuResults.h:62-75); the host copies
 *                               rows straight into PictureParentControlSet_t::me_results[sb][pu].
 *   SvtMeParams                 carries what MotionEstimateLcu reads from PCS/SCS/MeContext
 *                               (slice type, temporal layer, HME enables, search-area tables set by
 *                               set_me_hme_params_oq, EbMotionEstimationProcess.c:94-156).
 *
 * Error behaviour mirrors EbErrorType (Source/API/EbErrorCodes.h): 0 = EB_ErrorNone,
 * SVTME_ERR_INSUFFICIENT_RESOURCES = 0x80001000, SVTME_ERR_BAD_PARAMETER = 0x80001005,
 * SVTME_ERR_UNDEFINED = 0x80001001 (any CUDA failure). There is no CPU fallback.
 */
#ifndef SVTME_H
#define SVTME_H
#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SVTME_OK                          0
#define SVTME_ERR_INSUFFICIENT_RESOURCES  ((int32_t)0x80001000)
#define SVTME_ERR_UNDEFINED               ((int32_t)0x80001001)
#define SVTME_ERR_BAD_PARAMETER           ((int32_t)0x80001005)

#define SVTME_SQUARE_PU_COUNT 85
#define SVTME_MAX_PU_COUNT    209   /* MAX_ME_PU_COUNT: PIC_ALL_DEPTH_MODE pictures (enc-mode 0-2 reference pictures) */
#define SVTME_MAX_BATCH       128

/* kernel kinds, for svtme_kernel_times */
#define SVTME_KERNEL_PYRAMID     0
#define SVTME_KERNEL_HME         1
#define SVTME_KERNEL_FULLPEL     2
#define SVTME_KERNEL_SUBPEL      3
#define SVTME_KERNEL_BIPRED_PACK 4
#define SVTME_KERNEL_OIS         5
#define SVTME_KERNEL_KINDS       6

typedef struct SvtMeContext SvtMeContext;          /* opaque: one per GPU / encoder instance */

/* == MeoParams of the test oracle; field meaning documented there and in DESIGN.md */
typedef struct SvtMeParams {
    int32_t width, height;
    int32_t num_lists;
    int32_t temporal_layer_index;
    int32_t hierarchical_levels;
    int32_t is_used_as_reference;
    int32_t same_ref_poc;
    int32_t enable_hme_flag, enable_hme_level0, enable_hme_level1, enable_hme_level2;
    int32_t use_subpel;
    int32_t num_pus;
    int32_t cu8x8_mode;
    int32_t num_hme_regions_w, num_hme_regions_h;
    int32_t hme_l0_total_w, hme_l0_total_h;
    int32_t hme_l0_w[2], hme_l0_h[2];
    int32_t hme_l1_w[2], hme_l1_h[2];
    int32_t hme_l2_w[2], hme_l2_h[2];
    int32_t search_area_w, search_area_h;
    int32_t hme_l0_mult_x, hme_l0_mult_y;
} SvtMeParams;

/* byte image of MeCuResults_t (40 bytes) */
typedef struct SvtMePuResult {
    int16_t  xMvL0, yMvL0, xMvL1, yMvL1;
    struct { uint32_t distortion; uint32_t direction; } cand[3];
    uint8_t  total_cand;
    uint8_t  pad[7];
} SvtMePuResult;

/* per-SB intermediate state (optional output, for stage-level parity tests) */
typedef struct SvtMeSbDebug {
    int16_t  search_center_x[2], search_center_y[2];
    int16_t  sa_origin_x[2], sa_origin_y[2];
    int16_t  sa_width[2], sa_height[2];
    uint32_t best_sad[2][SVTME_SQUARE_PU_COUNT];
    uint32_t best_mv[2][SVTME_SQUARE_PU_COUNT];
} SvtMeSbDebug;

/* rc_me_distortion[sb] as MotionEstimateLcu leaves it when rate control is on (EbMotionEstimation.c:7150-7157): the sum
 * of the best distortions of the SB's sixteen 16x16 PUs (pu_index 5..20). sb_rows = the SB's result rows. */
uint32_t svtme_rc_me_distortion(const SvtMePuResult *sb_rows);

/* One picture to motion-estimate: slots name pictures previously uploaded into the context. */
typedef struct SvtMeJob {
    int32_t        cur_slot, ref0_slot, ref1_slot;   /* ref1_slot ignored when params.num_lists == 1 */
    /* SB rows [sb_row_begin, sb_row_end) of the picture to search; 0, 0 = the whole picture. The rows of one
     * picture are independent (each SB reads only source pictures), which is how a 4K picture is sharded across
     * GPUs: every GPU holds the pictures, searches its rows and returns only those rows (the ME segments of
     * EbMotionEstimationProcess.c:478-556 are the same partition on the CPU). */
    int32_t        sb_row_begin, sb_row_end;
    /* 0: the range above counts SB rows. 1: it counts SBs in raster order, [sb_row_begin, sb_row_end) of the picture's
     * SBs -- the finer shard (a 4K picture has 34 SB rows but 2040 SBs: eight GPUs take 255 SBs each, whole rows would
     * leave the busiest GPU with 5 rows against 4.25 on average). Results / debug hold the SBs of the range either way. */
    int32_t        range_in_sbs;
    SvtMeParams    params;
    SvtMePuResult *results;                          /* HOST memory, [SBs of the row range][params.num_pus]; may be NULL */
    SvtMeSbDebug  *debug;                            /* HOST memory, [SBs of the row range]; may be NULL */
} SvtMeJob;

/* Default search parameters for a resolution: the search sizes of column M0 of the reference's tables (every legal preset
 * of this version uses them, EbMotionEstimationProcess.c:102) with the HME level enables of ENC-MODE 3 (the only entry
 * that differs between presets is HME level 2 at 4K: on for enc-modes 0-1, off for 2-3, EbResourceCoordinationProcess.c:
 * 366-380) and 85 PUs. A hook that serves an encoder must copy the picture's own flags (enable_hme_*, num_pus, cu8x8_mode)
 * over these, as oracle/enc_driver.c does. hierarchical_levels is clamped to 0..5 and temporal_layer_index to
 * 0..hierarchical_levels. */
void    svtme_default_params(SvtMeParams *p, int32_t width, int32_t height, int32_t num_lists,
                             int32_t temporal_layer_index, int32_t hierarchical_levels,
                             int32_t is_used_as_reference);

/* Context: owns the stream, the picture slots (pyramids in HBM) and all scratch. */
int32_t svtme_create(SvtMeContext **ctx, int32_t device, int32_t width, int32_t height, int32_t num_slots);
void    svtme_destroy(SvtMeContext *ctx);
const char *svtme_last_error(const SvtMeContext *ctx);

/* Upload one 8-bit luma plane (host memory, 'stride' bytes per row) into a slot and build its padded
 * full / quarter / sixteenth planes on the device. Asynchronous, on the context's upload stream; a slot is
 * rewritten only after the batches that read its previous picture have finished. */
int32_t svtme_picture_upload(SvtMeContext *ctx, int32_t slot, const uint8_t *luma, int32_t stride);
/* 10-bit input: 'luma16' holds one sample per uint16_t ('stride' in samples). The 8-bit plane ME works on is
 * sample >> 2 and the 2-bit remainder plane is (uint8_t)(sample << 6): UnPack2D, EbEncHandle.c:2954-2963 ->
 * EB_ENC_msbUnPack2D, C_DEFAULT/EbPackUnPack_C.c:127-151. The split runs on the device, fused in front of the
 * pyramid build; when 'bit_inc' is not NULL it receives the remainder plane (HOST memory, width bytes per row,
 * height rows; in place by svtme_synchronize / svtme_join). */
int32_t svtme_picture_upload_16bit(SvtMeContext *ctx, int32_t slot, const uint16_t *luma16, int32_t stride, uint8_t *bit_inc);
/* Replicate a picture that another context of THIS process (normally on another GPU) already holds: with peer access
 * (enabled on first use) the local pyramid kernel reads the peer slot's picture straight over NVLink while it builds the
 * three padded planes -- no staging copy; without peer access the luma is staged by cudaMemcpy3DPeerAsync first. Ordered
 * on-device after the peer's upload; the peer's next rewrite of that slot waits for this read. This is how reference
 * pictures are replicated when pictures / SB rows of one stream are sharded over the GPUs of a box (no collective,
 * north_star). Threading: neither context may be used by another thread during the call (the peer's slot table is read
 * without a lock); svtme_destroy of either context must not overlap calls on the other. */
int32_t svtme_picture_upload_peer(SvtMeContext *ctx, int32_t slot, SvtMeContext *peer, int32_t peer_slot);
/* Same across PROCESSES (one process per GPU): the owner exports a slot once, ships the 152-byte token to the other
 * ranks by any host channel, and they import pictures from it. Ordering is the caller's: the owner must have ISSUED
 * its upload into the slot before an importer calls svtme_picture_upload_ipc (an inter-process event then orders the
 * copy behind that upload on the device), and must not rewrite the slot until the importers have synchronised. */
typedef struct SvtMeIpcSlot {
    uint8_t  mem_handle[64], event_handle[64];   /* cudaIpcMemHandle_t of the slot allocation, cudaIpcEventHandle_t of its ready event */
    int64_t  luma_offset;                        /* byte offset of pixel (0,0) of the padded full plane inside the allocation */
    int32_t  pitch, device, width, height;
} SvtMeIpcSlot;
int32_t svtme_slot_export(SvtMeContext *ctx, int32_t slot, SvtMeIpcSlot *out);
/* Wait (host) until every upload / import issued so far has landed; searches keep running. After it an importer may
 * tell the owner that the exported slots can be rewritten. */
int32_t svtme_synchronize_uploads(SvtMeContext *ctx);
int32_t svtme_picture_upload_ipc(SvtMeContext *ctx, int32_t slot, const SvtMeIpcSlot *src);
/* Same, from a luma plane that already lives in device memory (no H2D copy). */
int32_t svtme_picture_upload_device(SvtMeContext *ctx, int32_t slot, const uint8_t *d_luma, int32_t stride);
/* A batch of host pictures (n <= SVTME_MAX_BATCH, same row stride): n host-to-device copies followed by ONE pyramid launch
 * for all of them (e.g. the pictures of a mini-GOP as the look-ahead releases them). Asynchronous like svtme_picture_upload. */
int32_t svtme_pictures_upload(SvtMeContext *ctx, int32_t n, const int32_t *slots, const uint8_t *const *lumas, int32_t stride);
/* A batch of pictures (n <= SVTME_MAX_BATCH) that already live in device memory, all with the same row stride: ONE
 * pyramid launch for all of them (a 1080p picture is about a microsecond of HBM time; per-picture launches are
 * launch-latency bound). slots[i] receives d_lumas[i]. */
int32_t svtme_pictures_upload_device(SvtMeContext *ctx, int32_t n, const int32_t *slots, const uint8_t *const *d_lumas, int32_t stride);
/* Read one padded plane back (level 0 = full, 1 = quarter, 2 = sixteenth) as tightly packed rows of
 * (width>>level + 2*pad) bytes; *out_stride / *out_rows receive its geometry. Synchronous. */
int32_t svtme_picture_download(SvtMeContext *ctx, int32_t slot, int32_t level, uint8_t *dst, size_t dst_bytes,
                               int32_t *out_stride, int32_t *out_rows, int32_t *out_pad);

/* Motion-estimate a batch of pictures (n_jobs <= SVTME_MAX_BATCH): kernels run on the compute stream
 * once the pictures they read are uploaded; results/debug are copied to the host buffers of each job
 * and the call returns when they are there. */
int32_t svtme_motion_estimate(SvtMeContext *ctx, const SvtMeJob *jobs, int32_t n_jobs);
/* Enqueue only: searches on the compute stream, then the copies into job.results / job.debug (when not
 * NULL; with NULL the results stay in HBM) on the result stream. Up to three batches are in flight, so
 * uploads, searches and result copies of consecutive batches overlap. svtme_fetch_results waits until the
 * results of every batch submitted so far are in the callers' buffers; svtme_synchronize waits for
 * everything, uploads included. */
int32_t svtme_motion_estimate_async(SvtMeContext *ctx, const SvtMeJob *jobs, int32_t n_jobs);
int32_t svtme_fetch_results(SvtMeContext *ctx, const SvtMeJob *jobs, int32_t n_jobs);
int32_t svtme_synchronize(SvtMeContext *ctx);
/* Make the compute stream wait for all transfers enqueued so far (for timing with events on svtme_stream). */
int32_t svtme_join(SvtMeContext *ctx);

/* Allocation-free variant of svtme_motion_estimate for ONE picture in the 85-PU mode. It has no counterpart in the
 * reference: it was added for callers that cannot call cudaMalloc (code under
 * CUDA graph capture, tight serving loops, harnesses that forbid it). The job table, the inter-kernel SB scratch, the
 * result rows and the debug rows live in ONE file-scope device array of the library (about 6 MB, resident on every
 * GPU that loads the library until the process exits) instead of the context's cudaMalloc'ed batch ring; the call
 * itself calls no cudaMalloc / cudaFree / cudaMallocHost and creates no stream or event. The caller still passes the
 * inputs (pictures uploaded into the context's slots beforehand) and the outputs (job->results / job->debug, host
 * memory). Limits: pictures of at most SVTME_STATIC_MAX_SBS 64x64 SBs (e.g. 1280x720 = 240) and num_pus == 85,
 * anything else returns SVTME_ERR_BAD_PARAMETER with the reason in svtme_last_error. Synchronous: it returns when the
 * rows are in the caller's buffers. Every context of the process on one device shares the array, which keeps the
 * previous call's contents (this call's output rows are cleared first); the library serialises calls of one process
 * with a mutex, and calls that overlap on the GPU by any other route would race on it. */
#define SVTME_STATIC_MAX_SBS 256
int32_t svtme_motion_estimate_static(SvtMeContext *ctx, const SvtMeJob *job);

/* ---- side product of the ME process: zero-motion SADs ---------------------------------------------------------------
 * Replaces ComputeDecimatedZzSad (EbMotionEstimationProcess.c:219-349, called per ME segment at :585-603 when look-ahead
 * is on and picture_number > 0): for every complete 64x64 SB the 16x16 SAD between the current picture's sixteenth plane
 * and the previous picture (display order) sampled every fourth pixel -- both planes are already resident in the slots'
 * pyramids -- classified into zz_cost (BEA_CLASS_*: 0 / 3 / 10 / 20 / 30, incomplete SBs 255) and non_moving_index
 * (0 / 10 / 20 / 30). The reference stores both in the PREVIOUS picture's control set (zz_cost_array,
 * non_moving_index_array); here they are host arrays of one byte per SB in raster order. Either pointer may be NULL.
 * Synchronous: the arrays are filled when the call returns. */
typedef struct SvtMeZzJob {
    int32_t  cur_slot, prev_slot;
    uint8_t *zz_cost;              /* [num_sb] */
    uint8_t *non_moving_index;     /* [num_sb] */
} SvtMeZzJob;
int32_t svtme_zz_sad(SvtMeContext *ctx, const SvtMeZzJob *jobs, int32_t n_jobs);

/* ---- side product of the ME process: open-loop intra search ----------------------------------------------------------
 * Replaces the OIS loop of MotionEstimationKernel (EbMotionEstimationProcess.c:559-583): OpenLoopIntraSearchLcu
 * (EbMotionEstimation.c:8047-8365) for every SB of a resident picture. Per SB and raster-scan CU index 1..84 (4 x 32x32,
 * 16 x 16x16, 64 x 8x8; the index equals the ME pu_index of the same square) the reference leaves
 * sorted_ois_candidate[cu][0..17] and total_intra_luma_mode[cu]; `cand` receives the byte images of those OisCandidate_t
 * (bit 0-19 distortion, bit 20 valid_distortion, bit 24-31 intra_mode; EbCodingUnit.h:303-313) and `total` the counts.
 * Entries the reference does not write for a CU (it leaves the previous picture's values there) are zero; index 0 of every
 * SB (the 64x64 CU, never searched) is zero. me_sad: the best ME distortion per square PU of the same picture,
 * me_results[sb][pu].distortionDirection[0].distortion, [num_sb][85]; read for inter pictures above temporal layer 0 (and
 * for every inter picture at 4K), may be NULL otherwise. Synchronous. */
typedef struct SvtMeOisParams {
    int32_t slice_type;                 /* 0 = I, 1 = P, 2 = B */
    int32_t temporal_layer_index, is_used_as_reference;
    int32_t cu8x8_mode;                 /* pcs->cu8x8_mode: 1 = no 8x8 CUs on inter pictures */
    int32_t input_resolution;           /* scs->input_resolution, 3 = INPUT_SIZE_4K_RANGE */
    int32_t limit_ois_to_dc;            /* pcs->limit_ois_to_dc_mode_flag */
    int32_t enc_mode;                   /* pcs->enc_mode (one quirk of the reference reads it) */
} SvtMeOisParams;
#define SVTME_OIS_CUS        85
#define SVTME_OIS_CANDIDATES 18
typedef struct SvtMeOisJob {
    int32_t         slot;
    SvtMeOisParams  params;
    const uint32_t *me_sad;             /* [num_sb][85] or NULL */
    uint32_t       *cand;               /* [num_sb][85][18] */
    uint8_t        *total;              /* [num_sb][85] */
} SvtMeOisJob;
int32_t svtme_ois(SvtMeContext *ctx, const SvtMeOisJob *jobs, int32_t n_jobs);

/* Introspection for benchmarks: the CUDA stream (cudaStream_t as void*) and the number of kernel
 * launches issued so far. */
void   *svtme_stream(SvtMeContext *ctx);
int64_t svtme_kernel_launches(const SvtMeContext *ctx);
/* Per-kernel device time: when enabled every launch is bracketed by CUDA events on the context stream.
 * svtme_kernel_times synchronises, then returns accumulated milliseconds and launch counts per kernel
 * kind (arrays of SVTME_KERNEL_KINDS; either may be NULL) and optionally resets the accumulators. */
int32_t svtme_set_profiling(SvtMeContext *ctx, int32_t enable);
int32_t svtme_kernel_times(SvtMeContext *ctx, double *ms, int64_t *launches, int32_t reset);

#ifdef __cplusplus
}
#endif
#endif /* SVTME_H */
