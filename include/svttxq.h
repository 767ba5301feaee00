/* include/svttxq.h -- C-ABI of the secondary kernel family in libsvtme_cuda.so: AV1 forward 2-D transform +
 * quantisation for batches of equally sized transform blocks on B200.
 *
 *   svttxq_fwd_txfm_quant   replaces, for every block of the batch, the pair
 *                           Av1EstimateTransform   Source/Lib/Codec/EbTransforms.c:4410-4728
 *                             (av1_fwd_txfm2d_WxH, zero-out / repack / three_quad_energy of 64-point sizes)
 *                           Av1QuantizeInvQuantize Source/Lib/Codec/EbFullLoop.c:671-875
 *                             -> av1_quantize_b_facade_II :596 -> aom_quantize_b / _32x32 / _64x64
 *                             (= aom_highbd_quantize_b*_avx2, aom_dsp_rtcd.h:2086-2096, 2664-2668)
 *                           as called per transform block from Av1EncodeLoop (EbCodingLoop.c:661-696) and
 *                           ProductFullLoop (EbFullLoop.c:976-990).
 *   SvtTxqQuant             the first two lanes (DC, AC) of MacroblockPlane::zbin_QTX / round_QTX / quant_QTX /
 *                           quant_shift_QTX / dequant_QTX, i.e. Quants::y_*[qindex] built by av1_build_quantizer
 *                           (EbModeDecisionConfigurationProcess.c:417); the host encoder keeps owning those tables.
 *
 * Blocks are the txw x txh tiles of a residual plane in raster order (block b covers columns
 * (b % blocks_x) * txw .., rows (b / blocks_x) * txh ..). Outputs per block: min(32,txw) * min(32,txh)
 * coefficients in the reference's packed row-major layout (64-point sizes keep the top-left 32 columns/rows).
 * Error codes are those of svtme.h. There is no CPU fallback.
 */
#ifndef SVTTXQ_H
#define SVTTXQ_H
#include <stdint.h>
#include <stddef.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct SvtTxqContext SvtTxqContext;

typedef struct SvtTxqQuant {           /* index 0 = DC, 1 = AC */
    int16_t zbin[2], round[2], quant[2], quant_shift[2], dequant[2];
} SvtTxqQuant;

int32_t svttxq_create(SvtTxqContext **ctx, int32_t device);
void    svttxq_destroy(SvtTxqContext *ctx);
const char *svttxq_last_error(const SvtTxqContext *ctx);

/* 1 when the reference has 1-D kernels for (txw, txh, tx_type): ADST/FLIPADST up to 16 points, 64-point DCT only. */
int32_t svttxq_supported(int32_t txw, int32_t txh, int32_t tx_type);

/* Host buffers. residual: int16 plane, `stride` elements per row, width/height multiples of txw/txh.
 * coeff / qcoeff / dqcoeff: [n_blocks][pw*ph] int32, eob: [n_blocks] uint16, energy: [n_blocks] uint64
 * (three_quad_energy; zero unless a dimension is 64). Any output pointer may be NULL. Synchronous. */
int32_t svttxq_fwd_txfm_quant(SvtTxqContext *ctx, const int16_t *residual, int32_t stride, int32_t width, int32_t height,
                              int32_t txw, int32_t txh, int32_t tx_type, const SvtTxqQuant *quant,
                              int32_t *coeff, int32_t *qcoeff, int32_t *dqcoeff, uint16_t *eob, uint64_t *energy);
/* Same with every buffer in device memory; asynchronous on svttxq_stream(). */
int32_t svttxq_fwd_txfm_quant_device(SvtTxqContext *ctx, const int16_t *d_residual, int32_t stride, int32_t width, int32_t height,
                                     int32_t txw, int32_t txh, int32_t tx_type, const SvtTxqQuant *quant,
                                     int32_t *d_coeff, int32_t *d_qcoeff, int32_t *d_dqcoeff, uint16_t *d_eob, uint64_t *d_energy);
/* Eob-bounded return path for host buffers: what the consumer of a quantised block reads is its eob and the `eob` leading
 * levels in scan order (the entropy coder walks the scan backwards from eob - 1; dqcoeff = (level * dequant[i != 0]) >>
 * log_scale with the level's sign is rebuilt where reconstruction needs it: av1_quantize_b_facade_II, EbFullLoop.c:596-662,
 * aom_highbd_quantize_b*_avx2, ASM_AVX2/highbd_quantize_intrin_avx2.c:82-200). So instead of three int32 planes this call
 * returns eob[n_blocks], level_offset[n_blocks + 1] (exclusive prefix sum of eob; may be NULL: the caller can sum eob itself)
 * and levels[level_offset[b] + i] = qcoeff of block b at scan position i, i < eob[b], as int16 -- 2 bytes per block + 2 bytes
 * per kept level over the host link. *n_levels receives the total; SVTME_ERR_INSUFFICIENT_RESOURCES when it exceeds
 * levels_cap (n_blocks * pw * ph always suffices), SVTME_ERR_BAD_PARAMETER when a level does not fit 16 bits (only
 * possible above 8-bit video: use svttxq_fwd_txfm_quant then). Synchronous. */
int32_t svttxq_fwd_txfm_quant_compact(SvtTxqContext *ctx, const int16_t *residual, int32_t stride, int32_t width, int32_t height,
                                      int32_t txw, int32_t txh, int32_t tx_type, const SvtTxqQuant *quant,
                                      uint16_t *eob, uint32_t *level_offset, int16_t *levels, size_t levels_cap, size_t *n_levels);
int32_t svttxq_synchronize(SvtTxqContext *ctx);
void   *svttxq_stream(SvtTxqContext *ctx);
int64_t svttxq_kernel_launches(const SvtTxqContext *ctx);

#ifdef __cplusplus
}
#endif
#endif /* SVTTXQ_H */
