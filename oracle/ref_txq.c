/* oracle/ref_txq.c -- TEST INFRASTRUCTURE ONLY.
 * Thin accessors around the UNMODIFIED reference objects for the transform / quantisation family
 * (compiled with the reference's headers where they lie under /root/reference, see oracle/Makefile):
 * Av1EstimateTransform, av1_quantize_b_facade_II, av1_scan_orders, av1_build_quantizer. The function-pointer
 * set-up mirrors setup_rtcd_internal for a CPU with AVX2 + SSE4.1 (aom_dsp_rtcd.h:2086-2096, 2660-2727), the only
 * configuration the shipped encoder accepts. */
#include "EbDefinitions.h"
#include "aom_dsp_rtcd.h"
#include "EbTransforms.h"
#include "EbPictureControlSet.h"
#include <string.h>
#include <stdlib.h>

extern void av1_quantize_b_facade_II(const tran_low_t *coeff_ptr, int32_t stride, int32_t width, int32_t height,
                                     intptr_t n_coeffs, const MacroblockPlane *p, tran_low_t *qcoeff_ptr,
                                     tran_low_t *dqcoeff_ptr, uint16_t *eob_ptr, const SCAN_ORDER *sc, const QUANT_PARAM *qparam);
extern void av1_build_quantizer(aom_bit_depth_t bit_depth, int32_t y_dc_delta_q, int32_t u_dc_delta_q, int32_t u_ac_delta_q,
                                int32_t v_dc_delta_q, int32_t v_ac_delta_q, Quants *const quants, Dequants *const deq);

void ref_txq_init(void) {
    aom_quantize_b = aom_highbd_quantize_b_avx2;
    aom_quantize_b_32x32 = aom_highbd_quantize_b_32x32_avx2;
#if INTRINSIC_OPT_2
    aom_quantize_b_64x64 = aom_highbd_quantize_b_64x64_avx2;
    av1_fwd_txfm2d_16x8 = av1_fwd_txfm2d_16x8_avx2;   av1_fwd_txfm2d_8x16 = av1_fwd_txfm2d_8x16_avx2;
    av1_fwd_txfm2d_16x4 = av1_fwd_txfm2d_16x4_sse4_1; av1_fwd_txfm2d_4x16 = av1_fwd_txfm2d_4x16_sse4_1;
    av1_fwd_txfm2d_8x4 = av1_fwd_txfm2d_8x4_sse4_1;   av1_fwd_txfm2d_4x8 = av1_fwd_txfm2d_4x8_sse4_1;
    av1_fwd_txfm2d_16x16 = av1_fwd_txfm2d_16x16_avx2;
#else
    aom_quantize_b_64x64 = aom_quantize_b_64x64_c_II;
    av1_fwd_txfm2d_16x8 = av1_fwd_txfm2d_16x8_c;   av1_fwd_txfm2d_8x16 = av1_fwd_txfm2d_8x16_c;
    av1_fwd_txfm2d_16x4 = av1_fwd_txfm2d_16x4_c;   av1_fwd_txfm2d_4x16 = av1_fwd_txfm2d_4x16_c;
    av1_fwd_txfm2d_8x4 = av1_fwd_txfm2d_8x4_c;     av1_fwd_txfm2d_4x8 = av1_fwd_txfm2d_4x8_c;
    av1_fwd_txfm2d_16x16 = Av1TransformTwoD_16x16_c;
#endif
    av1_fwd_txfm2d_32x16 = av1_fwd_txfm2d_32x16_avx2; av1_fwd_txfm2d_32x8 = av1_fwd_txfm2d_32x8_avx2;
    av1_fwd_txfm2d_8x32 = av1_fwd_txfm2d_8x32_avx2;   av1_fwd_txfm2d_16x32 = av1_fwd_txfm2d_16x32_avx2;
    av1_fwd_txfm2d_32x64 = av1_fwd_txfm2d_32x64_avx2; av1_fwd_txfm2d_64x32 = av1_fwd_txfm2d_64x32_avx2;
    av1_fwd_txfm2d_16x64 = av1_fwd_txfm2d_16x64_avx2; av1_fwd_txfm2d_64x16 = av1_fwd_txfm2d_64x16_avx2;
    av1_fwd_txfm2d_64x64 = av1_fwd_txfm2d_64x64_avx2; av1_fwd_txfm2d_32x32 = av1_fwd_txfm2d_32x32_avx2;
    av1_fwd_txfm2d_8x8 = av1_fwd_txfm2d_8x8_avx2;     av1_fwd_txfm2d_4x4 = av1_fwd_txfm2d_4x4_sse4_1;
}

static int tx_size_of(int w, int h) {
    for (int t = 0; t < TX_SIZES_ALL; t++) if (tx_size_wide[t] == w && tx_size_high[t] == h) return t;
    return -1;
}

/* Av1EstimateTransform on one block; coeff must hold w*h int32. Returns the TxSize used. */
int ref_fwd_txfm(int16_t *residual, uint32_t stride, int32_t *coeff, int w, int h, int tx_type, int bit_increment,
                 uint64_t *three_quad_energy) {
    const int ts = tx_size_of(w, h);
    *three_quad_energy = 0;
    Av1EstimateTransform(residual, stride, coeff, (uint32_t)w, (TxSize)ts, three_quad_energy, NULL, (uint32_t)bit_increment,
                         (TxType)tx_type, ASM_AVX2, PLANE_TYPE_Y, DEFAULT_SHAPE);
    return ts;
}

/* q10 = { zbin_dc, zbin_ac, round_dc, round_ac, quant_dc, quant_ac, shift_dc, shift_ac, dequant_dc, dequant_ac } */
void ref_quantize(const int32_t *coeff, int w, int h, int tx_type, const int16_t *q10, int32_t *qcoeff, int32_t *dqcoeff,
                  uint16_t *eob) {
    const int ts = tx_size_of(w, h);
    int16_t lanes[5][8];
    for (int k = 0; k < 5; k++) for (int i = 0; i < 8; i++) lanes[k][i] = q10[2 * k + (i != 0)];
    MacroblockPlane p; memset(&p, 0, sizeof(p));
    p.zbin_QTX = lanes[0]; p.round_QTX = lanes[1]; p.quant_QTX = lanes[2]; p.quant_shift_QTX = lanes[3]; p.dequant_QTX = lanes[4];
    QUANT_PARAM qp; memset(&qp, 0, sizeof(qp));
    qp.log_scale = av1_get_tx_scale((TxSize)ts); qp.tx_size = (TxSize)ts; qp.qmatrix = NULL; qp.iqmatrix = NULL;
    const int pw = w < 32 ? w : 32, ph = h < 32 ? h : 32;
    av1_quantize_b_facade_II(coeff, pw, pw, ph, av1_get_max_eob((TxSize)ts), &p, qcoeff, dqcoeff, eob,
                             &av1_scan_orders[ts][tx_type], &qp);
}

const int16_t *ref_scan(int w, int h, int tx_type) { return av1_scan_orders[tx_size_of(w, h)][tx_type].scan; }
const int16_t *ref_iscan(int w, int h, int tx_type) { return av1_scan_orders[tx_size_of(w, h)][tx_type].iscan; }
int ref_max_eob(int w, int h) { return av1_get_max_eob((TxSize)tx_size_of(w, h)); }

/* luma quantiser of one qindex as built by av1_build_quantizer (EbModeDecisionConfigurationProcess.c:417-520) */
void ref_quant_params(int bit_depth, int qindex, int16_t *q10) {
    static Quants *q[2]; static Dequants *d[2];
    const int b = bit_depth == 10;
    if (!q[b]) {
        q[b] = (Quants *)calloc(1, sizeof(Quants)); d[b] = (Dequants *)calloc(1, sizeof(Dequants));
        av1_build_quantizer(b ? AOM_BITS_10 : AOM_BITS_8, 0, 0, 0, 0, 0, q[b], d[b]);
    }
    for (int i = 0; i < 2; i++) {
        q10[0 + i] = q[b]->y_zbin[qindex][i]; q10[2 + i] = q[b]->y_round[qindex][i]; q10[4 + i] = q[b]->y_quant[qindex][i];
        q10[6 + i] = q[b]->y_quant_shift[qindex][i]; q10[8 + i] = d[b]->y_dequant_QTX[qindex][i];
    }
}

/* Whole residual plane through the reference's own functions, block by block (the CPU arm of bench.py's
 * transform/quantisation sweep and a second pin for meo_txq_plane). Buffers must be 32-byte aligned and the stride a
 * multiple of 16 (AVX2 aligned loads). Output layout as in include/svttxq.h. Returns the number of blocks. */
int ref_txq_plane(int16_t *res, int stride, int width, int height, int w, int h, int tx_type, const int16_t *q10,
                  int32_t *coeff, int32_t *qcoeff, int32_t *dqcoeff, uint16_t *eob, uint64_t *energy) {
    const int pw = w < 32 ? w : 32, ph = h < 32 ? h : 32, n = pw * ph;
    static __attribute__((aligned(64))) int32_t tmp[64 * 64];
    int b = 0;
    for (int by = 0; by < height / h; by++)
        for (int bx = 0; bx < width / w; bx++, b++) {
            uint64_t e = 0;
            ref_fwd_txfm(res + (size_t)by * h * stride + bx * w, (uint32_t)stride, tmp, w, h, tx_type, 0, &e);
            memcpy(coeff + (size_t)b * n, tmp, sizeof(int32_t) * n);
            ref_quantize(coeff + (size_t)b * n, w, h, tx_type, q10, qcoeff + (size_t)b * n, dqcoeff + (size_t)b * n, &eob[b]);
            energy[b] = e;
        }
    return b;
}

/* ---- transform / quantisation dump of real encodes (BASELINE configs[0], "txfm/quant coeffs as golden") ---------------
 * libsvtref.so is linked with -Wl,--wrap=Av1EstimateTransform,--wrap=Av1QuantizeInvQuantize (oracle/Makefile), so the
 * calls the encoder's EncDec stage makes (EbCodingLoop.c:661-700 and its chroma twins) arrive here first. The real
 * functions run unchanged; when a dump is open, a forward transform followed on the same thread by the quantisation of
 * the buffer it wrote is recorded as one block: residual in, quantiser lanes, coefficients / three-quad energy /
 * levels / dequantised levels / eob out. Records are thinned per (size, type, plane) so the fixture stays small.
 * Layout: int32 header[16] then the arrays, see tests/golden/make_encoder_golden.py. */
#include <pthread.h>
#include <stdio.h>
EbErrorType __real_Av1EstimateTransform(int16_t *residual_buffer, uint32_t residualStride, int32_t *coeffBuffer, uint32_t coeffStride,
                                        TxSize transform_size, uint64_t *three_quad_energy, int16_t *transform_inner_array_ptr,
                                        uint32_t bitIncrement, TxType transform_type, EbAsm asm_type, PLANE_TYPE componentType,
                                        EB_TRANS_COEFF_SHAPE transCoeffShape);
void __real_Av1QuantizeInvQuantize(PictureControlSet_t *pcs, int32_t *coeff, const uint32_t coeffStride, int32_t *quantCoeff,
                                   int32_t *reconCoeff, uint32_t qp, uint32_t width, uint32_t height, TxSize txsize, uint16_t *eob,
                                   MacroblockPlane candidate_plane, EbAsm asm_type, uint32_t *y_count_non_zero_coeffs, EbPfMode pfMode,
                                   uint8_t enableContouringQCUpdateFlag, uint32_t componentType, uint32_t bitIncrement, TxType tx_type,
                                   EbBool cleanSparseCoeffFlag);

/* Hook (oracle/enc_driver.c --cuda-txq): when installed, 8-bit blocks are NOT computed by the reference's functions; the hook
 * gets the residual block and the quantiser (q10 NULL = transform only wanted) and fills coefficients, three-quad energy,
 * levels, dequantised levels and eob. Used by the encoder-in-the-loop test of the secondary kernel family. */
typedef int (*ref_txq_hook_fn)(void *user, const int16_t *residual, int w, int h, int tx_type, const int16_t *q10, int32_t *coeff,
                               int32_t *qcoeff, int32_t *dqcoeff, uint16_t *eob, uint64_t *energy);
static ref_txq_hook_fn g_txq_hook; static void *g_txq_hook_user;
static long g_txq_hook_calls[2];
void ref_txq_hook_set(ref_txq_hook_fn fn, void *user) { g_txq_hook = fn; g_txq_hook_user = user; }
void ref_txq_hook_stats(long *transforms, long *quants) { *transforms = g_txq_hook_calls[0]; *quants = g_txq_hook_calls[1]; }
static void luma_or_chroma_q10(const PictureControlSet_t *pcs, uint32_t componentType, uint32_t bitIncrement, int16_t *q10);

static FILE *g_txq_dump;
static pthread_mutex_t g_txq_mu = PTHREAD_MUTEX_INITIALIZER;
static int g_txq_every = 29, g_txq_cap = 24;
static uint32_t g_txq_seen[TX_SIZES_ALL][TX_TYPES][3], g_txq_kept[TX_SIZES_ALL][TX_TYPES][3];
/* the last two transforms of this thread: EncDec transforms Cb and Cr back to back, then quantises Cb and Cr */
static __thread struct TLast { int32_t *coeff; int ts, type, bit_inc; uint64_t energy; int16_t residual[64 * 64]; } t_ring[2];
static __thread int t_next;
#define t_last t_ring[t_slot]

int ref_txq_dump_open(const char *path, int every, int cap) {
    g_txq_dump = fopen(path, "wb");
    if (every > 0) g_txq_every = every;
    if (cap > 0) g_txq_cap = cap;
    memset(g_txq_seen, 0, sizeof(g_txq_seen)); memset(g_txq_kept, 0, sizeof(g_txq_kept));
    return g_txq_dump ? 0 : -1;
}
void ref_txq_dump_close(void) { if (g_txq_dump) fclose(g_txq_dump); g_txq_dump = NULL; }

EbErrorType __wrap_Av1EstimateTransform(int16_t *residual_buffer, uint32_t residualStride, int32_t *coeffBuffer, uint32_t coeffStride,
                                        TxSize transform_size, uint64_t *three_quad_energy, int16_t *transform_inner_array_ptr,
                                        uint32_t bitIncrement, TxType transform_type, EbAsm asm_type, PLANE_TYPE componentType,
                                        EB_TRANS_COEFF_SHAPE transCoeffShape) {
    const int t_slot = t_next;
    if (g_txq_hook && bitIncrement == 0) {
        const int w = tx_size_wide[transform_size], h = tx_size_high[transform_size], n = av1_get_max_eob(transform_size);
        for (int y = 0; y < h; y++) memcpy(t_last.residual + y * w, residual_buffer + (size_t)y * residualStride, sizeof(int16_t) * w);
        int32_t q[1024], dq[1024]; uint16_t eob = 0; uint64_t energy = 0;
        if (g_txq_hook(g_txq_hook_user, t_last.residual, w, h, (int)transform_type, NULL, coeffBuffer, q, dq, &eob, &energy) != 0) {
            fprintf(stderr, "ref_txq: the transform hook failed (%dx%d type %d)\n", w, h, (int)transform_type);
            abort();
        }
        (void)n;
        if (w == 64 || h == 64) *three_quad_energy = energy;
        t_last.coeff = coeffBuffer; t_last.ts = transform_size; t_last.type = transform_type; t_last.bit_inc = 0;
        t_last.energy = (w == 64 || h == 64) ? energy : 0;
        t_next ^= 1;
        __sync_fetch_and_add(&g_txq_hook_calls[0], 1);
        return EB_ErrorNone;
    }
    if (g_txq_dump) {                                     /* the input, before the call (in place is not used, but be safe) */
        const int w = tx_size_wide[transform_size], h = tx_size_high[transform_size];
        for (int y = 0; y < h; y++) memcpy(t_last.residual + y * w, residual_buffer + (size_t)y * residualStride, sizeof(int16_t) * w);
        if (w == 64 || h == 64) *three_quad_energy = 0;
    }
    const EbErrorType rc = __real_Av1EstimateTransform(residual_buffer, residualStride, coeffBuffer, coeffStride, transform_size,
                                                       three_quad_energy, transform_inner_array_ptr, bitIncrement, transform_type,
                                                       asm_type, componentType, transCoeffShape);
    if (g_txq_dump) {
        const int w = tx_size_wide[transform_size], h = tx_size_high[transform_size];
        t_last.coeff = coeffBuffer; t_last.ts = transform_size; t_last.type = transform_type; t_last.bit_inc = (int)bitIncrement;
        t_last.energy = (w == 64 || h == 64) ? *three_quad_energy : 0;
        t_next ^= 1;
    }
    return rc;
}

void __wrap_Av1QuantizeInvQuantize(PictureControlSet_t *pcs, int32_t *coeff, const uint32_t coeffStride, int32_t *quantCoeff,
                                   int32_t *reconCoeff, uint32_t qp, uint32_t width, uint32_t height, TxSize txsize, uint16_t *eob,
                                   MacroblockPlane candidate_plane, EbAsm asm_type, uint32_t *y_count_non_zero_coeffs, EbPfMode pfMode,
                                   uint8_t enableContouringQCUpdateFlag, uint32_t componentType, uint32_t bitIncrement, TxType tx_type,
                                   EbBool cleanSparseCoeffFlag) {
    int t_slot = -1;
    for (int k = 0; k < 2 && (g_txq_dump || g_txq_hook) && componentType < 3; k++)
        if (t_ring[k].coeff == coeff && t_ring[k].ts == (int)txsize && t_ring[k].type == (int)tx_type && t_ring[k].bit_inc == (int)bitIncrement)
            t_slot = k;
    const int paired = t_slot >= 0;
    const int n = av1_get_max_eob(txsize);
    if (g_txq_hook && paired && bitIncrement == 0) {
        int16_t q10[10];
        luma_or_chroma_q10(pcs, componentType, bitIncrement, q10);
        int32_t c[1024]; uint64_t energy = 0;
        for (uint32_t i = 0; i < height; i++) {           /* what Av1QuantizeInvQuantize clears first, EbFullLoop.c:903-907 */
            memset(quantCoeff + i * width, 0, width * sizeof(int32_t));
            memset(reconCoeff + i * width, 0, width * sizeof(int32_t));
        }
        if (g_txq_hook(g_txq_hook_user, t_last.residual, tx_size_wide[txsize], tx_size_high[txsize], (int)tx_type, q10, c, quantCoeff,
                       reconCoeff, eob, &energy) != 0 || memcmp(c, coeff, sizeof(int32_t) * n) != 0) {
            fprintf(stderr, "ref_txq: the quantisation hook failed\n");
            abort();
        }
        *y_count_non_zero_coeffs = *eob;
        t_last.coeff = NULL;
        __sync_fetch_and_add(&g_txq_hook_calls[1], 1);
        return;
    }
    int32_t *coeff_in = NULL;
    if (paired) { coeff_in = (int32_t *)malloc(sizeof(int32_t) * n); memcpy(coeff_in, coeff, sizeof(int32_t) * n); }
    __real_Av1QuantizeInvQuantize(pcs, coeff, coeffStride, quantCoeff, reconCoeff, qp, width, height, txsize, eob, candidate_plane,
                                  asm_type, y_count_non_zero_coeffs, pfMode, enableContouringQCUpdateFlag, componentType, bitIncrement,
                                  tx_type, cleanSparseCoeffFlag);
    if (!paired) return;
    t_last.coeff = NULL;
    pthread_mutex_lock(&g_txq_mu);
    const uint32_t seen = g_txq_seen[txsize][tx_type][componentType]++;
    if (g_txq_dump && seen % (uint32_t)g_txq_every == 0 && g_txq_kept[txsize][tx_type][componentType] < (uint32_t)g_txq_cap) {
        g_txq_kept[txsize][tx_type][componentType]++;
        const PictureParentControlSet_t *pp = pcs->parent_pcs_ptr;
        const uint32_t qi = pp->base_qindex;
        int16_t q10[10];
        luma_or_chroma_q10(pcs, componentType, bitIncrement, q10);
        const int w = tx_size_wide[txsize], h = tx_size_high[txsize];
        int32_t hd[16]; memset(hd, 0, sizeof(hd));
        hd[0] = 0x54584647;                               /* 'TXFG' */
        hd[1] = txsize; hd[2] = w; hd[3] = h; hd[4] = tx_type; hd[5] = (int32_t)componentType; hd[6] = (int32_t)bitIncrement;
        hd[7] = n; hd[8] = (int32_t)qi; hd[9] = (int32_t)pp->picture_number; hd[10] = *eob;
        hd[11] = (int32_t)(t_last.energy & 0xffffffffu); hd[12] = (int32_t)(t_last.energy >> 32);
        fwrite(hd, sizeof(hd), 1, g_txq_dump);
        fwrite(q10, sizeof(q10), 1, g_txq_dump);
        fwrite(t_last.residual, sizeof(int16_t), (size_t)w * h, g_txq_dump);
        fwrite(coeff_in, sizeof(int32_t), n, g_txq_dump);
        fwrite(quantCoeff, sizeof(int32_t), n, g_txq_dump);
        fwrite(reconCoeff, sizeof(int32_t), n, g_txq_dump);
    }
    pthread_mutex_unlock(&g_txq_mu);
    free(coeff_in);
}

/* the quantiser Av1QuantizeInvQuantize_II picks: EbFullLoop.c:721-800 (ADD_DELTA_QP_SUPPORT is 0) */
static void luma_or_chroma_q10(const PictureControlSet_t *pcs, uint32_t componentType, uint32_t bitIncrement, int16_t *q10) {
    const PictureParentControlSet_t *pp = pcs->parent_pcs_ptr;
    const uint32_t qi = pp->base_qindex;
    const Quants *q = bitIncrement == 0 ? &pp->quantsMd : &pp->quants;
    const Dequants *d = bitIncrement == 0 ? &pp->deqMd : &pp->deq;
    for (int i = 0; i < 2; i++) {
        if (componentType == COMPONENT_LUMA) {
            q10[0 + i] = q->y_zbin[qi][i]; q10[2 + i] = q->y_round[qi][i]; q10[4 + i] = q->y_quant[qi][i];
            q10[6 + i] = q->y_quant_shift[qi][i]; q10[8 + i] = d->y_dequant_QTX[qi][i];
        } else if (componentType == COMPONENT_CHROMA_CB) {
            q10[0 + i] = q->u_zbin[qi][i]; q10[2 + i] = q->u_round[qi][i]; q10[4 + i] = q->u_quant[qi][i];
            q10[6 + i] = q->u_quant_shift[qi][i]; q10[8 + i] = d->u_dequant_QTX[qi][i];
        } else {
            q10[0 + i] = q->v_zbin[qi][i]; q10[2 + i] = q->v_round[qi][i]; q10[4 + i] = q->v_quant[qi][i];
            q10[6 + i] = q->v_quant_shift[qi][i]; q10[8 + i] = d->v_dequant_QTX[qi][i];
        }
    }
}
