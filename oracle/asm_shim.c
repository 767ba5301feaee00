/* oracle/asm_shim.c -- TEST INFRASTRUCTURE ONLY.
 *
 * The reference ships five NASM files (Source/Lib/ASM_SSE2/*.asm); this image has no
 * nasm/yasm, so the handful of symbols they export are restated here in plain C so that
 * the reference's C/intrinsics objects link into oracle/_ref/libsvtref.so.
 *
 * On the ME / transform path only these matter (semantics read from
 * Source/Lib/ASM_SSE2/EbPictureOperators_SSE2.asm):
 *   Log2f_SSE2                 bsr           (asm :629-632)
 *   PictureAverageKernel_SSE2  pavgb loop    (a+b+1)>>1
 *   PictureCopyKernel_SSE2     row copy      (asm :11-200)
 *   ZeroOutCoeff*              zero a w x h int16 area
 *   RunEmms                    emms (no-op for our purposes)
 * Everything else (intra DC/V/H predictors, high-bitdepth variance) is off the hot path
 * and aborts loudly if ever reached.
 */
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

uint32_t Log2f_SSE2(uint32_t x) {
    /* bsr: index of highest set bit (undefined for 0; callers never pass 0) */
    uint32_t r = 0;
    while (x >>= 1) r++;
    return r;
}

void RunEmms(void) {}

void PictureCopyKernel_SSE2(uint8_t *src, uint32_t src_stride, uint8_t *dst, uint32_t dst_stride,
                            uint32_t areaWidth, uint32_t areaHeight) {
    for (uint32_t y = 0; y < areaHeight; y++)
        memcpy(dst + (size_t)y * dst_stride, src + (size_t)y * src_stride, areaWidth);
}

void PictureAverageKernel_SSE2(uint8_t *src0, uint32_t src0Stride, uint8_t *src1, uint32_t src1Stride,
                               uint8_t *dst, uint32_t dst_stride, uint32_t areaWidth, uint32_t areaHeight) {
    for (uint32_t y = 0; y < areaHeight; y++) {
        for (uint32_t x = 0; x < areaWidth; x++)
            dst[x] = (uint8_t)((src0[x] + src1[x] + 1) >> 1);
        src0 += src0Stride; src1 += src1Stride; dst += dst_stride;
    }
}

static void zero_area(int16_t *buf, uint32_t stride, uint32_t origin, uint32_t w, uint32_t h) {
    for (uint32_t y = 0; y < h; y++)
        memset(buf + origin + (size_t)y * stride, 0, w * sizeof(int16_t));
}
void ZeroOutCoeff4x4_SSE(int16_t *b, uint32_t s, uint32_t o, uint32_t w, uint32_t h) { (void)w; (void)h; zero_area(b, s, o, 4, 4); }
void ZeroOutCoeff8x8_SSE2(int16_t *b, uint32_t s, uint32_t o, uint32_t w, uint32_t h) { (void)w; (void)h; zero_area(b, s, o, 8, 8); }
void ZeroOutCoeff16x16_SSE2(int16_t *b, uint32_t s, uint32_t o, uint32_t w, uint32_t h) { (void)w; (void)h; zero_area(b, s, o, 16, 16); }
void ZeroOutCoeff32x32_SSE2(int16_t *b, uint32_t s, uint32_t o, uint32_t w, uint32_t h) { (void)w; (void)h; zero_area(b, s, o, 32, 32); }

#define OFF_PATH(name) \
    void name(void) { fprintf(stderr, "asm_shim: off-path NASM symbol %s reached\n", #name); abort(); }

/* NASM intra predictors: forwarded to the reference's own C versions (bit-identical by construction), so that the whole
 * encoder can run from libsvtref.so (oracle/enc_driver.c dumps the ME stage of real encodes). */
#include <stddef.h>
#define FWD8(name, sz) \
    void name##_##sz##_c(uint8_t *dst, ptrdiff_t stride, const uint8_t *above, const uint8_t *left); \
    void name##_##sz##_sse2(uint8_t *dst, ptrdiff_t stride, const uint8_t *above, const uint8_t *left) { name##_##sz##_c(dst, stride, above, left); }
#define FWD8_ALL(name) FWD8(name, 4x4) FWD8(name, 8x8) FWD8(name, 16x16)
FWD8_ALL(aom_dc_128_predictor) FWD8_ALL(aom_dc_left_predictor) FWD8_ALL(aom_dc_predictor) FWD8_ALL(aom_dc_top_predictor)
FWD8_ALL(aom_h_predictor) FWD8_ALL(aom_v_predictor)
#define FWD16(name, sz) \
    void name##_##sz##_c(uint16_t *dst, ptrdiff_t stride, const uint16_t *above, const uint16_t *left, int32_t bd); \
    void name##_##sz##_sse2(uint16_t *dst, ptrdiff_t stride, const uint16_t *above, const uint16_t *left, int32_t bd) { name##_##sz##_c(dst, stride, above, left, bd); }
FWD16(aom_highbd_dc_predictor, 4x4) FWD16(aom_highbd_dc_predictor, 8x8) FWD16(aom_highbd_v_predictor, 4x4) FWD16(aom_highbd_v_predictor, 8x8)
OFF_PATH(aom_highbd_calc16x16var_sse2) OFF_PATH(aom_highbd_calc8x8var_sse2)
