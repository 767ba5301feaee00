// me_common.cuh -- device-side structures and helpers shared by the ME kernels (sm_100a).
//
// Data layout in HBM:
//   picture slot  = three padded 8-bit planes (full pad 68, quarter pad 32, sixteenth pad 16),
//                   row pitch rounded up to 64 bytes, plane base 256-byte aligned
//   SbScratch     = per (job, SB) intermediate state handed from kernel to kernel
//   SvtMePuResult = per (job, SB, PU) output, byte image of the reference's MeCuResults_t
#pragma once
#include <cstdint>
#include <cuda_runtime.h>
#include "../../include/svtme.h"
#include "ois_core.cuh"

namespace svtme {

constexpr int kPadFull = 68, kPadQuarter = 32, kPadSixteenth = 16;
constexpr int kNumPu = SVTME_SQUARE_PU_COUNT;
constexpr int kMaxPu = SVTME_MAX_PU_COUNT;              // 209-PU mode (PIC_ALL_DEPTH_MODE pictures)
constexpr int kNumNsq = kMaxPu - kNumPu;                // the 124 non-square PUs
constexpr int kNumLevels = 4, kMaxLevels = 14;          // PU classes ("levels"): 4 square + 10 non-square
constexpr uint32_t kMaxSadValue = 128u * 128u * 255u;   // EbMotionEstimation.h:78

struct DPlane {
    uint8_t* buf;            // first byte of the padded allocation
    int stride;              // bytes per row (multiple of 64)
    int ox, oy;              // origin of pixel (0,0) inside the allocation
    int w, h;                // picture size at this level
    __device__ __forceinline__ const uint8_t* at(int x, int y) const {
        return buf + (size_t)(oy + y) * stride + ox + x;
    }
};

struct DPyramid { DPlane full, quarter, sixteenth; };

// State handed between the per-list kernels of one SB.
struct SbScratch {
    int32_t  sa_x[2], sa_y[2], sa_w[2], sa_h[2];     // clipped search area per list
    int32_t  center_x[2], center_y[2];
    int32_t  hme_x[3][2][2], hme_y[3][2][2];         // per-region HME centres of list 0, inherited by list 1 (quirk Q7)
    uint32_t best_sad[2][kNumPu];                    // internal (z-order) PU index, like p_sb_best_sad
    uint32_t best_mv[2][kNumPu];
    uint8_t  pred[4][64 * 64];                       // final list-0 uni-prediction blocks per PU size level (read by the list-1 sub-pel pass: bi-prediction)
};

// Extra state of the 209-PU mode: the non-square PUs (internal index 85..208) and their uni-prediction blocks.
struct SbScratchExt {
    uint32_t best_sad[2][kNumNsq];
    uint32_t best_mv[2][kNumNsq];
    uint8_t  pred[kMaxLevels - kNumLevels][64 * 64];         // list 0 only, as in SbScratch
};

struct DJob {
    DPyramid cur, ref[2];
    SvtMeParams prm;
    SbScratch* scratch;          // [n_sb]
    SbScratchExt* scratch_ext;   // [n_sb], only for jobs with prm.num_pus == 209
    SvtMePuResult* results;      // [n_sb][prm.num_pus]  (device)
    SvtMeSbDebug* debug;         // [n_sb]      (device)
    int sb_cols, sb_rows;
    int sb_begin, sb_end;        // SBs [sb_begin, sb_end) of the raster are searched (an SB-row range)
    const void* ref_tmap[2];     // TMA descriptors (CUtensorMap in global memory) of ref[l].full, box kTmaBoxBytes x kTmaBoxRows; null = cp.async staging
};
constexpr int kTmaBoxBytes = 160, kTmaBoxRows = 126;   // the raw full-pel window: ten 16-byte granules x (64 search rows + 62)

// ---- TMA staging of the raw full-pel window (fullpel.cu, fullpel_ext.cu): ONE 2-D tensor box (kTmaBoxBytes x kTmaBoxRows from
// the 16-byte aligned address below the window's first byte) issued by one thread; completion is counted in bytes on an
// mbarrier. SVTME_TMA=0 (or a driver without cuTensorMapEncodeTiled) leaves DJob.ref_tmap null: 16-byte cp.async granules.
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(bar), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t done = 0, spins = 0;
    while (!done) {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done) : "r"(bar), "r"(parity) : "memory");
        if (!done && ++spins > (1u << 26)) __trap();          // a copy that never completes is an error, not a hang
    }
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const void* tmap, int x, int y, uint32_t bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 :: "r"(dst), "l"(tmap), "r"(bar), "r"(x), "r"(y) : "memory");
}

// raster index -> internal index (EbMotionEstimation.h:89-128)
__device__ __constant__ const uint8_t kTab16[16] = { 0, 1, 4, 5, 2, 3, 6, 7, 8, 9, 12, 13, 10, 11, 14, 15 };
__device__ __constant__ const uint8_t kTab8[64] = {
    0, 1, 4, 5, 16, 17, 20, 21, 2, 3, 6, 7, 18, 19, 22, 23, 8, 9, 12, 13, 24, 25, 28, 29, 10, 11, 14, 15, 26, 27, 30, 31,
    32, 33, 36, 37, 48, 49, 52, 53, 34, 35, 38, 39, 50, 51, 54, 55, 40, 41, 44, 45, 56, 57, 60, 61, 42, 43, 46, 47, 58, 59, 62, 63 };

// PU classes of the 209-PU mode, in pu_index order (EbMotionEstimation.h:176-262, EbMotionEstimationContext.h:45-261):
// first pu_index, width and height in 8x8 blocks. Inside a class pu_index runs in raster order.
__device__ __constant__ const uint8_t kLevelBase[kMaxLevels] = { 0, 1, 5, 21, 85, 87, 95, 127, 129, 137, 169, 185, 201, 205 };
__device__ __constant__ const uint8_t kLevelWb[kMaxLevels]   = { 8, 4, 2, 1, 8, 4, 2, 4, 2, 1, 4, 1, 8, 2 };
__device__ __constant__ const uint8_t kLevelHb[kMaxLevels]   = { 8, 4, 2, 1, 4, 2, 1, 8, 4, 2, 1, 4, 2, 8 };
// pu_index -> internal index (p_sb_best_sad order): tab16x16 / tab8x8 / tab32x16 / tab16x8 / tab16x32 / tab8x16 / tab32x8 /
// tab8x32 of EbMotionEstimation.h:89-175 with the class bases added
__device__ __constant__ const uint8_t kInternal[kMaxPu] = {
    0, 1, 2, 3, 4,
    5, 6, 9, 10, 7, 8, 11, 12, 13, 14, 17, 18, 15, 16, 19, 20,
    21, 22, 25, 26, 37, 38, 41, 42, 23, 24, 27, 28, 39, 40, 43, 44, 29, 30, 33, 34, 45, 46, 49, 50, 31, 32, 35, 36, 47, 48, 51, 52,
    53, 54, 57, 58, 69, 70, 73, 74, 55, 56, 59, 60, 71, 72, 75, 76, 61, 62, 65, 66, 77, 78, 81, 82, 63, 64, 67, 68, 79, 80, 83, 84,
    85, 86,
    87, 89, 88, 90, 91, 93, 92, 94,
    95, 97, 103, 105, 96, 98, 104, 106, 99, 101, 107, 109, 100, 102, 108, 110,
    111, 113, 119, 121, 112, 114, 120, 122, 115, 117, 123, 125, 116, 118, 124, 126,
    127, 128,
    129, 130, 131, 132, 133, 134, 135, 136,
    137, 138, 139, 140, 145, 146, 147, 148, 141, 142, 143, 144, 149, 150, 151, 152,
    153, 154, 155, 156, 161, 162, 163, 164, 157, 158, 159, 160, 165, 166, 167, 168,
    169, 173, 170, 174, 171, 175, 172, 176, 177, 181, 178, 182, 179, 183, 180, 184,
    185, 186, 187, 188, 189, 190, 191, 192, 193, 194, 195, 196, 197, 198, 199, 200,
    201, 202, 203, 204, 205, 206, 207, 208 };
// internal index of the PU of class L that contains 8x8 block (bx, by)
__device__ __forceinline__ int pu_of_block_any(int L, int bx, int by) {
    const int wb = kLevelWb[L], hb = kLevelHb[L];
    return kInternal[kLevelBase[L] + (by / hb) * (8 / wb) + bx / wb];
}

__device__ __forceinline__ int mv_x(uint32_t mv) { return (int16_t)(mv & 0xFFFFu); }
__device__ __forceinline__ int mv_y(uint32_t mv) { return (int16_t)(mv >> 16); }
__device__ __forceinline__ uint32_t make_mv(int x, int y) { return ((uint32_t)(uint16_t)y << 16) | (uint16_t)x; }

// SAD of four packed bytes accumulated into c: one VABSDIFF4.U8.ACC on sm_100a.
// (inline PTX: nvcc lowers `__vsadu4(a, b) + c` to VABSDIFF4 with a zero addend plus a separate IADD3.)
__device__ __forceinline__ uint32_t sad4(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t d;
    asm("vabsdiff4.u32.u32.u32.add %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

// 32-bit word at an arbitrary byte address of global memory (two aligned loads + funnel shift).
__device__ __forceinline__ uint32_t ldg_u32_unaligned(const uint8_t* p) {
    const uintptr_t a = (uintptr_t)p;
    const uint32_t* w = (const uint32_t*)(a & ~(uintptr_t)3);
    const uint32_t sh = ((uint32_t)a & 3u) * 8u;
    const uint32_t lo = __ldg(w);
    if (sh == 0) return lo;
    return __funnelshift_r(lo, __ldg(w + 1), sh);
}

// The reference's four-step search-area edge correction, statement order preserved
// (quirk: left/top clipping moves the origin but never shrinks the size).  ME.c:6690-6723
__device__ __forceinline__ void clip_search_area(int org, int& sao, int& sa, int pad, int dim) {
    int o = sao, a = sa;
    o = ((org + o) < -pad) ? -pad - org : o;
    a = ((org + o) < -pad) ? a - (-pad - (org + o)) : a;
    o = ((org + o) > dim - 1) ? o - ((org + o) - (dim - 1)) : o;
    a = ((org + o + a) > dim) ? max(1, a - ((org + o + a) - dim)) : a;
    sao = o; sa = a;
}
__device__ __forceinline__ int clip_center(int org, int c, int pad, int dim) {
    c = ((org + c) < -pad) ? -pad - org : c;
    c = ((org + c) > dim - 1) ? c - ((org + c) - (dim - 1)) : c;
    return c;
}

// 64-bit min across a warp / block
__device__ __forceinline__ unsigned long long warp_min_u64(unsigned long long v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        unsigned long long t = __shfl_xor_sync(0xffffffffu, v, o);
        v = t < v ? t : v;
    }
    return v;
}

// Search areas beyond 64 x 64 positions (the reference searches MIN(search_area, 127), ME.c:6644-6645) run in the Wide
// instantiations of the full-pel and sub-pel kernels. Decided on the job's parameters (uniform per picture), not on the
// clipped area of an SB.
__host__ __device__ inline bool wide_search(const SvtMeParams& p) { return p.search_area_w > 64 || p.search_area_h > 64; }

// Pictures of one batched pyramid launch: source luma and slot base pointer per picture
struct PyramidBatch {
    const uint8_t* luma[SVTME_MAX_BATCH];
    uint8_t* base[SVTME_MAX_BATCH];
};

// host-side launchers (one per .cu)
void launch_pyramid_batch(const PyramidBatch& batch, int n, int luma_stride, const DPyramid& geom, const uint8_t* geom_base, cudaStream_t st);
void launch_pyramid(const uint8_t* d_luma, int luma_stride, const DPyramid& pyr, cudaStream_t st);
void launch_unpack16(const uint16_t* d_in, int in_stride, uint8_t* d_msb, uint8_t* d_rem, int width, int height, cudaStream_t st);
void launch_hme(const DJob* d_jobs, int n_jobs, int max_sb, int list, cudaStream_t st);
// The square-mode and ext-mode kernels are separate launches over the same grid; a CTA whose job is of the other
// mode exits at once. `ext` selects the 209-PU variant, `wide` the one for search areas beyond 64 x 64.
void launch_fullpel(const DJob* d_jobs, int n_jobs, int max_sb, int list, bool ext, bool wide, cudaStream_t st);
// `worklist`: room for n_jobs * max_sb 32-bit entries preceded by one 32-bit counter (worklist[-1]); owned by this launch
// sequence until it has run (the small-window sub-pel pass defers the SBs whose window does not fit to a second pass)
void launch_subpel(const DJob* d_jobs, int n_jobs, int max_sb, int list, bool ext, bool wide, uint32_t* worklist, cudaStream_t st);
void launch_bipred_pack(const DJob* d_jobs, int n_jobs, int max_sb, bool ext, cudaStream_t st);
// open-loop intra search (ois.cu): one job = one picture; me_sad [n_sb][85] (device, may be null for intra pictures),
// out_cand [n_sb][85][18] byte images of OisCandidate_t, out_total [n_sb][85]
struct OisJob { DPlane plane; ois::Params prm; const uint32_t* me_sad; uint32_t* out_cand; uint8_t* out_total; };
void launch_ois(const OisJob* d_jobs, int n_jobs, int sb_cols, int n_sb, int variant, cudaStream_t st);   // variant 0: line-per-lane kernel; 1: pixel-wise; 2: word-wise
// zero-motion SADs (zz_sad.cu): one job = one (current, previous) picture pair
struct ZzJob { DPlane cur16, prev16; uint8_t* out; };        // out: [2][n_sb] (zz cost, non-moving index), device
void launch_zz_sad(const ZzJob* d_jobs, int n_jobs, int width, int height, int sb_cols, int n_sb, cudaStream_t st);
cudaError_t configure_kernels();

}  // namespace svtme
