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

namespace svtme {

constexpr int kPadFull = 68, kPadQuarter = 32, kPadSixteenth = 16;
constexpr int kNumPu = SVTME_SQUARE_PU_COUNT;
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
    uint8_t  pred[2][4][64 * 64];                    // final uni-pred blocks per list and PU size level
};

struct DJob {
    DPyramid cur, ref[2];
    SvtMeParams prm;
    SbScratch* scratch;          // [n_sb]
    SvtMePuResult* results;      // [n_sb][85]  (device)
    SvtMeSbDebug* debug;         // [n_sb]      (device)
    int sb_cols, sb_rows;
    int sb_begin, sb_end;        // SBs [sb_begin, sb_end) of the raster are searched (an SB-row range)
};

// raster index -> internal index (EbMotionEstimation.h:89-128)
__device__ __constant__ const uint8_t kTab16[16] = { 0, 1, 4, 5, 2, 3, 6, 7, 8, 9, 12, 13, 10, 11, 14, 15 };
__device__ __constant__ const uint8_t kTab8[64] = {
    0, 1, 4, 5, 16, 17, 20, 21, 2, 3, 6, 7, 18, 19, 22, 23, 8, 9, 12, 13, 24, 25, 28, 29, 10, 11, 14, 15, 26, 27, 30, 31,
    32, 33, 36, 37, 48, 49, 52, 53, 34, 35, 38, 39, 50, 51, 54, 55, 40, 41, 44, 45, 56, 57, 60, 61, 42, 43, 46, 47, 58, 59, 62, 63 };

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

// host-side launchers (one per .cu)
void launch_pyramid(const uint8_t* d_luma, int luma_stride, const DPyramid& pyr, cudaStream_t st);
void launch_unpack16(const uint16_t* d_in, int in_stride, uint8_t* d_msb, uint8_t* d_rem, int width, int height, cudaStream_t st);
void launch_hme(const DJob* d_jobs, int n_jobs, int max_sb, int list, cudaStream_t st);
void launch_fullpel(const DJob* d_jobs, int n_jobs, int max_sb, int list, cudaStream_t st);
void launch_subpel(const DJob* d_jobs, int n_jobs, int max_sb, int list, cudaStream_t st);
void launch_bipred_pack(const DJob* d_jobs, int n_jobs, int max_sb, cudaStream_t st);
cudaError_t configure_kernels();

}  // namespace svtme
