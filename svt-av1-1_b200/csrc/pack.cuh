// pack.cuh -- the tail of MotionEstimateLcu for one PU (reference Source/Lib/Codec/EbMotionEstimation.c:6973-7146):
// sum of the bi-prediction SAD over the PU's 8x8 blocks, Sort3Elements (:5434-5463) over {L0, L1, Bi}, and the 40-byte
// MeCuResults_t image. Shared by bipred_pack_kernel (pictures with one list) and by the list-1 pass of the sub-pel kernel,
// which computes the bi-prediction SADs while the list-1 prediction blocks are still in registers.
#pragma once
#include "me_common.cuh"

namespace svtme {

// pu = output pu_index (raster inside each size class). part: [levels][64] bi-prediction SADs per 8x8 block (read only when the
// picture has two lists). best_sad(l, n) / best_mv(l, n): per list and INTERNAL index n.
template <bool Ext, typename SadFn, typename MvFn>
__device__ __forceinline__ void pack_pu_result(const DJob& job, int sb, int pu, const uint32_t* part, SadFn best_sad, MvFn best_mv) {
    constexpr int NPU = Ext ? kMaxPu : kNumPu;
    const SvtMeParams& p = job.prm;
    int L, bx, by, nbx, nby, n;
    if constexpr (Ext) {
        L = kMaxLevels - 1;
        while (pu < kLevelBase[L]) L--;
        const int i = pu - kLevelBase[L];
        nbx = kLevelWb[L]; nby = kLevelHb[L];
        bx = (i % (8 / nbx)) * nbx; by = (i / (8 / nbx)) * nby; n = kInternal[pu];
    } else {
        if (pu == 0) { L = 0; bx = 0; by = 0; nbx = 8; n = 0; }
        else if (pu < 5) { const int i = pu - 1; L = 1; bx = (i & 1) * 4; by = (i >> 1) * 4; nbx = 4; n = pu; }
        else if (pu < 21) { const int i = pu - 5; L = 2; bx = (i & 3) * 2; by = (i >> 2) * 2; nbx = 2; n = 5 + kTab16[i]; }
        else { const int i = pu - 21; L = 3; bx = i & 7; by = i >> 3; nbx = 1; n = 21 + kTab8[i]; }
        nby = nbx;
    }
    SvtMePuResult o;
    uint32_t* ow = reinterpret_cast<uint32_t*>(&o);
#pragma unroll
    for (int i = 0; i < 10; i++) ow[i] = 0;
    int total = p.num_lists;
    uint32_t bi = 0;
    if (p.num_lists == 2 && (p.cu8x8_mode == 0 || pu < 21 || Ext)) {
        for (int yy = 0; yy < nby; yy++) for (int xx = 0; xx < nbx; xx++) bi += part[L * 64 + (by + yy) * 8 + bx + xx];
        total = 3;
    }
    const uint32_t a = best_sad(0, n), b = p.num_lists == 2 ? best_sad(1, n) : 0u;
    const uint32_t mv0 = best_mv(0, n);
    o.total_cand = (uint8_t)total;
    o.xMvL0 = (int16_t)mv_x(mv0); o.yMvL0 = (int16_t)mv_y(mv0);
    if (p.num_lists == 2) { const uint32_t mv1 = best_mv(1, n); o.xMvL1 = (int16_t)mv_x(mv1); o.yMvL1 = (int16_t)mv_y(mv1); }
#define SVTME_SET(k, d, dirv) do { o.cand[k].distortion = (d); o.cand[k].direction = (dirv); } while (0)
    if (total == 3) {                                   // Sort3Elements, ME.c:5434-5463
        if (a <= b && a <= bi) {
            if (b <= bi) { SVTME_SET(0, a, 0); SVTME_SET(1, b, 1); SVTME_SET(2, bi, 2); } else { SVTME_SET(0, a, 0); SVTME_SET(1, bi, 2); SVTME_SET(2, b, 1); }
        } else if (b <= a && b <= bi) {
            if (a <= bi) { SVTME_SET(0, b, 1); SVTME_SET(1, a, 0); SVTME_SET(2, bi, 2); } else { SVTME_SET(0, b, 1); SVTME_SET(1, bi, 2); SVTME_SET(2, a, 0); }
        } else if (a <= b) { SVTME_SET(0, bi, 2); SVTME_SET(1, a, 0); SVTME_SET(2, b, 1); }
        else               { SVTME_SET(0, bi, 2); SVTME_SET(1, b, 1); SVTME_SET(2, a, 0); }
    } else if (total == 2) {
        if (a <= b) { SVTME_SET(0, a, 0); SVTME_SET(1, b, 1); } else { SVTME_SET(0, b, 1); SVTME_SET(1, a, 0); }
    } else SVTME_SET(0, a, 0);
#undef SVTME_SET
    uint32_t* dst = reinterpret_cast<uint32_t*>(&job.results[(size_t)sb * NPU + pu]);
#pragma unroll
    for (int i = 0; i < 10; i++) dst[i] = ow[i];
}

}  // namespace svtme
