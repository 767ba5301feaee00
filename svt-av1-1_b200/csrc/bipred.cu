// bipred.cu -- bi-prediction SAD, candidate ordering and result packing for every PU of every SB:
// the tail of MotionEstimateLcu (reference Source/Lib/Codec/EbMotionEstimation.c:6973-7146) with
// BiPredictionSearch / BiPredictionCompensation / BiPredAverging (:4933-5342) and Sort3Elements (:5434).
//
// The two uni-prediction blocks per PU were materialised by subpel_kernel (one 64x64 byte image per
// PU-size level and list), so the bi-pred distortion is SAD(src, (p0 + p1 + 1) >> 1) over 8x8-block
// tasks; one thread per PU then orders {L0, L1, Bi} and writes the 40-byte MeCuResults_t image.
#include "me_common.cuh"

namespace svtme {

constexpr int kBpThreads = 256;

// Ext = the 209-PU mode (14 PU classes; the non-square PUs and their prediction blocks live in SbScratchExt)
template <bool Ext>
__global__ void __launch_bounds__(kBpThreads) bipred_pack_kernel(const DJob* __restrict__ jobs) {
    constexpr int NPU = Ext ? kMaxPu : kNumPu, NLEV = Ext ? kMaxLevels : kNumLevels;
    __shared__ uint32_t part[NLEV][64];
    const DJob& job = jobs[blockIdx.y];
    const int sb = job.sb_begin + blockIdx.x;
    if (sb >= job.sb_end) return;
    const SvtMeParams& p = job.prm;
    if ((p.num_pus == kMaxPu) != Ext) return;
    const int tid = threadIdx.x;
    const int ox = (sb % job.sb_cols) * 64, oy = (sb / job.sb_cols) * 64;
    const SbScratch& scr = job.scratch[sb];

    if (p.num_lists == 2) for (int t = tid; t < NLEV * 64; t += kBpThreads) {
        const int blk = t & 63, L = t >> 6, bx = blk & 7, by = blk >> 3;
        const uint8_t* q0 = scr.pred[0][L < kNumLevels ? L : 0];
        const uint8_t* q1 = scr.pred[1][L < kNumLevels ? L : 0];
        if constexpr (Ext) { if (L >= kNumLevels) { q0 = job.scratch_ext[sb].pred[0][L - kNumLevels]; q1 = job.scratch_ext[sb].pred[1][L - kNumLevels]; } }
        const uint32_t* p0 = reinterpret_cast<const uint32_t*>(q0) + (by * 8) * 16 + bx * 2;
        const uint32_t* p1 = reinterpret_cast<const uint32_t*>(q1) + (by * 8) * 16 + bx * 2;
        uint32_t sad = 0;
#pragma unroll
        for (int r = 0; r < 8; r++) {
            const uint32_t* s = reinterpret_cast<const uint32_t*>(job.cur.full.at(ox + bx * 8, oy + by * 8 + r));
            sad = sad4(s[0], __vavgu4(p0[r * 16], p1[r * 16]), sad);
            sad = sad4(s[1], __vavgu4(p0[r * 16 + 1], p1[r * 16 + 1]), sad);
        }
        part[L][blk] = sad;
    }
    __syncthreads();

    if (tid < NPU) {
        // tid = output pu_index (raster inside each size class); n = internal index
        int L, bx, by, nbx, nby, n;
        if constexpr (Ext) {
            L = kMaxLevels - 1;
            while (tid < kLevelBase[L]) L--;
            const int i = tid - kLevelBase[L];
            nbx = kLevelWb[L]; nby = kLevelHb[L];
            bx = (i % (8 / nbx)) * nbx; by = (i / (8 / nbx)) * nby; n = kInternal[tid];
        } else {
            if (tid == 0) { L = 0; bx = 0; by = 0; nbx = 8; n = 0; }
            else if (tid < 5) { const int i = tid - 1; L = 1; bx = (i & 1) * 4; by = (i >> 1) * 4; nbx = 4; n = tid; }
            else if (tid < 21) { const int i = tid - 5; L = 2; bx = (i & 3) * 2; by = (i >> 2) * 2; nbx = 2; n = 5 + kTab16[i]; }
            else { const int i = tid - 21; L = 3; bx = i & 7; by = i >> 3; nbx = 1; n = 21 + kTab8[i]; }
            nby = nbx;
        }
        // best SAD / MV of the PU per list (internal index n)
        auto best_sad = [&](int l) { if constexpr (Ext) { if (n >= kNumPu) return job.scratch_ext[sb].best_sad[l][n - kNumPu]; } return scr.best_sad[l][n]; };
        auto best_mv = [&](int l) { if constexpr (Ext) { if (n >= kNumPu) return job.scratch_ext[sb].best_mv[l][n - kNumPu]; } return scr.best_mv[l][n]; };
        SvtMePuResult o;
        uint32_t* ow = reinterpret_cast<uint32_t*>(&o);
#pragma unroll
        for (int i = 0; i < 10; i++) ow[i] = 0;
        int total = p.num_lists;
        uint32_t bi = 0;
        if (p.num_lists == 2 && (p.cu8x8_mode == 0 || tid < 21 || Ext)) {
            for (int yy = 0; yy < nby; yy++) for (int xx = 0; xx < nbx; xx++) bi += part[L][(by + yy) * 8 + bx + xx];
            total = 3;
        }
        const uint32_t a = best_sad(0), b = p.num_lists == 2 ? best_sad(1) : 0u;
        o.total_cand = (uint8_t)total;
        o.xMvL0 = (int16_t)mv_x(best_mv(0)); o.yMvL0 = (int16_t)mv_y(best_mv(0));
        if (p.num_lists == 2) { o.xMvL1 = (int16_t)mv_x(best_mv(1)); o.yMvL1 = (int16_t)mv_y(best_mv(1)); }
#define SET(k, d, dirv) do { o.cand[k].distortion = (d); o.cand[k].direction = (dirv); } while (0)
        if (total == 3) {                                   // Sort3Elements, ME.c:5434-5463
            if (a <= b && a <= bi) {
                if (b <= bi) { SET(0, a, 0); SET(1, b, 1); SET(2, bi, 2); } else { SET(0, a, 0); SET(1, bi, 2); SET(2, b, 1); }
            } else if (b <= a && b <= bi) {
                if (a <= bi) { SET(0, b, 1); SET(1, a, 0); SET(2, bi, 2); } else { SET(0, b, 1); SET(1, bi, 2); SET(2, a, 0); }
            } else if (a <= b) { SET(0, bi, 2); SET(1, a, 0); SET(2, b, 1); }
            else               { SET(0, bi, 2); SET(1, b, 1); SET(2, a, 0); }
        } else if (total == 2) {
            if (a <= b) { SET(0, a, 0); SET(1, b, 1); } else { SET(0, b, 1); SET(1, a, 0); }
        } else SET(0, a, 0);
#undef SET
        uint32_t* dst = reinterpret_cast<uint32_t*>(&job.results[(size_t)sb * NPU + tid]);
#pragma unroll
        for (int i = 0; i < 10; i++) dst[i] = ow[i];
    }
    if (job.debug != nullptr) {
        SvtMeSbDebug& d = job.debug[sb];
        if (tid < 2) {
            const bool on = tid < p.num_lists;
            d.search_center_x[tid] = on ? (int16_t)scr.center_x[tid] : 0; d.search_center_y[tid] = on ? (int16_t)scr.center_y[tid] : 0;
            d.sa_origin_x[tid] = on ? (int16_t)scr.sa_x[tid] : 0; d.sa_origin_y[tid] = on ? (int16_t)scr.sa_y[tid] : 0;
            d.sa_width[tid] = on ? (int16_t)scr.sa_w[tid] : 0; d.sa_height[tid] = on ? (int16_t)scr.sa_h[tid] : 0;
        }
        for (int t = tid; t < 2 * kNumPu; t += kBpThreads) {
            const int l = t / kNumPu, i = t - l * kNumPu;
            d.best_sad[l][i] = l < p.num_lists ? scr.best_sad[l][i] : 0u;
            d.best_mv[l][i] = l < p.num_lists ? scr.best_mv[l][i] : 0u;
        }
    }
}

void launch_bipred_pack(const DJob* d_jobs, int n_jobs, int max_sb, bool ext, cudaStream_t st) {
    if (ext) bipred_pack_kernel<true><<<dim3(max_sb, n_jobs), kBpThreads, 0, st>>>(d_jobs);
    else     bipred_pack_kernel<false><<<dim3(max_sb, n_jobs), kBpThreads, 0, st>>>(d_jobs);
}

}  // namespace svtme
