// bipred.cu -- bi-prediction SAD, candidate ordering and result packing for every PU of every SB:
// the tail of MotionEstimateLcu (reference Source/Lib/Codec/EbMotionEstimation.c:6973-7146) with
// BiPredictionSearch / BiPredictionCompensation / BiPredAverging (:4933-5342) and Sort3Elements (:5434).
//
// The list-0 uni-prediction block of every PU is materialised by the list-0 pass of subpel_kernel (one 64x64 byte image per
// PU-size level); the list-1 pass computes SAD(src, (p0 + p1 + 1) >> 1) per 8x8 block from it and its own prediction blocks
// and packs the rows (pack.cuh). This kernel is what is left: packing for pictures with ONE list, and the debug records.
#include "me_common.cuh"
#include "pack.cuh"

namespace svtme {

constexpr int kBpThreads = 256;

// Ext = the 209-PU mode (14 PU classes; the non-square PUs and their prediction blocks live in SbScratchExt)
template <bool Ext>
__global__ void __launch_bounds__(kBpThreads) bipred_pack_kernel(const DJob* __restrict__ jobs) {
    constexpr int NPU = Ext ? kMaxPu : kNumPu;
    const DJob& job = jobs[blockIdx.y];
    const int sb = job.sb_begin + blockIdx.x;
    if (sb >= job.sb_end) return;
    const SvtMeParams& p = job.prm;
    if ((p.num_pus == kMaxPu) != Ext) return;
    const int tid = threadIdx.x;
    const SbScratch& scr = job.scratch[sb];

    // Pictures with two lists were packed by the list-1 pass of the sub-pel kernel (it computes the bi-prediction SADs while the
    // list-1 prediction blocks are in registers: the 16 KB of them per SB never travel through memory); this kernel packs the
    // pictures with one list and writes the optional debug records.
    if (p.num_lists != 2 && tid < NPU) {
        auto best_sad = [&](int l, int n) { if constexpr (Ext) { if (n >= kNumPu) return job.scratch_ext[sb].best_sad[l][n - kNumPu]; } return scr.best_sad[l][n]; };
        auto best_mv = [&](int l, int n) { if constexpr (Ext) { if (n >= kNumPu) return job.scratch_ext[sb].best_mv[l][n - kNumPu]; } return scr.best_mv[l][n]; };
        pack_pu_result<Ext>(job, sb, tid, nullptr, best_sad, best_mv);
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
