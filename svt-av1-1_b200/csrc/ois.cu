// ois.cu -- open-loop intra search of the ME process for every SB of a picture (OpenLoopIntraSearchLcu, reference
// Source/Lib/Codec/EbMotionEstimation.c:8047-8365, called per ME segment at EbMotionEstimationProcess.c:559-583).
//
// Three generations of the kernel, bit-identical, the last one shipped: ois_kernel (pixel-wise, below),
// ois_words_kernel (4 pixels per lane), ois_lines_kernel (a line of the CU per lane, 32 / n modes side by side).
// ois_kernel: one CTA (8 warps) per SB. Phase 1: the 84 CUs of the SB (4 x 32x32, 16 x 16x16, 64 x 8x8) are dealt to the warps; a warp
// loads the CU's 4n + 1 neighbours from the source picture into shared memory, and for every mode the path asks for
// (1, 7, 3..9 or all 35) the lanes predict the CU's pixels (ois_core.cuh: one formula per pixel, no prediction buffer),
// take |src - pred| and reduce the SAD over the warp. Phase 2: one thread per CU turns the SADs into the reference's
// candidate list (injected lists per stage-1 winner, best-18 selection + the reference's swap sort on layer-0 pictures) and
// writes the byte image of OisCandidate_t. The reference's `bestMode` local survives from CU to CU; that only shows when no
// mode of a CU beats the initial 32 * 32 * 255, which phase 2 resolves by looking back over the earlier CUs.
// Work: sum of n^2 over the CUs = 12288 pixels per mode and SB; HBM traffic is the source SB + its border once (L2 serves the
// re-reads), 6.2 KB of results per SB: the kernel is bound by ALU issue at ~35 modes x 12288 pixels x ~12 instructions.
#include "me_common.cuh"
#include "ois_core.cuh"

namespace svtme {

using namespace ois;

constexpr int kOisThreads = 256;

// Phase 2 of every kernel below, entered by all threads after the barrier that ends phase 1 (s_sad / s_point complete):
// thread cu = 1..84 turns the CU's SADs into its candidate list.
__device__ __forceinline__ void write_candidate_lists(const OisJob& job, const Params& q, int sb, const uint32_t (*s_sad)[kNumModes],
                                                      const int8_t* s_point, int8_t* s_best, int tid) {
    // ---- 2a: the winner of every CU that ran a stage-1 / intra-picture search
    const int cu = tid;
    const bool searched = cu >= 1 && cu < kCus && s_point[cu] >= 0;
    int cx = 0, cy = 0, n = 8;
    if (cu >= 1 && cu < kCus) cu_geometry(cu, cx, cy, n);
    if (searched) s_best[cu] = (int8_t)cu_winner(q, n, s_sad[cu], s_point[cu]);
    __syncthreads();
    if (cu < 1 || cu >= kCus) return;
    // ---- 2b: candidate lists
    uint32_t c[kMaxCand];
    int total = 0;
    if (searched) {
        int bm = s_best[cu];
        for (int k = cu - 1; bm < 0 && k >= 1; k--) bm = s_best[k];                      // the reference's local survives from CU to CU
        if (bm < 0) bm = kPlanar;
        total = build_lists(q, n, s_sad[cu], s_point[cu], bm, c);
    } else {
#pragma unroll
        for (int i = 0; i < kMaxCand; i++) c[i] = 0;
    }
    uint32_t* out = job.out_cand + ((size_t)sb * kCus + cu) * kMaxCand;
#pragma unroll
    for (int i = 0; i < kMaxCand; i++) out[i] = c[i];
    job.out_total[(size_t)sb * kCus + cu] = (uint8_t)total;
}

__global__ void __launch_bounds__(kOisThreads) ois_kernel(const OisJob* __restrict__ jobs, int sb_cols, int n_sb) {
    __shared__ uint8_t s_r[8][132];                  // neighbour array of the warp's current CU
    __shared__ uint8_t s_refbuf[8][104];             // main reference of the current angular mode: index 32 = top-left
    __shared__ uint32_t s_sad[kCus][kNumModes];      // phase 1 -> phase 2 (inter path: [0..8] stage 1, [9] DC)
    __shared__ int8_t s_point[kCus];                 // inter path: OIS point; -1 = the CU was not searched
    __shared__ int8_t s_best[kCus];                  // stage-1 / intra winner of the CU, -1 = none found
    const OisJob& job = jobs[blockIdx.y];
    const int sb = blockIdx.x;
    if (sb >= n_sb) return;
    const Params q = job.prm;
    const Picture pic = { job.plane.at(0, 0), job.plane.stride, job.plane.w, job.plane.h };
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int ox = (sb % sb_cols) * 64, oy = (sb / sb_cols) * 64;
    const bool all_modes = all_modes_path(q);
    if (tid < kCus) { s_point[tid] = -1; s_best[tid] = -1; }
    __syncthreads();

    uint8_t* r = s_r[warp];
    uint8_t* ref = s_refbuf[warp] + 32;
    for (int cu = 1 + warp; cu < kCus; cu += 8) {
        int cx, cy, n;
        cu_geometry(cu, cx, cy, n);
        const int x0 = ox + cx, y0 = oy + cy, lg = n == 8 ? 3 : n == 16 ? 4 : 5;
        if (x0 + n > pic.w || y0 + n > pic.h) continue;                                  // raster_scan_cu_validity
        if (q.slice_type != 0 && q.cu8x8_mode == 1 && n == 8) continue;
        __syncwarp();
        for (int i = lane; i <= 4 * n; i += 32) r[i] = (uint8_t)neighbour(pic, x0, y0, n, i);
        __syncwarp();
        int dc = 0;
        {
            int part = 0;
            for (int i = lane; i < 2 * n; i += 32) part += i < n ? r[2 * n + 1 + i] : r[2 * n - 1 - (i - n)];
            dc = (__reduce_add_sync(0xffffffffu, part) + n) >> (lg + 1);
        }
        const uint8_t* src = pic.pix + (size_t)y0 * pic.stride + x0;
        auto mode_sad = [&](int mode) -> uint32_t {
            if (mode >= 2) {
                __syncwarp();
                for (int i = ref_first(n, mode) + lane; i <= 2 * n; i += 32) ref[i] = (uint8_t)ref_sample(r, n, mode, i);
                __syncwarp();
            }
            uint32_t part = 0;
            for (int px = lane; px < n * n; px += 32) {
                const int x = px & (n - 1), y = px >> lg;
                const int d = (int)src[(size_t)y * pic.stride + x] - pred_pixel(r, ref, n, lg, mode, x, y, dc);
                part += (uint32_t)(d < 0 ? -d : d);
            }
            return __reduce_add_sync(0xffffffffu, part);
        };
        if (q.slice_type == 0) {
            if (n == 32) { const uint32_t s = mode_sad(kPlanar); if (lane == 0) s_sad[cu][0] = s; }
            else for (int i = 0; i < 7; i++) { const uint32_t s = mode_sad(islice_mode(i)); if (lane == 0) s_sad[cu][i] = s; }
            if (lane == 0) s_point[cu] = 0;
        } else if (all_modes) {
            for (int m = 0; m < kNumModes; m++) { const uint32_t s = mode_sad(m); if (lane == 0) s_sad[cu][m] = s; }
            if (lane == 0) s_point[cu] = 0;
        } else {
            const uint32_t dcs = mode_sad(kDc);
            int point = 0;
            if (!q.limit_ois_to_dc) {
                point = ois_point(q, job.me_sad[(size_t)sb * kCus + cu], dcs);
                if (point) for (int i = 0; i < points_modes(point); i++) { const uint32_t s = mode_sad(stage1_mode(i)); if (lane == 0) s_sad[cu][i] = s; }
            }
            if (lane == 0) { s_sad[cu][9] = dcs; s_point[cu] = (int8_t)point; }
        }
    }
    __syncthreads();

    write_candidate_lists(job, q, sb, s_sad, s_point, s_best, tid);
}

// ---- the word-wise kernel (second generation, SVTME_OIS_KERNEL=2) ------------------------------------------------------------
// Phase 2 as above; phase 1 differs: the warp keeps the CU's source pixels in registers (as
// row words and, for the horizontal modes, as column words), a lane predicts 4 adjacent pixels at a time (ois_core.cuh:
// angular_word -- two aligned shared-memory words, two pixels per multiply-add) and VABSDIFF4-accumulates them; CUs are handed
// out largest first from a shared counter, so that the four 32x32 CUs do not make four of the eight warps twice as long.
template <int LG>
__device__ __forceinline__ void search_cu_words(const OisJob& job, const Params& q, const Picture& pic, int sb, int cu, int x0, int y0,
                                                uint8_t* r, uint8_t* vbuf, uint8_t* hbuf, uint32_t* sad_row, int8_t* s_point, int lane, bool all_modes) {
    constexpr int n = 1 << LG, W = (n * n) / (LG == 3 ? 2 : 4) / 32;
    __syncwarp();
    for (int i = lane; i <= 4 * n; i += 32) r[i] = (uint8_t)neighbour(pic, x0, y0, n, i);
    __syncwarp();
    int dc;
    {
        int part = 0;
        for (int i = lane; i < 2 * n; i += 32) part += i < n ? r[2 * n + 1 + i] : r[2 * n - 1 - (i - n)];
        dc = (__reduce_add_sync(0xffffffffu, part) + n) >> (LG + 1);
    }
    // main references of the vertical / horizontal modes, [32 + i], i = 0..2n; [32 + i], i < 0, is projected per mode
    for (int i = lane; i <= 2 * n; i += 32) {
        vbuf[32 + i] = (uint8_t)straight_sample(r, n, true, i);
        hbuf[32 + i] = (uint8_t)straight_sample(r, n, false, i);
    }
    __syncwarp();
    const uint8_t* src = pic.pix + (size_t)y0 * pic.stride + x0;
    uint32_t srow[W], scol[W];
#pragma unroll
    for (int k = 0; k < W; k++) {
        int j, i0;
        word_pos(LG, lane, k, j, i0);
        srow[k] = source_word(src, pic.stride, LG, j, i0, false);
        scol[k] = source_word(src, pic.stride, LG, j, i0, true);
    }
    auto mode_sad = [&](int mode) -> uint32_t {
        uint32_t part = 0;
        if (mode < 2) {
#pragma unroll
            for (int k = 0; k < W; k++) {
                int j, i0;
                word_pos(LG, lane, k, j, i0);
                part += sad_word(flat_word(r, LG, mode, j, i0, dc), srow[k]);
            }
        } else {
            const int a = angle_of(mode);
            const bool vertical = mode >= 18;
            uint8_t* refbuf = vertical ? vbuf : hbuf;
            if (a < 0) {                                                                 // at most n <= 32 projected samples
                __syncwarp();
                const int i = ref_first(n, mode) + lane;
                if (i < 0) refbuf[32 + i] = (uint8_t)ref_sample(r, n, mode, i);
                __syncwarp();
            }
#pragma unroll
            for (int k = 0; k < W; k++) {
                int j, i0;
                word_pos(LG, lane, k, j, i0);
                part += sad_word(angular_word(refbuf, r, LG, vertical, a, j, i0), vertical ? srow[k] : scol[k]);
            }
        }
        return __reduce_add_sync(0xffffffffu, part);
    };
    // one call site for mode_sad (the unrolled word loops are large): the path only decides which mode comes i-th and where its
    // SAD goes; the inter path starts with DC (slot 9), whose SAD against the ME SAD sets how many stage-1 directions follow
    const bool intra = q.slice_type == 0;
    int count = intra ? (n == 32 ? 1 : 7) : all_modes ? kNumModes : 1, point = 0;
    for (int i = 0; i < count; i++) {
        int mode, slot;
        if (intra) { mode = islice_mode(i); slot = i; }
        else if (all_modes) { mode = i; slot = i; }
        else if (i == 0) { mode = kDc; slot = 9; }
        else { mode = stage1_mode(i - 1); slot = i - 1; }
        const uint32_t s = mode_sad(mode);
        if (lane == 0) sad_row[slot] = s;
        if (!intra && !all_modes && i == 0 && !q.limit_ois_to_dc) {
            point = ois_point(q, job.me_sad[(size_t)sb * kCus + cu], s);
            if (point) count = 1 + points_modes(point);
        }
    }
    if (lane == 0) s_point[cu] = (int8_t)point;
}

__global__ void __launch_bounds__(kOisThreads, 4) ois_words_kernel(const OisJob* __restrict__ jobs, int sb_cols, int n_sb) {
    __shared__ uint8_t s_r[8][132];
    __shared__ uint32_t s_refbuf[8][2][26];          // per warp: vertical / horizontal main reference, 104 bytes, word aligned
    __shared__ uint32_t s_sad[kCus][kNumModes];
    __shared__ int8_t s_point[kCus];
    __shared__ int8_t s_best[kCus];
    __shared__ int s_next;
    const OisJob& job = jobs[blockIdx.y];
    const int sb = blockIdx.x;
    if (sb >= n_sb) return;
    const Params q = job.prm;
    const Picture pic = { job.plane.at(0, 0), job.plane.stride, job.plane.w, job.plane.h };
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int ox = (sb % sb_cols) * 64, oy = (sb / sb_cols) * 64;
    const bool all_modes = all_modes_path(q);
    if (tid < kCus) { s_point[tid] = -1; s_best[tid] = -1; }
    if (tid == 0) s_next = 1;
    __syncthreads();

    for (;;) {
        int cu = 0;
        if (lane == 0) cu = atomicAdd(&s_next, 1);
        cu = __shfl_sync(0xffffffffu, cu, 0);
        if (cu >= kCus) break;
        int cx, cy, n;
        cu_geometry(cu, cx, cy, n);
        const int x0 = ox + cx, y0 = oy + cy;
        if (x0 + n > pic.w || y0 + n > pic.h) continue;                                  // raster_scan_cu_validity
        if (q.slice_type != 0 && q.cu8x8_mode == 1 && n == 8) continue;
        uint8_t* vbuf = reinterpret_cast<uint8_t*>(s_refbuf[warp][0]);
        uint8_t* hbuf = reinterpret_cast<uint8_t*>(s_refbuf[warp][1]);
        if (n == 32) search_cu_words<5>(job, q, pic, sb, cu, x0, y0, s_r[warp], vbuf, hbuf, s_sad[cu], s_point, lane, all_modes);
        else if (n == 16) search_cu_words<4>(job, q, pic, sb, cu, x0, y0, s_r[warp], vbuf, hbuf, s_sad[cu], s_point, lane, all_modes);
        else search_cu_words<3>(job, q, pic, sb, cu, x0, y0, s_r[warp], vbuf, hbuf, s_sad[cu], s_point, lane, all_modes);
    }
    __syncthreads();

    write_candidate_lists(job, q, sb, s_sad, s_point, s_best, tid);
}

// ---- the line-per-lane kernel ----------------------------------------------------------------------------------------------
// As the word-wise kernel, but a CU of n x n pixels occupies n lanes, lane gl owning line gl of it (n / 4 words: a row for the
// vertical modes, planar and DC, a column for the horizontal ones), so that 32 / n modes of the CU run side by side in a warp: 4
// on the 8x8 CUs, where the per-mode set-up (angle, projected side reference, reduction, store) otherwise costs twice the
// prediction. Each group of lanes has its own pair of main references; the position along the mode list is base + group.
template <int LG>
__device__ __forceinline__ void search_cu_lines(const OisJob& job, const Params& q, const Picture& pic, int sb, int cu, int x0, int y0,
                                                uint8_t* r, uint8_t* refs, uint32_t* sad_row, int8_t* s_point, int lane, bool all_modes,
                                                const int8_t* s_list, const int8_t* s_angle) {
    constexpr int n = 1 << LG, M = 32 / n, W = n / 4;
    const int g = lane >> LG, gl = lane & (n - 1);
    __syncwarp();
    for (int i = lane; i <= 4 * n; i += 32) r[i] = (uint8_t)neighbour(pic, x0, y0, n, i);
    __syncwarp();
    int dc;
    {
        int part = 0;
        for (int i = lane; i < 2 * n; i += 32) part += i < n ? r[2 * n + 1 + i] : r[2 * n - 1 - (i - n)];
        dc = (__reduce_add_sync(0xffffffffu, part) + n) >> (LG + 1);
    }
    uint8_t* vbuf = refs + g * 208;
    uint8_t* hbuf = vbuf + 104;
    for (int i = gl; i <= 2 * n; i += n) {
        vbuf[32 + i] = (uint8_t)straight_sample(r, n, true, i);
        hbuf[32 + i] = (uint8_t)straight_sample(r, n, false, i);
    }
    const uint8_t* src = pic.pix + (size_t)y0 * pic.stride + x0;
    uint32_t srow[W], scol[W];
#pragma unroll
    for (int k = 0; k < W; k++) {
        srow[k] = source_pixels(src, pic.stride, 4, gl, 4 * k, false);
        scol[k] = source_pixels(src, pic.stride, 4, gl, 4 * k, true);
    }
    const bool intra = q.slice_type == 0, staged = !intra && !all_modes;        // staged: DC first, then the stage-1 directions
    int count = intra ? (n == 32 ? 1 : 7) : all_modes ? kNumModes : 1, point = 0, base = 0;
    while (base < count) {
        const bool dc_pass = staged && base == 0;
        const int p = base + g;
        const bool active = dc_pass ? g == 0 : p < count;
        int mode = kDc, slot = 9;
        if (active && !dc_pass) {
            slot = staged ? p - 1 : p;
            mode = s_list[slot];
        }
        const bool angular = mode >= 2, vertical = mode >= 18;
        const int a = s_angle[mode];
        uint8_t* refbuf = vertical ? vbuf : hbuf;
        __syncwarp();
        if (active && a < 0) {                                                   // at most n projected samples: one per lane of the group
            const int i = ((n * a) >> 5) + gl;                                   // ref_first(n, mode) + gl
            if (i < 0) refbuf[32 + i] = (uint8_t)ref_sample(r, n, mode, i);
        }
        __syncwarp();
        uint32_t part = 0;
        if (active) {
            if (angular) {
#pragma unroll
                for (int k = 0; k < W; k++) part += sad_word(angular_pixels(refbuf, r, LG, vertical, a, gl, 4 * k, 4), vertical ? srow[k] : scol[k]);
            } else if (mode == kDc) {
#pragma unroll
                for (int k = 0; k < W; k++) part += sad_word(dc_line_word(r, LG, gl, 4 * k, dc), srow[k]);
            } else {
#pragma unroll
                for (int k = 0; k < W; k++) part += sad_word(planar_line_word(r, LG, gl, 4 * k), srow[k]);
            }
        }
#pragma unroll
        for (int d = n / 2; d >= 1; d >>= 1) part += __shfl_xor_sync(0xffffffffu, part, d);
        if (active && gl == 0) sad_row[slot] = part;
        if (dc_pass) {
            base = 1;
            if (!q.limit_ois_to_dc) {
                point = ois_point(q, job.me_sad[(size_t)sb * kCus + cu], __shfl_sync(0xffffffffu, part, 0));
                if (point) count = 1 + points_modes(point);
            }
        } else {
            base += M;
        }
    }
    if (lane == 0) s_point[cu] = (int8_t)point;
}

__global__ void __launch_bounds__(kOisThreads, 4) ois_lines_kernel(const OisJob* __restrict__ jobs, int sb_cols, int n_sb) {
    __shared__ uint8_t s_r[8][132];
    __shared__ uint32_t s_refs[8][4][2][26];         // per warp and group of lanes: vertical / horizontal main reference, 104 bytes each
    __shared__ int8_t s_list[kNumModes];             // the path's modes in the order they are tried (slot -> mode)
    __shared__ int8_t s_angle[kNumModes];            // mode -> intraPredAngle (0 for planar / DC)
    __shared__ uint32_t s_sad[kCus][kNumModes];
    __shared__ int8_t s_point[kCus];
    __shared__ int8_t s_best[kCus];
    __shared__ int s_next;
    const OisJob& job = jobs[blockIdx.y];
    const int sb = blockIdx.x;
    if (sb >= n_sb) return;
    const Params q = job.prm;
    const Picture pic = { job.plane.at(0, 0), job.plane.stride, job.plane.w, job.plane.h };
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int ox = (sb % sb_cols) * 64, oy = (sb / sb_cols) * 64;
    const bool all_modes = all_modes_path(q);
    if (tid < kCus) { s_point[tid] = -1; s_best[tid] = -1; }
    if (tid < kNumModes) {
        s_list[tid] = (int8_t)(q.slice_type == 0 ? (tid < 7 ? islice_mode(tid) : kDc) : all_modes ? tid : tid < 9 ? stage1_mode(tid) : kDc);
        s_angle[tid] = (int8_t)(tid >= 2 ? angle_of(tid) : 0);
    }
    if (tid == 0) s_next = 1;
    __syncthreads();

    for (;;) {
        int cu = 0;
        if (lane == 0) cu = atomicAdd(&s_next, 1);
        cu = __shfl_sync(0xffffffffu, cu, 0);
        if (cu >= kCus) break;
        int cx, cy, n;
        cu_geometry(cu, cx, cy, n);
        const int x0 = ox + cx, y0 = oy + cy;
        if (x0 + n > pic.w || y0 + n > pic.h) continue;                                  // raster_scan_cu_validity
        if (q.slice_type != 0 && q.cu8x8_mode == 1 && n == 8) continue;
        uint8_t* refs = reinterpret_cast<uint8_t*>(s_refs[warp]);
        if (n == 32) search_cu_lines<5>(job, q, pic, sb, cu, x0, y0, s_r[warp], refs, s_sad[cu], s_point, lane, all_modes, s_list, s_angle);
        else if (n == 16) search_cu_lines<4>(job, q, pic, sb, cu, x0, y0, s_r[warp], refs, s_sad[cu], s_point, lane, all_modes, s_list, s_angle);
        else search_cu_lines<3>(job, q, pic, sb, cu, x0, y0, s_r[warp], refs, s_sad[cu], s_point, lane, all_modes, s_list, s_angle);
    }
    __syncthreads();

    write_candidate_lists(job, q, sb, s_sad, s_point, s_best, tid);
}

void launch_ois(const OisJob* d_jobs, int n_jobs, int sb_cols, int n_sb, int variant, cudaStream_t st) {
    if (variant == 1) ois_kernel<<<dim3(n_sb, n_jobs), kOisThreads, 0, st>>>(d_jobs, sb_cols, n_sb);
    else if (variant == 2) ois_words_kernel<<<dim3(n_sb, n_jobs), kOisThreads, 0, st>>>(d_jobs, sb_cols, n_sb);
    else ois_lines_kernel<<<dim3(n_sb, n_jobs), kOisThreads, 0, st>>>(d_jobs, sb_cols, n_sb);      // the shipped one; the other two are kept for A/B
}

}  // namespace svtme
