// hme.cu -- search-centre derivation for one reference list: hme_mv_center_check, HME level 0/1/2,
// region selection, CheckZeroZeroCenter and the search-area clipping of MotionEstimateLcu
// (reference Source/Lib/Codec/EbMotionEstimation.c:5882-6145, 4306-4758, 6300-6733, 5466-5552).
//
// One CTA (256 threads) per SB. The 2x2 HME regions are independent chains (level 0 -> 1 -> 2 of the
// same region), so each region is owned by a GROUP of 64 threads (two warps) that synchronises on its
// own named barrier and stages its own reference window in shared memory; the four chains run
// concurrently without CTA-wide barriers. Inside a group one thread evaluates 4 consecutive search
// positions of one row: the window is staged word-aligned, so the four byte offsets are constant funnel
// shifts of the same two words (3 SHF + 4 VABSDIFF4.ACC per source word). Two neighbouring lanes form one
// of the reference's 8-position SIMD groups, which is what quirk Q3 (HME level 1:
// min(lo+hi, 65535 + min_group(lo))) needs. argmin: 32-bit key (value, row, col) -> smallest value,
// first in raster order; thread-local min, CREDUX.MIN per warp, one shared atomicMin per warp.
#include "me_common.cuh"

namespace svtme {

constexpr int kHmeThreads = 256;
constexpr int kGroupThreads = 64;
constexpr int kRegionWinWords = 2560;
constexpr int kSrcPitch = 18, kQPitch = 10;      // padded so that lanes on different block rows hit different banks          // 10 KB window buffer per region group

struct HmeShared {
    uint32_t src_even[32 * kSrcPitch];   // even rows of the 64x64 source SB
    uint32_t q_even[16 * kQPitch];       // even rows of the 32x32 quarter SB
    uint32_t s_even[8 * 4];         // even rows of the 16x16 sixteenth SB
    uint32_t win[4][kRegionWinWords];
    uint32_t best[12];              // [level][region] winning keys
    uint32_t cand_sad[8];
    int x[3][2][2], y[3][2][2];
    uint32_t sad[3][2][2];
    int center[2];
};

// Named barrier of one 64-thread region group (ids 1..4; constant ids so ptxas reserves 5 barriers, not 16).
__device__ __forceinline__ void group_bar(int id) {
    switch (id) {
    case 1: asm volatile("bar.sync 1, %0;" :: "n"(kGroupThreads) : "memory"); break;
    case 2: asm volatile("bar.sync 2, %0;" :: "n"(kGroupThreads) : "memory"); break;
    case 3: asm volatile("bar.sync 3, %0;" :: "n"(kGroupThreads) : "memory"); break;
    default: asm volatile("bar.sync 4, %0;" :: "n"(kGroupThreads) : "memory"); break;
    }
}

// SAD of the (sub-sampled) SB against the reference picture at displacement (dx,dy): one warp, lane = source row.
// NxMSadKernel with doubled strides, ME.c:5918-5926. Split in two so that the reference words can be in flight while the
// source SB is still being staged (the centre check is L2-latency bound: three dependent round trips per CTA before).
struct SbSadRow { uint32_t v[17]; uint32_t sh; };
__device__ __forceinline__ void warp_sb_sad_load(SbSadRow& row, const DPlane& rp, int x, int y, int w_words, int rows) {
    const int lane = threadIdx.x & 31;
    row.sh = 0;
#pragma unroll
    for (int c = 0; c < 17; c++) row.v[c] = 0u;
    if (lane < rows) {
        // 17 independent aligned loads in flight, then one funnel shift per word
        const uint8_t* r = rp.at(x, y + 2 * lane);
        const uint32_t* w = reinterpret_cast<const uint32_t*>((uintptr_t)r & ~(uintptr_t)3);
        row.sh = ((uint32_t)(uintptr_t)r & 3u) * 8u;
#pragma unroll
        for (int c = 0; c < 17; c++) if (c <= w_words) row.v[c] = __ldg(w + c);
    }
}
__device__ __forceinline__ uint32_t warp_sb_sad_finish(const SbSadRow& row, const uint32_t* src_even, int w_words, int rows) {
    const int lane = threadIdx.x & 31;
    uint32_t acc = 0;
    if (lane < rows) {
#pragma unroll
        for (int c = 0; c < 16; c++)
            if (c < w_words) acc = sad4(src_even[lane * kSrcPitch + c], __funnelshift_r(row.v[c], row.v[c + 1], row.sh), acc);
    }
    return __reduce_add_sync(0xffffffffu, acc);
}
__device__ uint32_t warp_sb_sad(const uint32_t* src_even, const DPlane& rp, int x, int y, int w_words, int rows) {
    SbSadRow row;
    warp_sb_sad_load(row, rp, x, y, w_words, rows);
    return warp_sb_sad_finish(row, src_even, w_words, rows);
}

__device__ __forceinline__ void cp_async16(uint32_t smem_addr, const void* gptr) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"(smem_addr), "l"(gptr) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() {
    asm volatile("cp.async.commit_group;\n\tcp.async.wait_group 0;" ::: "memory");
}

// Exhaustive search of one region by one 64-thread group. Returns the winning key
// (value << VS | row << RS | col) to every thread of the group.
//   src: even rows of the block, SP words per row;  W words x `rows` rows
//   (rx, ry): top-left search position in level coordinates; saw x sah positions
// Thread tile = one of the reference's 8-position SIMD groups (8 consecutive x on one search row): every
// window word is funnel-shifted to its three other byte phases once (3 SHF) and feeds 8 VABSDIFF4 (4 phases x
// 2 word offsets), and every lane reads DIFFERENT window words, so a shared-memory wavefront carries 32 useful
// words (sharing words between lanes made the search wavefront-bound). RSPLIT lanes split the block rows of a
// tile (interleaved rows, odd window pitch -> conflict-free banks) and add their partial sums with shuffles:
// level 2 has only 64 positions but 2048 pixels each.
template <int W, int SP, bool QUIRK, int VS, int RS, int RSPLIT>
__device__ uint32_t group_search(uint32_t* win, uint32_t* best_slot, int bar_id, int gtid, const uint32_t* src, int rows,
                                 const DPlane& rp, int rx, int ry, int saw, int sah) {
    const int tpr = (saw + 7) >> 3;                     // 8-position tiles per search row
    const int pitch = (2 * tpr + W) | 1;                // window words per row (odd)
    const int extra = 2 * (rows - 1);
    // (2560 / pitch by an approximate float division: a non-integer quotient is at least 1/pitch >= 1/131 away from an
    // integer, the approximation error is below 1e-3, and +0.002 keeps exact quotients from truncating one too low)
    const int chunk = max(1, (int)(__fdividef((float)kRegionWinWords, (float)pitch) + 0.002f) - extra);
    const int rs = gtid & (RSPLIT - 1);
    constexpr int kTilesPerPass = kGroupThreads / RSPLIT;
    // t / d as umulhi(t, 2^32 / d + 1): exact while t * d < 2^32 (t, d are a few thousand at most here); d == 1 would
    // need the multiplier 2^32, so a one-tile row (search width <= 8) keeps the plain index
    // reciprocal over-estimated by 2e-6 relative (>= the 2.4e-7 error of the approximate division): umulhi(t, inv) is then
    // floor(t / tpr + t / tpr * 2.3e-6), exact for the t < 2^13, tpr <= 63 of a window chunk; an integer division here
    // cost every thread ~25 instructions per level
    const uint32_t inv_tpr = tpr > 1 ? (uint32_t)(__fdividef(4294967296.0f, (float)tpr) * 1.000002f) : 0u;
    uint32_t best = 0xFFFFFFFFu;
    if (gtid == 0) *best_slot = 0xFFFFFFFFu;
    for (int i0 = 0; i0 < sah; i0 += chunk) {
        const int i1 = min(sah, i0 + chunk);
        const int wrows = i1 - i0 + extra;
        const int wcols = 2 * tpr + W;
        group_bar(bar_id);                              // previous users of the window are done
        {   // lane = (row within the warp's row bundle, window column): narrow windows (<= 16 / <= 8 words) put 2 / 4 rows
            // on a warp so that few lanes idle; the two warps of the group take alternate bundles. The row pointer advances
            // by a constant and the misalignment of the window's first byte is the same on every row (the pitch is a
            // multiple of 64), so a word costs two loads, one funnel shift and one store (the per-word division and
            // 64-bit address arithmetic of a flat index were a quarter of this kernel's instructions)
            const uint8_t* p0 = rp.at(rx, ry + i0);
            const uint32_t m = (uint32_t)(uintptr_t)p0 & 3u, sh8 = m * 8u;
            const int csh = wcols <= 8 ? 3 : wcols <= 16 ? 4 : 5, rpi = 32 >> csh;       // rows per warp iteration
            const int lane = gtid & 31, c0 = lane & ((1 << csh) - 1), rfirst = (gtid >> 5) * rpi + (lane >> csh);
            const int wstride = rp.stride >> 2;
            for (int c = c0; c < wcols; c += 32) {
                const uint32_t* a = reinterpret_cast<const uint32_t*>(p0 - m) + c + rfirst * wstride;
                uint32_t* d = win + rfirst * pitch + c;
#pragma unroll 4
                for (int r = rfirst; r < wrows; r += 2 * rpi, a += 2 * rpi * wstride, d += 2 * rpi * pitch)
                    *d = __funnelshift_r(__ldg(a), __ldg(a + 1), sh8);
            }
        }
        group_bar(bar_id);
        const int ntiles = (i1 - i0) * tpr;
        for (int t0 = 0; t0 < ntiles; t0 += kTilesPerPass) {
            const int t = t0 + gtid / RSPLIT;
            const bool live = t < ntiles;
            const int i = !live ? 0 : tpr > 1 ? (int)__umulhi((uint32_t)t, inv_tpr) : t, jt = live ? t - i * tpr : 0;
            uint32_t lo[2][4], hi[2][4];
#pragma unroll
            for (int k = 0; k < 2; k++)
#pragma unroll
                for (int ph = 0; ph < 4; ph++) { lo[k][ph] = 0; hi[k][ph] = 0; }
            if (live) {
                const uint32_t* wr = win + (i + 2 * rs) * pitch + 2 * jt;
                const uint32_t* sr = src + rs * SP;
                for (int r = rs; r < rows; r += RSPLIT, wr += 2 * RSPLIT * pitch, sr += RSPLIT * SP) {
                    uint32_t a = wr[0], sprev = 0;
                    uint2 spair = make_uint2(0u, 0u);
#pragma unroll
                    for (int c = 0; c <= W; c++) {
                        const uint32_t b = wr[c + 1];
                        const uint32_t v1 = __funnelshift_r(a, b, 8), v2 = __funnelshift_r(a, b, 16), v3 = __funnelshift_r(a, b, 24);
                        if (c < W) {                    // positions 8jt + phase: source word c
                            if ((c & 1) == 0) spair = *reinterpret_cast<const uint2*>(sr + c);   // rows are 8-byte aligned: W, SP even
                            const uint32_t sw = (c & 1) ? spair.y : spair.x;
                            if (QUIRK && c >= 4) { hi[0][0] = sad4(sw, a, hi[0][0]); hi[0][1] = sad4(sw, v1, hi[0][1]); hi[0][2] = sad4(sw, v2, hi[0][2]); hi[0][3] = sad4(sw, v3, hi[0][3]); }
                            else                 { lo[0][0] = sad4(sw, a, lo[0][0]); lo[0][1] = sad4(sw, v1, lo[0][1]); lo[0][2] = sad4(sw, v2, lo[0][2]); lo[0][3] = sad4(sw, v3, lo[0][3]); }
                            if (c >= 1) {               // positions 8jt + 4 + phase: source word c - 1
                                if (QUIRK && c - 1 >= 4) { hi[1][0] = sad4(sprev, a, hi[1][0]); hi[1][1] = sad4(sprev, v1, hi[1][1]); hi[1][2] = sad4(sprev, v2, hi[1][2]); hi[1][3] = sad4(sprev, v3, hi[1][3]); }
                                else                     { lo[1][0] = sad4(sprev, a, lo[1][0]); lo[1][1] = sad4(sprev, v1, lo[1][1]); lo[1][2] = sad4(sprev, v2, lo[1][2]); lo[1][3] = sad4(sprev, v3, lo[1][3]); }
                            }
                            sprev = sw;
                        } else {
                            if (QUIRK && W - 1 >= 4) { hi[1][0] = sad4(sprev, a, hi[1][0]); hi[1][1] = sad4(sprev, v1, hi[1][1]); hi[1][2] = sad4(sprev, v2, hi[1][2]); hi[1][3] = sad4(sprev, v3, hi[1][3]); }
                            else                     { lo[1][0] = sad4(sprev, a, lo[1][0]); lo[1][1] = sad4(sprev, v1, lo[1][1]); lo[1][2] = sad4(sprev, v2, lo[1][2]); lo[1][3] = sad4(sprev, v3, lo[1][3]); }
                        }
                        a = b;
                    }
                }
            }
            if (RSPLIT > 1) {
#pragma unroll
                for (int o = 1; o < RSPLIT; o <<= 1)
#pragma unroll
                    for (int k = 0; k < 2; k++)
#pragma unroll
                        for (int ph = 0; ph < 4; ph++) {
                            lo[k][ph] += __shfl_xor_sync(0xffffffffu, lo[k][ph], o);
                            if (QUIRK) hi[k][ph] += __shfl_xor_sync(0xffffffffu, hi[k][ph], o);
                        }
            }
            if (live) {
                uint32_t mlo = 65535u;                   // quirk Q3: min of the low halves over the 8-position group
                if (QUIRK) {
#pragma unroll
                    for (int k = 0; k < 2; k++)
#pragma unroll
                        for (int ph = 0; ph < 4; ph++) mlo = min(mlo, (8 * jt + 4 * k + ph < saw) ? lo[k][ph] : 65535u);
                }
#pragma unroll
                for (int k = 0; k < 2; k++)
#pragma unroll
                    for (int ph = 0; ph < 4; ph++) {
                        const int x = 8 * jt + 4 * k + ph;
                        const uint32_t val = QUIRK ? min(lo[k][ph] + hi[k][ph], 65535u + mlo) : lo[k][ph];
                        const uint32_t key = (val << VS) | ((uint32_t)(i0 + i) << RS) | (uint32_t)x;
                        if (x < saw) best = min(best, key);
                    }
            }
        }
    }
    best = __reduce_min_sync(0xffffffffu, best);
    if ((gtid & 31) == 0) atomicMin(best_slot, best);
    group_bar(bar_id);
    return *best_slot;
}

// The same search for ANY block width (edge SBs 8 / 16 / 24 / 48 pixels wide, i.e. 2 .. 48-pixel blocks on the decimated
// levels; the reference runs them through the C SadLoopKernel or the 4 / 8 / 16 / 24 / 48 cases of the AVX2 one,
// ME.c:4417-4496, 4579-4615, 4711-4750). One search position per thread, the block width in pixels at run time, the last
// partial word masked. Only the right-most SB column of such pictures comes here, so this path is written for
// size, not speed. Quirk Q12: 4-pixel-wide blocks with rows % 4 == 0 -- inside the full 8-position groups the AVX2
// kernels pair reference row 2 with source row 0 and reference row 0 with source row 2 of every four rows
// (EbComputeSAD_Intrinsic_AVX2.c:1650-1652, 3973-3975); the leftover positions (search width & 7) are right.
template <int VS, int RS>
__device__ uint32_t group_search_any(uint32_t* win, uint32_t* best_slot, int bar_id, int gtid, const uint32_t* src, int sp, int wpx, int rows,
                                     const DPlane& rp, int rx, int ry, int saw, int sah) {
    const int W = (wpx + 3) >> 2;
    const uint32_t last_mask = (wpx & 3) ? (0xFFFFFFFFu >> (32 - 8 * (wpx & 3))) : 0xFFFFFFFFu;
    const bool q12 = wpx == 4 && (rows & 3) == 0;
    const int full = saw & ~7;                          // positions below this belong to full 8-position groups
    const int wcols = ((saw + 3) >> 2) + W + 1;         // window words per row
    const int pitch = wcols | 1;
    const int extra = 2 * (rows - 1);
    const int chunk = max(1, kRegionWinWords / pitch - extra);
    uint32_t best = 0xFFFFFFFFu;
    if (gtid == 0) *best_slot = 0xFFFFFFFFu;
    for (int i0 = 0; i0 < sah; i0 += chunk) {
        const int i1 = min(sah, i0 + chunk);
        const int wrows = i1 - i0 + extra;
        group_bar(bar_id);
        {
            const uint8_t* p0 = rp.at(rx, ry + i0);
            const uint32_t m = (uint32_t)(uintptr_t)p0 & 3u, sh8 = m * 8u;
            const int wstride = rp.stride >> 2;
            for (int t = gtid; t < wrows * wcols; t += kGroupThreads) {
                const int r = t / wcols, c = t - r * wcols;
                const uint32_t* a = reinterpret_cast<const uint32_t*>(p0 - m) + c + r * wstride;
                win[r * pitch + c] = __funnelshift_r(__ldg(a), __ldg(a + 1), sh8);
            }
        }
        group_bar(bar_id);
        for (int t = gtid; t < (i1 - i0) * saw; t += kGroupThreads) {
            const int i = t / saw, x = t - i * saw;
            const uint32_t sh8 = (x & 3) * 8u;
            const bool swap = q12 && x < full;
            uint32_t sad = 0;
            for (int r = 0; r < rows; r++) {
                const int rr = (swap && !(r & 1)) ? (r ^ 2) : r;
                const uint32_t* wr = win + (i + 2 * rr) * pitch + (x >> 2);
                const uint32_t* sr = src + r * sp;
                for (int c = 0; c < W; c++) {
                    uint32_t v = __funnelshift_r(wr[c], wr[c + 1], sh8), sw = sr[c];
                    if (c == W - 1) { v &= last_mask; sw &= last_mask; }
                    sad = sad4(sw, v, sad);
                }
            }
            best = min(best, (sad << VS) | ((uint32_t)(i0 + i) << RS) | (uint32_t)x);
        }
    }
    best = __reduce_min_sync(0xffffffffu, best);
    if ((gtid & 31) == 0) atomicMin(best_slot, best);
    group_bar(bar_id);
    return *best_slot;
}

__device__ __forceinline__ int round_sa_width(int w) { return (w < 8) ? 8 : (w & 7) ? w + (w - ((w >> 3) << 3)) : w; }

__global__ void __launch_bounds__(kHmeThreads, 4) hme_kernel(const DJob* __restrict__ jobs, int list) {
    __shared__ __align__(16) HmeShared sh;
    const DJob& job = jobs[blockIdx.y];
    const int sb = job.sb_begin + blockIdx.x;
    if (sb >= job.sb_end) return;
    const SvtMeParams& p = job.prm;
    if (list >= p.num_lists) return;
    const int tid = threadIdx.x, warp = tid >> 5;
    const int ox = (sb % job.sb_cols) * 64, oy = (sb / job.sb_cols) * 64;
    const int sb_w = min(64, p.width - ox), sb_h = min(64, p.height - oy);
    SbScratch& scr = job.scratch[sb];
    const DPyramid& ref = job.ref[list];

    // ---- hme_mv_center_check, ME.c:5882-6145: the candidates' reference rows are requested first, so that their L2 round
    // trip overlaps the staging of the source SB below (position A is evaluated at the zero-MV address: quirk Q2)
    const bool do_center = (p.temporal_layer_index > 0 || list == 0);
    const int tw = p.hme_l0_total_w, th = p.hme_l0_total_h;
    int dxv = 0, dyv = 0;
    SbSadRow cand_row;
    const bool cand_live = do_center && (warp < 4 || (warp == 4 && list == 1));
    if (do_center && list == 1) {
        const uint32_t mv0 = scr.best_mv[0][0];
        dxv = (int16_t)(0 - (mv_x(mv0) >> 2)); dyv = (int16_t)(0 - (mv_y(mv0) >> 2));
    }
    if (cand_live) {
        int cx = 0, cy = 0;
        if (warp == 1) cx = tw; else if (warp == 2) cy = -th; else if (warp == 3) cy = th;
        else if (warp == 4) { cx = dxv; cy = dyv; }
        warp_sb_sad_load(cand_row, ref.full, ox + clip_center(ox, cx, 63, ref.full.w), oy + clip_center(oy, cy, 63, ref.full.h), sb_w >> 2, sb_h >> 1);
    }
    // ---- stage the source SB (even rows only: every SAD on this path is row-subsampled)
    for (int t = tid; t < 32 * 16; t += kHmeThreads)
        sh.src_even[(t >> 4) * kSrcPitch + (t & 15)] = *reinterpret_cast<const uint32_t*>(job.cur.full.at(ox + 4 * (t & 15), oy + 2 * (t >> 4)));
    if (tid < 16 * 8)
        sh.q_even[(tid >> 3) * kQPitch + (tid & 7)] = *reinterpret_cast<const uint32_t*>(job.cur.quarter.at((ox >> 1) + 4 * (tid & 7), (oy >> 1) + 2 * (tid >> 3)));
    if (tid < 8 * 4)
        sh.s_even[tid] = *reinterpret_cast<const uint32_t*>(job.cur.sixteenth.at((ox >> 2) + 4 * (tid & 3), (oy >> 2) + 2 * (tid >> 2)));
    __syncthreads();

    int xsc = 0, ysc = 0;
    if (do_center) {
        if (warp < 5) {
            const uint32_t s = cand_live ? warp_sb_sad_finish(cand_row, sh.src_even, sb_w >> 2, sb_h >> 1) : 0xFFFFFFFFu;
            if ((tid & 31) == 0) sh.cand_sad[warp] = s;
        }
        __syncthreads();
        {
            const uint32_t zero = sh.cand_sad[0], b = sh.cand_sad[1], c = sh.cand_sad[2], d = sh.cand_sad[3], dir = sh.cand_sad[4];
            const uint32_t best = min(zero, min(b, min(c, min(d, dir))));
            if (best == zero) { xsc = 0; ysc = 0; }
            else if (best == b) { xsc = tw; ysc = 0; }
            else if (best == c) { xsc = 0; ysc = -th; }
            else if (best == dir) { xsc = dxv; ysc = dyv; }
            else { xsc = 0; ysc = th; }
        }
        // ---- HME, ME.c:6323-6637
        if (p.enable_hme_flag && sb_h == 64) {
            const int nw = p.num_hme_regions_w, nh = p.num_hme_regions_h;
            // quirk Q7 (ME.c:6323-6345): the reference initialises the per-region centres for list 0 only;
            // list 1 inherits whatever list 0 left there (visible when an HME level is disabled).
            const int g = tid >> 6, gtid = tid & 63;             // region group
            if (g < nw * nh) {
                const int w = g % nw, h = g / nw, bar = 1 + g;
                uint32_t* win = sh.win[g];
                if (gtid < 3) {                                   // the group initialises its own region (a group barrier, not a CTA one)
                    sh.x[gtid][w][h] = list == 0 ? xsc : scr.hme_x[gtid][w][h];
                    sh.y[gtid][w][h] = list == 0 ? ysc : scr.hme_y[gtid][w][h];
                    sh.sad[gtid][w][h] = 0;
                }
                group_bar(bar);
                const bool half = sb_w == 32;                     // right-edge SB of a width % 64 == 32 picture
                const bool odd = sb_w != 64 && sb_w != 32;        // 8 / 16 / 24 / 48-wide edge SB: the any-width search
                if (p.enable_hme_level0) {                        // HmeLevel0, ME.c:4306-4503 (sixteenth picture)
                    const DPlane& rp = ref.sixteenth;
                    int saw = (int16_t)((p.hme_l0_w[w] * p.hme_l0_mult_x) / 100);
                    int sah = (int16_t)((p.hme_l0_h[h] * p.hme_l0_mult_y) / 100);
                    int xd = xsc >> 2, yd = ysc >> 2;
                    for (int i = w; i > 0; i--) xd += (int16_t)((p.hme_l0_w[i - 1] * p.hme_l0_mult_x) / 100);
                    for (int i = h; i > 0; i--) yd += (int16_t)((p.hme_l0_h[i - 1] * p.hme_l0_mult_y) / 100);
                    int sax = -(int16_t)(((p.hme_l0_total_w * p.hme_l0_mult_x) / 100) >> 1) + xd;
                    int say = -(int16_t)(((p.hme_l0_total_h * p.hme_l0_mult_y) / 100) >> 1) + yd;
                    clip_search_area(ox >> 2, sax, saw, rp.ox - 1, rp.w);
                    clip_search_area(oy >> 2, say, sah, rp.oy - 1, rp.h);
                    const uint32_t k = odd ? group_search_any<17, 9>(win, &sh.best[g], bar, gtid, sh.s_even, 4, sb_w >> 2, 8, rp, (ox >> 2) + sax, (oy >> 2) + say, saw, sah) : half
                        ? group_search<2, 4, false, 17, 9, 1>(win, &sh.best[g], bar, gtid, sh.s_even, 8, rp, (ox >> 2) + sax, (oy >> 2) + say, saw, sah)
                        : group_search<4, 4, false, 17, 9, 1>(win, &sh.best[g], bar, gtid, sh.s_even, 8, rp, (ox >> 2) + sax, (oy >> 2) + say, saw, sah);
                    if (gtid == 0) {
                        sh.sad[0][w][h] = (k >> 17) * 2;
                        sh.x[0][w][h] = (int16_t)(((int)(k & 511) + sax) * 4);
                        sh.y[0][w][h] = (int16_t)(((int)((k >> 9) & 255) + say) * 4);
                    }
                    group_bar(bar);
                }
                if (p.enable_hme_level1) {                        // HmeLevel1, ME.c:4505-4625 (quarter picture)
                    const DPlane& rp = ref.quarter;
                    int saw = round_sa_width(p.hme_l1_w[w]), sah = p.hme_l1_h[h];
                    int sax = -(saw >> 1) + (sh.x[0][w][h] >> 1), say = -(sah >> 1) + (sh.y[0][w][h] >> 1);
                    clip_search_area(ox >> 1, sax, saw, rp.ox - 1, rp.w);
                    clip_search_area(oy >> 1, say, sah, rp.oy - 1, rp.h);
                    const uint32_t k = odd ? group_search_any<13, 7>(win, &sh.best[4 + g], bar, gtid, sh.q_even, kQPitch, sb_w >> 1, 16, rp, (ox >> 1) + sax, (oy >> 1) + say, saw, sah) : half
                        ? group_search<4, kQPitch, false, 13, 7, 2>(win, &sh.best[4 + g], bar, gtid, sh.q_even, 16, rp, (ox >> 1) + sax, (oy >> 1) + say, saw, sah)
                        : group_search<8, kQPitch, true, 13, 7, 2>(win, &sh.best[4 + g], bar, gtid, sh.q_even, 16, rp, (ox >> 1) + sax, (oy >> 1) + say, saw, sah);
                    if (gtid == 0) {
                        sh.sad[1][w][h] = (k >> 13) * 2;
                        sh.x[1][w][h] = (int16_t)(((int)(k & 127) + sax) * 2);
                        sh.y[1][w][h] = (int16_t)(((int)((k >> 7) & 63) + say) * 2);
                    }
                    group_bar(bar);
                }
                if (p.enable_hme_level2) {                        // HmeLevel2, ME.c:4627-4758 (full picture)
                    const DPlane& rp = ref.full;
                    int saw = round_sa_width(p.hme_l2_w[w]), sah = p.hme_l2_h[h];
                    int sax = -(saw >> 1) + sh.x[1][w][h], say = -(sah >> 1) + sh.y[1][w][h];
                    clip_search_area(ox, sax, saw, 63, rp.w);
                    clip_search_area(oy, say, sah, 63, rp.h);
                    const uint32_t k = odd ? group_search_any<13, 7>(win, &sh.best[8 + g], bar, gtid, sh.src_even, kSrcPitch, sb_w, 32, rp, ox + sax, oy + say, saw, sah) : half
                        ? group_search<8, kSrcPitch, false, 13, 7, 8>(win, &sh.best[8 + g], bar, gtid, sh.src_even, 32, rp, ox + sax, oy + say, saw, sah)
                        : group_search<16, kSrcPitch, false, 13, 7, 8>(win, &sh.best[8 + g], bar, gtid, sh.src_even, 32, rp, ox + sax, oy + say, saw, sah);
                    if (gtid == 0) {
                        sh.sad[2][w][h] = (k >> 13) * 2;
                        sh.x[2][w][h] = (int16_t)((int)(k & 127) + sax);
                        sh.y[2][w][h] = (int16_t)((int)((k >> 7) & 63) + say);
                    }
                }
            }
            __syncthreads();
            // What is left is one warp's work: the other seven retire here instead of sitting in four more CTA barriers
            if (warp != 0) return;
            if (list == 0 && p.num_lists == 2 && tid < 12) {
                const int l = tid >> 2, w = (tid >> 1) & 1, h = tid & 1;
                if (w < nw && h < nh) { scr.hme_x[l][w][h] = sh.x[l][w][h]; scr.hme_y[l][w][h] = sh.y[l][w][h]; }
                else { scr.hme_x[l][w][h] = xsc; scr.hme_y[l][w][h] = ysc; }   // regions no group owns keep the list-0 initialisation (Q7)
            }
            int xh = 0, yh = 0;
            if (tid == 0) {
                const int lvl = p.enable_hme_level2 ? 2 : p.enable_hme_level1 ? 1 : p.enable_hme_level0 ? 0 : -1;
                if (lvl >= 0) {
                    xh = sh.x[lvl][0][0]; yh = sh.y[lvl][0][0];
                    uint32_t sh_ = sh.sad[lvl][0][0];
                    for (int h = 0; h < nh; h++) for (int w = (h == 0 ? 1 : 0); w < nw; w++)
                        if (sh.sad[lvl][w][h] < sh_) { xh = sh.x[lvl][w][h]; yh = sh.y[lvl][w][h]; sh_ = sh.sad[lvl][w][h]; }
                    const int tot = nw * nh;
                    if (lvl == 2 && p.same_ref_poc && list == 1 && tot > 1) {           // ME.c:6606-6631
                        for (int q = 0; q < tot - 1; q++) for (int n = q + 1; n < tot; n++) {
                            const int qa = q / nw, qb = q % nw, na = n / nw, nb = n % nw;
                            if (sh.sad[2][qa][qb] > sh.sad[2][na][nb]) {
                                const int tx = sh.x[2][qa][qb], ty = sh.y[2][qa][qb]; const uint32_t ts = sh.sad[2][qa][qb];
                                sh.x[2][qa][qb] = sh.x[2][na][nb]; sh.y[2][qa][qb] = sh.y[2][na][nb]; sh.sad[2][qa][qb] = sh.sad[2][na][nb];
                                sh.x[2][na][nb] = tx; sh.y[2][na][nb] = ty; sh.sad[2][na][nb] = ts;
                            }
                        }
                        xh = sh.x[2][0][1]; yh = sh.y[2][0][1];
                    }
                }
            }
            xsc = __shfl_sync(0xffffffffu, xh, 0); ysc = __shfl_sync(0xffffffffu, yh, 0);
        }
    }
    if (warp != 0) return;                                        // (paths without HME: every thread holds the same xsc / ysc)

    // ---- CheckZeroZeroCenter, ME.c:5466-5552: warp 0 alone. The zero-MV SAD is the centre check's first candidate (a non-zero
    // centre only ever comes out of that check)
    int saw = min(p.search_area_w, 127), sah = min(p.search_area_h, 127);
    if ((xsc != 0 || ysc != 0) && p.is_used_as_reference) {
        xsc = clip_center(ox, xsc, 63, ref.full.w);
        ysc = clip_center(oy, ysc, 63, ref.full.h);
        const uint32_t at_center = warp_sb_sad(sh.src_even, ref.full, ox + xsc, oy + ysc, sb_w >> 2, sb_h >> 1);
        if (sh.cand_sad[0] <= at_center) { xsc = 0; ysc = 0; }
    }
    if (tid == 0) {
        int sax = xsc - (saw >> 1), say = ysc - (sah >> 1);
        clip_search_area(ox, sax, saw, 63, p.width);
        clip_search_area(oy, say, sah, 63, p.height);
        scr.sa_x[list] = sax; scr.sa_y[list] = say; scr.sa_w[list] = saw; scr.sa_h[list] = sah;
        scr.center_x[list] = xsc; scr.center_y[list] = ysc;
    }
}

void launch_hme(const DJob* d_jobs, int n_jobs, int max_sb, int list, cudaStream_t st) {
    hme_kernel<<<dim3(max_sb, n_jobs), kHmeThreads, 0, st>>>(d_jobs, list);
}

}  // namespace svtme
