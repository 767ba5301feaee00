// subpel.cu -- half-pel and quarter-pel refinement of the 85 square PUs of one SB for one list, and the
// final uni-prediction blocks later consumed by the bi-prediction kernel.
// Reference: InterpolateSearchRegionAVC / HalfPelSearch_LCU / PU_HalfPelRefinement /
// QuarterPelSearch_LCU / SetQuarterPelRefinementInputsOnTheFly / PU_QuarterPelRefinementOnTheFly
// (Source/Lib/Codec/EbMotionEstimation.c:1707-1835, 1842-2415, 2824-3585), with the leaves
// AvcStyleLumaInterpolationFilter{Horizontal,Vertical}_SSSE3 (4-tap {-2,18,18,-2}, +16 >> 5),
// SpatialFullDistortionKernel*_SSSE3 (8-bit wrapping "SSD", quirk Q5), CombinedAveragingSSD and
// CombinedAveraging*SAD_AVX2.
//
// B200 mapping: one CTA per SB keeps four byte planes of the search window in shared memory --
// A (integer), B (half-pel in x), H (half-pel in y), J (vertical filter of the clipped B) -- exactly the
// reference's pos_b/pos_h/pos_j buffers, but addressed by absolute position so that every candidate
// of every PU is a plain unaligned 8x8-block read. Work is split into (PU level, 8x8 block, candidate)
// tasks: distortions are additive over 8x8 blocks, so 64x64..8x8 PUs all share one task shape.
// The wrapping SSD is  dp4a.s8.s8(t,t)  with t = per-byte (a-b) mod 256; the true SSD is dp4a.u8.u8
// of VABSDIFF4; SAD is VABSDIFF4.ACC.
#include "me_common.cuh"

namespace svtme {

constexpr int kSpThreads = 256;
constexpr int kPlRows = 134;            // 64 + 63 + 7
constexpr int kPlPitchW = 36;           // words (144 bytes >= 64 + 70 + one spare word for the last B taps)
constexpr int kPlWords = kPlRows * kPlPitchW;
constexpr int kMargin = 3;              // window starts 3 pixels left/above the search-area origin

struct SpShared {
    uint32_t plane[4][kPlWords];        // A, B, H, J
    uint32_t src[64 * 16];
    uint32_t sum_ssd[9][kNumPu];        // per (candidate, PU) distortion sums, accumulated with shared atomics
    uint32_t sum_sad[9][kNumPu];
    uint32_t pu_sad[kNumPu], pu_mv[kNumPu], pu_ssd[kNumPu];
    int32_t  pu_dir[kNumPu];
    uint32_t pu_imv[kNumPu];            // MV at entry of the current stage
};
constexpr int kSpSmemBytes = (int)sizeof(SpShared);

enum { S_A = 0, S_B = 1, S_H = 2, S_J = 3 };
enum { D_TL = 0, D_T = 1, D_TR = 2, D_R = 3, D_BR = 4, D_B = 5, D_BL = 6, D_L = 7 };   // EbMotionEstimation.h:62-69

// half-pel candidates, order: (integer), L R T B TL TR BR BL
__device__ __constant__ const int8_t kHalf[9][5] = {   // type, dx, dy, mvdx, mvdy
    { S_A, 0, 0, 0, 0 },
    { S_B, -1, 0, -2, 0 }, { S_B, 0, 0, 2, 0 }, { S_H, 0, -1, 0, -2 }, { S_H, 0, 0, 0, 2 },
    { S_J, -1, -1, -2, -2 }, { S_J, 0, -1, 2, -2 }, { S_J, 0, 0, 2, 2 }, { S_J, -1, 0, -2, 2 } };
// quarter-pel: [method][L R T B TL TR BR BL] -> {type1,dx1,dy1,type2,dx2,dy2}   (ME.c:3246-3331)
__device__ __constant__ const int8_t kQuarter[4][8][6] = {
    { { S_B, -1, 0, S_A, 0, 0 }, { S_A, 0, 0, S_B, 0, 0 }, { S_H, 0, -1, S_A, 0, 0 }, { S_A, 0, 0, S_H, 0, 0 },
      { S_B, -1, 0, S_H, 0, -1 }, { S_H, 0, -1, S_B, 0, 0 }, { S_H, 0, 0, S_B, 0, 0 }, { S_B, -1, 0, S_H, 0, 0 } },
    { { S_A, -1, 0, S_B, -1, 0 }, { S_B, -1, 0, S_A, 0, 0 }, { S_J, -1, -1, S_B, -1, 0 }, { S_B, -1, 0, S_J, -1, 0 },
      { S_H, -1, -1, S_B, -1, 0 }, { S_B, -1, 0, S_H, 0, -1 }, { S_B, -1, 0, S_H, 0, 0 }, { S_H, -1, 0, S_B, -1, 0 } },
    { { S_J, -1, -1, S_H, 0, -1 }, { S_H, 0, -1, S_J, 0, -1 }, { S_A, 0, -1, S_H, 0, -1 }, { S_H, 0, -1, S_A, 0, 0 },
      { S_B, -1, -1, S_H, 0, -1 }, { S_H, 0, -1, S_B, 0, -1 }, { S_H, 0, -1, S_B, 0, 0 }, { S_B, -1, 0, S_H, 0, -1 } },
    { { S_H, -1, -1, S_J, -1, -1 }, { S_J, -1, -1, S_H, 0, -1 }, { S_B, -1, -1, S_J, -1, -1 }, { S_J, -1, -1, S_B, -1, 0 },
      { S_H, -1, -1, S_B, -1, -1 }, { S_B, -1, -1, S_H, 0, -1 }, { S_B, -1, 0, S_H, 0, -1 }, { S_H, -1, -1, S_B, -1, 0 } } };
__device__ __constant__ const int8_t kQmv[8][2] = { { -1, 0 }, { 1, 0 }, { 0, -1 }, { 0, 1 }, { -1, -1 }, { 1, -1 }, { 1, 1 }, { -1, 1 } };
// final prediction by fractional position fx + 4*fy: {type1,dx1,dy1,type2(-1 = none),dx2,dy2}   (ME.c:4762-4920)
__device__ __constant__ const int8_t kFrac[16][6] = {
    { S_A, 0, 0, -1, 0, 0 },  { S_A, 0, 0, S_B, 0, 0 }, { S_B, 0, 0, -1, 0, 0 },  { S_B, 0, 0, S_A, 1, 0 },
    { S_A, 0, 0, S_H, 0, 0 }, { S_B, 0, 0, S_H, 0, 0 }, { S_B, 0, 0, S_J, 0, 0 }, { S_B, 0, 0, S_H, 1, 0 },
    { S_H, 0, 0, -1, 0, 0 },  { S_H, 0, 0, S_J, 0, 0 }, { S_J, 0, 0, -1, 0, 0 },  { S_J, 0, 0, S_H, 1, 0 },
    { S_H, 0, 0, S_A, 0, 1 }, { S_H, 0, 0, S_B, 0, 1 }, { S_J, 0, 0, S_B, 0, 1 }, { S_H, 1, 0, S_B, 0, 1 } };

// PU (internal index) that contains 8x8 block (bx,by) at level L (0: 64x64 .. 3: 8x8)
__device__ __forceinline__ int pu_of_block(int L, int bx, int by) {
    if (L == 0) return 0;
    if (L == 1) return 1 + (by >> 2) * 2 + (bx >> 2);
    if (L == 2) return 5 + kTab16[(by >> 1) * 4 + (bx >> 1)];
    return 21 + kTab8[by * 8 + bx];
}
// internal PU index -> level, origin in 8x8-block units, size in blocks
__device__ __forceinline__ void pu_blocks(int pu, int& L, int& bx, int& by, int& nb) {
    if (pu == 0) { L = 0; bx = 0; by = 0; nb = 8; return; }
    if (pu < 5) { const int i = pu - 1; L = 1; bx = (i & 1) * 4; by = (i >> 1) * 4; nb = 4; return; }
    if (pu < 21) {
        int i = 0;
        for (int r = 0; r < 16; r++) if (kTab16[r] == pu - 5) i = r;
        L = 2; bx = (i & 3) * 2; by = (i >> 2) * 2; nb = 2; return;
    }
    int i = 0;
    for (int r = 0; r < 64; r++) if (kTab8[r] == pu - 21) i = r;
    L = 3; bx = i & 7; by = i >> 3; nb = 1;
}

// 8 bytes at byte column c of row r of a plane (unaligned)
__device__ __forceinline__ uint2 load8(const uint32_t* plane, int r, int c) {
    const uint32_t* w = plane + r * kPlPitchW + (c >> 2);
    const uint32_t sh = (c & 3) * 8;
    const uint32_t a = w[0], b = w[1], d = w[2];
    return make_uint2(__funnelshift_r(a, b, sh), __funnelshift_r(b, d, sh));
}
__device__ __forceinline__ uint32_t wrap_ssd4(uint32_t a, uint32_t b, uint32_t acc) {
    const uint32_t t = __vsub4(a, b);                     // per-byte (a-b) mod 256, read as int8: quirk Q5
    return (uint32_t)__dp4a((int)t, (int)t, (int)acc);
}
__device__ __forceinline__ uint32_t true_ssd4(uint32_t a, uint32_t b, uint32_t acc) {
    const uint32_t d = __vabsdiffu4(a, b);
    return __dp4a(d, d, acc);
}
// 4-tap {-2,18,18,-2}, +16 >> 5, clip to u8 on four packed pixels given the four tap words.
// SIMD-in-word on two 16-bit lanes (even bytes / odd bytes): lane = 18*(b+c) - 2*(a+d) + 16 + 256*32 stays in
// [7188, 17388], so nothing crosses a lane; after >> 5 the lane is 256 + (v >> 5), VIMNMX.U16x2 clamps it to
// [256, 511] and the low byte of each lane is the clipped pixel.
__device__ __forceinline__ uint32_t filt4_lanes(uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
    uint32_t v = b * 18u + 0x20102010u;
    v = c * 18u + v;
    v = v - 2u * (a + d);
    v = (v >> 5) & 0x03FF03FFu;
    v = __vmaxu2(v, 0x01000100u);
    return __vminu2(v, 0x01FF01FFu);
}
__device__ __forceinline__ uint32_t filt4(uint32_t t0, uint32_t t1, uint32_t t2, uint32_t t3) {
    const uint32_t e = filt4_lanes(__byte_perm(t0, 0, 0x4240), __byte_perm(t1, 0, 0x4240), __byte_perm(t2, 0, 0x4240), __byte_perm(t3, 0, 0x4240));
    const uint32_t o = filt4_lanes(__byte_perm(t0, 0, 0x4341), __byte_perm(t1, 0, 0x4341), __byte_perm(t2, 0, 0x4341), __byte_perm(t3, 0, 0x4341));
    return __byte_perm(e, o, 0x6240);
}
// validity of quarter-pel candidate k (order L R T B TL TR BR BL) given the best half-pel direction (ME.c:2859-2881):
// the candidate itself or one of its two neighbours in the 8-cycle TL,T,TR,R,BR,B,BL,L; mirrored when the
// half-pel MV has a fractional part (method != 0)
__device__ __forceinline__ bool quarter_valid(int method, int dir, int k) {
    const int q8 = k == 0 ? D_L : k == 1 ? D_R : k == 2 ? D_T : k == 3 ? D_B : k == 4 ? D_TL : k == 5 ? D_TR : k == 6 ? D_BR : D_BL;
    const int centre = method ? ((q8 + 4) & 7) : q8;
    return dir == centre || dir == ((centre + 1) & 7) || dir == ((centre + 7) & 7);
}

__global__ void __launch_bounds__(kSpThreads, 2) subpel_kernel(const DJob* __restrict__ jobs, int list) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    SpShared& sh = *reinterpret_cast<SpShared*>(smem_raw);
    const DJob& job = jobs[blockIdx.y];
    const int sb = blockIdx.x;
    if (sb >= job.sb_cols * job.sb_rows) return;
    const SvtMeParams& p = job.prm;
    if (list >= p.num_lists) return;
    const int tid = threadIdx.x;
    const int ox = (sb % job.sb_cols) * 64, oy = (sb / job.sb_cols) * 64;
    SbScratch& scr = job.scratch[sb];
    const DPlane& rp = job.ref[list].full;
    const int sax = scr.sa_x[list], say = scr.sa_y[list], saw = scr.sa_w[list], sah = scr.sa_h[list];
    const int ww = ((saw + 63 + 7 + 3) >> 2) + 1;        // window words per row actually used (+1: taps of the last B word)
    const int wr = sah + 63 + 7;                          // window rows

    // ---- stage: source SB, PU state, plane A
    for (int t = tid; t < 64 * 16; t += kSpThreads)
        sh.src[t] = *reinterpret_cast<const uint32_t*>(job.cur.full.at(ox + 4 * (t & 15), oy + (t >> 4)));
    if (tid < kNumPu) { sh.pu_sad[tid] = scr.best_sad[list][tid]; sh.pu_mv[tid] = scr.best_mv[list][tid]; sh.pu_imv[tid] = sh.pu_mv[tid]; }
    for (int t = tid; t < wr * ww; t += kSpThreads) {
        const int r = t / ww, c = t - r * ww;
        sh.plane[S_A][r * kPlPitchW + c] = ldg_u32_unaligned(rp.at(ox + sax - kMargin + 4 * c, oy + say - kMargin + r));
    }
    __syncthreads();
    if (p.use_subpel == 1) {
        // ---- planes B (horizontal) and H (vertical) from A; then J (vertical) from B
        for (int t = tid; t < wr * ww; t += kSpThreads) {
            const int r = t / ww, c = t - r * ww;
            const uint32_t* a = sh.plane[S_A] + r * kPlPitchW + c;
            if (c >= 1 && c < ww - 1) {                   // B(x) = f(A[x-1], A[x], A[x+1], A[x+2])
                const uint32_t w0 = a[-1], w1 = a[0], w2 = a[1];
                sh.plane[S_B][r * kPlPitchW + c] = filt4(__funnelshift_r(w0, w1, 24), w1, __funnelshift_r(w1, w2, 8), __funnelshift_r(w1, w2, 16));
            } else if (c == 0) {
                const uint32_t w1 = a[0], w2 = a[1];
                sh.plane[S_B][r * kPlPitchW] = filt4(w1 << 8, w1, __funnelshift_r(w1, w2, 8), __funnelshift_r(w1, w2, 16));
            }
            if (r >= 1 && r < wr - 2)                     // H(y) = f(A[y-1], A[y], A[y+1], A[y+2])
                sh.plane[S_H][r * kPlPitchW + c] = filt4(a[-kPlPitchW], a[0], a[kPlPitchW], a[2 * kPlPitchW]);
        }
        __syncthreads();
        for (int t = tid; t < wr * ww; t += kSpThreads) {
            const int r = t / ww, c = t - r * ww;
            const uint32_t* b = sh.plane[S_B] + r * kPlPitchW + c;
            if (r >= 1 && r < wr - 2)
                sh.plane[S_J][r * kPlPitchW + c] = filt4(b[-kPlPitchW], b[0], b[kPlPitchW], b[2 * kPlPitchW]);
        }
        __syncthreads();

        for (int t = tid; t < 9 * kNumPu; t += kSpThreads) { (&sh.sum_ssd[0][0])[t] = 0; (&sh.sum_sad[0][0])[t] = 0; }
        __syncthreads();
        // ---- half-pel: (level, block, candidate) tasks -> wrapping-SSD and SAD summed per (candidate, PU)
        for (int t = tid; t < 4 * 64 * 9; t += kSpThreads) {
            const int k = t % 9, blk = (t / 9) & 63, L = t / (9 * 64);
            const int bx = blk & 7, by = blk >> 3;
            const int pu = pu_of_block(L, bx, by);
            const uint32_t mv = sh.pu_imv[pu];
            const int X = kMargin + (mv_x(mv) >> 2) - sax + bx * 8 + kHalf[k][1];
            const int Y = kMargin + (mv_y(mv) >> 2) - say + by * 8 + kHalf[k][2];
            const uint32_t* pl = sh.plane[kHalf[k][0]];
            uint32_t ssd = 0, sad = 0;
#pragma unroll
            for (int r = 0; r < 8; r++) {
                const uint2 c = load8(pl, Y + r, X);
                const uint32_t s0 = sh.src[(by * 8 + r) * 16 + bx * 2], s1 = sh.src[(by * 8 + r) * 16 + bx * 2 + 1];
                ssd = wrap_ssd4(s0, c.x, ssd); ssd = wrap_ssd4(s1, c.y, ssd);
                sad = sad4(s0, c.x, sad); sad = sad4(s1, c.y, sad);
            }
            atomicAdd(&sh.sum_ssd[k][pu], ssd); atomicAdd(&sh.sum_sad[k][pu], sad);
        }
        __syncthreads();
        // ---- per PU: nine candidates in the reference's order, strict '<' on the SSD  (ME.c:1909-2239)
        if (tid < kNumPu && (tid < 21 || p.cu8x8_mode != 1)) {
            uint32_t dist[9], sadv[9];
#pragma unroll
            for (int k = 0; k < 9; k++) { dist[k] = sh.sum_ssd[k][tid]; sadv[k] = sh.sum_sad[k][tid]; }
            const uint32_t imv = sh.pu_imv[tid];
            uint32_t bssd = dist[0], bsad = sh.pu_sad[tid], bmv = imv;
            for (int k = 1; k < 9; k++)
                if (dist[k] < bssd) { bssd = dist[k]; bsad = sadv[k]; bmv = make_mv((int16_t)(mv_x(imv) + kHalf[k][3]), (int16_t)(mv_y(imv) + kHalf[k][4])); }
            uint32_t best = dist[1];
            for (int k = 2; k < 9; k++) best = min(best, dist[k]);
            // tie order L R T B TL TR BL BR
            const int dir = best == dist[1] ? D_L : best == dist[2] ? D_R : best == dist[3] ? D_T : best == dist[4] ? D_B :
                            best == dist[5] ? D_TL : best == dist[6] ? D_TR : best == dist[8] ? D_BL : D_BR;
            sh.pu_ssd[tid] = bssd; sh.pu_sad[tid] = bsad; sh.pu_mv[tid] = bmv; sh.pu_dir[tid] = dir;
        }
        __syncthreads();
        if (tid < kNumPu) sh.pu_imv[tid] = sh.pu_mv[tid];
        for (int t = tid; t < 9 * kNumPu; t += kSpThreads) { (&sh.sum_ssd[0][0])[t] = 0; (&sh.sum_sad[0][0])[t] = 0; }
        __syncthreads();

        // ---- quarter-pel: (level, block, candidate) tasks -> partial true-SSD and SAD of the averaged pair
        // (exactly three of the eight positions are valid for a given half-pel direction: task index j picks one)
        for (int t = tid; t < 4 * 64 * 3; t += kSpThreads) {
            const int j = t % 3, blk = (t / 3) & 63, L = t / 192;
            const int bx = blk & 7, by = blk >> 3;
            uint32_t ssd = 0, sad = 0;
            const int pu = pu_of_block(L, bx, by);
            const bool in_area = (L != 0) || (bx < 4 && by < 4);          // quirk Q6: 64x64 refined on its top-left 32x32
            if (in_area && (L != 3 || p.cu8x8_mode != 1)) {
                const uint32_t mv = sh.pu_imv[pu];
                const int xm = mv_x(mv), ym = mv_y(mv), dir = sh.pu_dir[pu];
                const int method = (ym & 2) + ((xm & 2) >> 1);
                int k = 0, cnt = 0;
#pragma unroll
                for (int kk = 0; kk < 8; kk++)
                    if (quarter_valid(method, dir, kk)) { if (cnt == j) k = kk; cnt++; }
                {
                    const int8_t* q = kQuarter[method][k];
                    const int X = kMargin + ((xm + 2) >> 2) - sax + bx * 8, Y = kMargin + ((ym + 2) >> 2) - say + by * 8;
                    const uint32_t* p1 = sh.plane[q[0]]; const uint32_t* p2 = sh.plane[q[3]];
#pragma unroll
                    for (int r = 0; r < 8; r++) {
                        const uint2 c1 = load8(p1, Y + q[2] + r, X + q[1]), c2 = load8(p2, Y + q[5] + r, X + q[4]);
                        const uint32_t a0 = __vavgu4(c1.x, c2.x), a1 = __vavgu4(c1.y, c2.y);
                        const uint32_t s0 = sh.src[(by * 8 + r) * 16 + bx * 2], s1 = sh.src[(by * 8 + r) * 16 + bx * 2 + 1];
                        ssd = true_ssd4(s0, a0, ssd); ssd = true_ssd4(s1, a1, ssd);
                        sad = sad4(s0, a0, sad); sad = sad4(s1, a1, sad);
                    }
                    atomicAdd(&sh.sum_ssd[k][pu], ssd); atomicAdd(&sh.sum_sad[k][pu], sad);
                }
            }
        }
        __syncthreads();
        if (tid < kNumPu && (tid < 21 || p.cu8x8_mode != 1)) {
            const uint32_t imv = sh.pu_imv[tid];
            const int method = (mv_y(imv) & 2) + ((mv_x(imv) & 2) >> 1), dir = sh.pu_dir[tid];
            uint32_t bssd = sh.pu_ssd[tid], bsad = sh.pu_sad[tid], bmv = imv;
#pragma unroll
            for (int k = 0; k < 8; k++) {
                if (!quarter_valid(method, dir, k)) continue;
                const uint32_t a = sh.sum_ssd[k][tid], b = sh.sum_sad[k][tid];
                if (a < bssd) { bssd = a; bsad = b; bmv = make_mv((int16_t)(mv_x(imv) + kQmv[k][0]), (int16_t)(mv_y(imv) + kQmv[k][1])); }
            }
            sh.pu_sad[tid] = bsad; sh.pu_mv[tid] = bmv;
        }
        __syncthreads();
        if (tid < kNumPu) { scr.best_sad[list][tid] = sh.pu_sad[tid]; scr.best_mv[list][tid] = sh.pu_mv[tid]; }
    } else {
        // planes B/H/J are not needed: integer MVs only use plane A below
    }

    // ---- final uni-prediction blocks of this list for every PU level (consumed by bipred_pack_kernel)
    if (p.num_lists == 2) {
        for (int t = tid; t < 4 * 64; t += kSpThreads) {
            const int blk = t & 63, L = t >> 6, bx = blk & 7, by = blk >> 3;
            const uint32_t mv = sh.pu_mv[pu_of_block(L, bx, by)];
            const int xm = mv_x(mv), ym = mv_y(mv);
            const int8_t* f = kFrac[(xm & 3) + ((ym & 3) << 2)];
            const int X = kMargin + (xm >> 2) - sax + bx * 8, Y = kMargin + (ym >> 2) - say + by * 8;
            uint32_t* dst = reinterpret_cast<uint32_t*>(scr.pred[list][L]) + (by * 8) * 16 + bx * 2;
#pragma unroll
            for (int r = 0; r < 8; r++) {
                uint2 c = load8(sh.plane[f[0]], Y + f[2] + r, X + f[1]);
                if (f[3] >= 0) {
                    const uint2 d = load8(sh.plane[f[3]], Y + f[5] + r, X + f[4]);
                    c.x = __vavgu4(c.x, d.x); c.y = __vavgu4(c.y, d.y);
                }
                dst[r * 16] = c.x; dst[r * 16 + 1] = c.y;
            }
        }
    }
}

cudaError_t configure_subpel() {
    return cudaFuncSetAttribute(subpel_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSpSmemBytes);
}

void launch_subpel(const DJob* d_jobs, int n_jobs, int max_sb, int list, cudaStream_t st) {
    subpel_kernel<<<dim3(max_sb, n_jobs), kSpThreads, kSpSmemBytes, st>>>(d_jobs, list);
}

}  // namespace svtme
