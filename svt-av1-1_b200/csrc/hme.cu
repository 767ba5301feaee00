// hme.cu -- search-centre derivation for one reference list: hme_mv_center_check, HME level 0/1/2,
// region selection, CheckZeroZeroCenter and the search-area clipping of MotionEstimateLcu
// (reference Source/Lib/Codec/EbMotionEstimation.c:5882-6145, 4306-4758, 6300-6733, 5466-5552).
//
// One CTA (256 threads) per SB. Each search level stages its reference window in shared memory
// (word aligned, chunked by rows when large) and gives every thread one search position at a time;
// an 8-lane group is one of the reference's 8-position SIMD groups, which is what quirk Q3
// (HME level 1: min(lo+hi, 65535 + min_group(lo))) needs. The argmin is a 64-bit key
// (value << 32 | row << 16 | col): smallest value, first in raster order.
#include "me_common.cuh"

namespace svtme {

constexpr int kHmeThreads = 256;
constexpr int kWinBytes = 24 * 1024;

struct HmeShared {
    uint32_t src_even[32 * 16];     // even rows of the 64x64 source SB
    uint32_t q_even[16 * 8];        // even rows of the 32x32 quarter SB
    uint32_t s_even[8 * 4];         // even rows of the 16x16 sixteenth SB
    uint32_t win[kWinBytes / 4];
    unsigned long long red[kHmeThreads / 32];
    unsigned long long result;
    uint32_t cand_sad[8];
};

// SAD of the (sub-sampled) SB against the reference picture at displacement (dx,dy): one warp.
// NxMSadKernel with doubled strides, ME.c:5918-5926.
__device__ uint32_t warp_sb_sad(const uint32_t* src_even, const DPlane& rp, int x, int y, int w_words, int rows) {
    const int lane = threadIdx.x & 31;
    uint32_t acc = 0;
    if (lane < rows) {
        const uint8_t* r = rp.at(x, y + 2 * lane);
        for (int c = 0; c < w_words; c++) acc = sad4(src_even[lane * 16 + c], ldg_u32_unaligned(r + 4 * c), acc);
    }
    return __reduce_add_sync(0xffffffffu, acc);
}

// Exhaustive search of one region. All threads of the CTA call this; returns the winning key in sh.result.
//   src: even rows of the block, 'src_pitch' words per row;  w_words x rows block
//   (rx, ry): top-left search position in level coordinates; saw x sah positions
__device__ void search_region(HmeShared& sh, const uint32_t* src, int src_pitch, int w_words, int rows,
                              const DPlane& rp, int rx, int ry, int saw, int sah, bool l1_quirk) {
    const int tid = threadIdx.x;
    const int sawp = (saw + 7) & ~7;
    const int bx0 = rp.ox + rx;                       // byte column of position 0 inside the allocation
    const int m = bx0 & 3;
    const int pitch = ((m + saw + 4 * w_words - 1 + 3) >> 2) + 1;   // words per window row
    const int max_rows = (kWinBytes / 4) / pitch;
    const int extra = 2 * (rows - 1);
    const int chunk = max(1, max_rows - extra);
    unsigned long long best = ~0ull;
    for (int i0 = 0; i0 < sah; i0 += chunk) {
        const int i1 = min(sah, i0 + chunk);
        const int wrows = i1 - i0 + extra;
        __syncthreads();                              // previous users of sh.win are done
        for (int t = tid; t < wrows * pitch; t += kHmeThreads) {
            const int r = t / pitch, c = t - r * pitch;
            const uint8_t* g = rp.buf + (size_t)(rp.oy + ry + i0 + r) * rp.stride + (bx0 - m) + 4 * c;
            sh.win[t] = __ldg(reinterpret_cast<const uint32_t*>(g));
        }
        __syncthreads();
        const int npos = (i1 - i0) * sawp;
        for (int t = tid; t < ((npos + kHmeThreads - 1) / kHmeThreads) * kHmeThreads; t += kHmeThreads) {
            const int i = t / sawp, j = t - i * sawp;
            const bool valid = t < npos && j < saw;
            uint32_t lo = 0, hi = 0;
            if (valid) {
                const int off = m + j, idx = off >> 2;
                const uint32_t shb = (off & 3) * 8;
                for (int r = 0; r < rows; r++) {
                    const uint32_t* wr = sh.win + (i + 2 * r) * pitch + idx;
                    const uint32_t* sr = src + r * src_pitch;
                    uint32_t a = wr[0];
                    for (int c = 0; c < w_words; c++) {
                        const uint32_t b = wr[c + 1];
                        const uint32_t v = __funnelshift_r(a, b, shb);
                        if (c < 4) lo = sad4(sr[c], v, lo); else hi = sad4(sr[c], v, hi);
                        a = b;
                    }
                }
            }
            uint32_t val;
            if (l1_quirk) {                            // quirk Q3: 16-bit halves + min-subtraction
                uint32_t l = valid ? lo : 65535u, mlo = l;
                mlo = min(mlo, __shfl_xor_sync(0xffffffffu, mlo, 1));
                mlo = min(mlo, __shfl_xor_sync(0xffffffffu, mlo, 2));
                mlo = min(mlo, __shfl_xor_sync(0xffffffffu, mlo, 4));
                val = min(lo + hi, 65535u + mlo);
            } else {
                val = lo + hi;
            }
            if (valid) {
                const unsigned long long key = ((unsigned long long)val << 32) | ((uint32_t)(i0 + i) << 16) | (uint32_t)j;
                best = key < best ? key : best;
            }
        }
    }
    best = warp_min_u64(best);
    if ((tid & 31) == 0) sh.red[tid >> 5] = best;
    __syncthreads();
    if (tid == 0) {
        unsigned long long b = sh.red[0];
        for (int w = 1; w < kHmeThreads / 32; w++) b = sh.red[w] < b ? sh.red[w] : b;
        sh.result = b;
    }
    __syncthreads();
}

__device__ __forceinline__ int round_sa_width(int w) { return (w < 8) ? 8 : (w & 7) ? w + (w - ((w >> 3) << 3)) : w; }

__global__ void __launch_bounds__(kHmeThreads) hme_kernel(const DJob* __restrict__ jobs, int list) {
    __shared__ HmeShared sh;
    __shared__ int s_x[3][2][2], s_y[3][2][2];
    __shared__ unsigned long long s_sad[3][2][2];
    __shared__ int s_center[2];
    const DJob& job = jobs[blockIdx.y];
    const int sb = blockIdx.x;
    if (sb >= job.sb_cols * job.sb_rows) return;
    const SvtMeParams& p = job.prm;
    if (list >= p.num_lists) return;
    const int tid = threadIdx.x, warp = tid >> 5;
    const int ox = (sb % job.sb_cols) * 64, oy = (sb / job.sb_cols) * 64;
    const int sb_w = min(64, p.width - ox), sb_h = min(64, p.height - oy);
    SbScratch& scr = job.scratch[sb];
    const DPyramid& ref = job.ref[list];

    // ---- stage the source SB (even rows only: every SAD on this path is row-subsampled)
    for (int t = tid; t < 32 * 16; t += kHmeThreads)
        sh.src_even[t] = *reinterpret_cast<const uint32_t*>(job.cur.full.at(ox + 4 * (t & 15), oy + 2 * (t >> 4)));
    if (tid < 16 * 8)
        sh.q_even[tid] = *reinterpret_cast<const uint32_t*>(job.cur.quarter.at((ox >> 1) + 4 * (tid & 7), (oy >> 1) + 2 * (tid >> 3)));
    if (tid < 8 * 4)
        sh.s_even[tid] = *reinterpret_cast<const uint32_t*>(job.cur.sixteenth.at((ox >> 2) + 4 * (tid & 3), (oy >> 2) + 2 * (tid >> 2)));
    __syncthreads();

    int xsc = 0, ysc = 0;
    const bool do_center = (p.temporal_layer_index > 0 || list == 0);
    if (do_center) {
        // ---- hme_mv_center_check, ME.c:5882-6145 (position A is evaluated at the zero-MV address: quirk Q2)
        const int tw = p.hme_l0_total_w, th = p.hme_l0_total_h;
        int dxv = 0, dyv = 0;
        if (list == 1) {
            const uint32_t mv0 = scr.best_mv[0][0];
            dxv = (int16_t)(0 - (mv_x(mv0) >> 2)); dyv = (int16_t)(0 - (mv_y(mv0) >> 2));
        }
        if (warp < 5) {
            int cx = 0, cy = 0;
            if (warp == 1) cx = tw; else if (warp == 2) cy = -th; else if (warp == 3) cy = th;
            else if (warp == 4) { cx = dxv; cy = dyv; }
            uint32_t s = 0xFFFFFFFFu;
            if (warp < 4 || list == 1)
                s = warp_sb_sad(sh.src_even, ref.full, ox + clip_center(ox, cx, 63, ref.full.w),
                                oy + clip_center(oy, cy, 63, ref.full.h), sb_w >> 2, sb_h >> 1);
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
            if (tid < 12) {
                const int l = tid >> 2, w = (tid >> 1) & 1, h = tid & 1;
                s_x[l][w][h] = list == 0 ? xsc : scr.hme_x[l][w][h];
                s_y[l][w][h] = list == 0 ? ysc : scr.hme_y[l][w][h];
                s_sad[l][w][h] = 0;
            }
            __syncthreads();
            if (p.enable_hme_level0) {
                const DPlane& rp = ref.sixteenth;
                for (int h = 0; h < nh; h++) for (int w = 0; w < nw; w++) {
                    int saw = (int16_t)((p.hme_l0_w[w] * p.hme_l0_mult_x) / 100);
                    int sah = (int16_t)((p.hme_l0_h[h] * p.hme_l0_mult_y) / 100);
                    int xd = xsc >> 2, yd = ysc >> 2;
                    for (int i = w; i > 0; i--) xd += (int16_t)((p.hme_l0_w[i - 1] * p.hme_l0_mult_x) / 100);
                    for (int i = h; i > 0; i--) yd += (int16_t)((p.hme_l0_h[i - 1] * p.hme_l0_mult_y) / 100);
                    int sax = -(int16_t)(((p.hme_l0_total_w * p.hme_l0_mult_x) / 100) >> 1) + xd;
                    int say = -(int16_t)(((p.hme_l0_total_h * p.hme_l0_mult_y) / 100) >> 1) + yd;
                    clip_search_area(ox >> 2, sax, saw, rp.ox - 1, rp.w);
                    clip_search_area(oy >> 2, say, sah, rp.oy - 1, rp.h);
                    search_region(sh, sh.s_even, 4, sb_w >> 4, sb_h >> 3, rp, (ox >> 2) + sax, (oy >> 2) + say, saw, sah, false);
                    if (tid == 0) {
                        const unsigned long long k = sh.result;
                        s_sad[0][w][h] = (k >> 32) * 2;
                        s_x[0][w][h] = (int16_t)((((int)(k & 0xFFFF)) + sax) * 4);
                        s_y[0][w][h] = (int16_t)((((int)((k >> 16) & 0xFFFF)) + say) * 4);
                    }
                    __syncthreads();
                }
            }
            if (p.enable_hme_level1) {
                const DPlane& rp = ref.quarter;
                for (int h = 0; h < nh; h++) for (int w = 0; w < nw; w++) {
                    int saw = round_sa_width(p.hme_l1_w[w]), sah = p.hme_l1_h[h];
                    int sax = -(saw >> 1) + (s_x[0][w][h] >> 1), say = -(sah >> 1) + (s_y[0][w][h] >> 1);
                    clip_search_area(ox >> 1, sax, saw, rp.ox - 1, rp.w);
                    clip_search_area(oy >> 1, say, sah, rp.oy - 1, rp.h);
                    const int ww = sb_w >> 3, rr = sb_h >> 2;
                    search_region(sh, sh.q_even, 8, ww, rr, rp, (ox >> 1) + sax, (oy >> 1) + say, saw, sah, ww == 8 && rr <= 16);
                    if (tid == 0) {
                        const unsigned long long k = sh.result;
                        s_sad[1][w][h] = (k >> 32) * 2;
                        s_x[1][w][h] = (int16_t)((((int)(k & 0xFFFF)) + sax) * 2);
                        s_y[1][w][h] = (int16_t)((((int)((k >> 16) & 0xFFFF)) + say) * 2);
                    }
                    __syncthreads();
                }
            }
            if (p.enable_hme_level2) {
                const DPlane& rp = ref.full;
                for (int h = 0; h < nh; h++) for (int w = 0; w < nw; w++) {
                    int saw = round_sa_width(p.hme_l2_w[w]), sah = p.hme_l2_h[h];
                    int sax = -(saw >> 1) + s_x[1][w][h], say = -(sah >> 1) + s_y[1][w][h];
                    clip_search_area(ox, sax, saw, 63, rp.w);
                    clip_search_area(oy, say, sah, 63, rp.h);
                    search_region(sh, sh.src_even, 16, sb_w >> 2, sb_h >> 1, rp, ox + sax, oy + say, saw, sah, false);
                    if (tid == 0) {
                        const unsigned long long k = sh.result;
                        s_sad[2][w][h] = (k >> 32) * 2;
                        s_x[2][w][h] = (int16_t)(((int)(k & 0xFFFF)) + sax);
                        s_y[2][w][h] = (int16_t)(((int)((k >> 16) & 0xFFFF)) + say);
                    }
                    __syncthreads();
                }
            }
            if (list == 0 && p.num_lists == 2 && tid < 12) {
                const int l = tid >> 2, w = (tid >> 1) & 1, h = tid & 1;
                scr.hme_x[l][w][h] = s_x[l][w][h]; scr.hme_y[l][w][h] = s_y[l][w][h];
            }
            if (tid == 0) {
                int xh = 0, yh = 0;
                const int lvl = p.enable_hme_level2 ? 2 : p.enable_hme_level1 ? 1 : p.enable_hme_level0 ? 0 : -1;
                if (lvl >= 0) {
                    xh = s_x[lvl][0][0]; yh = s_y[lvl][0][0];
                    unsigned long long sh_ = s_sad[lvl][0][0];
                    for (int h = 0; h < nh; h++) for (int w = (h == 0 ? 1 : 0); w < nw; w++)
                        if (s_sad[lvl][w][h] < sh_) { xh = s_x[lvl][w][h]; yh = s_y[lvl][w][h]; sh_ = s_sad[lvl][w][h]; }
                    const int tot = nw * nh;
                    if (lvl == 2 && p.same_ref_poc && list == 1 && tot > 1) {           // ME.c:6606-6631
                        for (int q = 0; q < tot - 1; q++) for (int n = q + 1; n < tot; n++) {
                            const int qa = q / nw, qb = q % nw, na = n / nw, nb = n % nw;
                            if (s_sad[2][qa][qb] > s_sad[2][na][nb]) {
                                const int tx = s_x[2][qa][qb], ty = s_y[2][qa][qb]; const unsigned long long ts = s_sad[2][qa][qb];
                                s_x[2][qa][qb] = s_x[2][na][nb]; s_y[2][qa][qb] = s_y[2][na][nb]; s_sad[2][qa][qb] = s_sad[2][na][nb];
                                s_x[2][na][nb] = tx; s_y[2][na][nb] = ty; s_sad[2][na][nb] = ts;
                            }
                        }
                        xh = s_x[2][0][1]; yh = s_y[2][0][1];
                    }
                }
                s_center[0] = xh; s_center[1] = yh;
            }
            __syncthreads();
            xsc = s_center[0]; ysc = s_center[1];
        }
    }

    // ---- CheckZeroZeroCenter, ME.c:5466-5552
    int saw = min(p.search_area_w, 127), sah = min(p.search_area_h, 127);
    if ((xsc != 0 || ysc != 0) && p.is_used_as_reference) {
        xsc = clip_center(ox, xsc, 63, ref.full.w);
        ysc = clip_center(oy, ysc, 63, ref.full.h);
        __syncthreads();
        if (warp < 2) {
            const uint32_t s = warp_sb_sad(sh.src_even, ref.full, ox + (warp ? xsc : 0), oy + (warp ? ysc : 0), sb_w >> 2, sb_h >> 1);
            if ((tid & 31) == 0) sh.cand_sad[warp] = s;
        }
        __syncthreads();
        if (sh.cand_sad[0] <= sh.cand_sad[1]) { xsc = 0; ysc = 0; }
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
