// fullpel_ext.cu -- integer full-pel search of one 64x64 SB against one reference list in the 209-PU mode
// (PIC_ALL_DEPTH_MODE pictures, i.e. enc-mode 0-2 reference pictures): SADs of the 85 square PUs AND of the 124
// non-square ones (64x32, 32x16, 16x8, 32x64, 16x32, 8x16, 32x8, 8x32, 64x16, 16x64) at every position of the
// search area, and the per-PU argmin. Bit-exact with open_loop_me_fullpel_search_sblock and its leaves
// (reference Source/Lib/Codec/EbMotionEstimation.c:159-1236, 1556-1595,
//  Source/Lib/ASM_SSE4_1/EbComputeSAD_Intrinsic_SSE4_1.c:21-164).
//
// Unlike the 85-PU path the reference visits ONE position at a time in raster order and every update is a plain
// strict '<', so every PU is "smallest SAD, first in raster order" = the minimum of key = sad * 4096 + (row << 6 | col)
// -- with one exception, quirk Q11: the update test of 32x16 PU #5 (x 0, y 48) reads a stale variable holding the
// bottom 64x32 SAD but stores its own SAD. That PU is a sequential state machine over the positions
//     best <- T(pos)  whenever  S(pos) < best          (S = bottom 64x32 SAD >= T = the PU's own SAD)
// which the kernel replays after the search from per-position S / T tables kept in shared memory.
//
// B200 mapping: same as fullpel.cu -- four byte-shifted copies of the reference window in shared memory, one thread =
// 4 positions x, x+4, x+8, x+12 of one row, VABSDIFF4.U8.ACC on the ALU pipe, keys built with IMAD on the FMA pipe and
// reduced across the warp by streaming transpose-reductions (pair_min) so that at the end of a tile every lane issues
// one RED.SHARED.MIN per PU class. The non-square SADs are sums of the 8x8 SADs: horizontal sums inside a band of
// eight rows, vertical sums carried across bands in registers (this is what costs the registers: 1 CTA per SM).
// Search areas beyond 64x64 (template parameter Wide, MIN(search_area, 127), ME.c:6644-6645) are cut into sub-areas of at
// most 64x64 positions exactly like fullpel.cu and merged per PU by (SAD, row, column); the S / T tables of quirk Q11
// then cover 128 x 128 positions and are replayed once after the last sub-area.
#include "me_common.cuh"
#include <cstdlib>

namespace svtme {
namespace {

constexpr int kWinRows = 126;          // 64 search rows + 62
constexpr int kWinPitch = 36;          // words; = 4 (mod 32): the 8 row-distinct LDS.128 of a warp are conflict-free
constexpr int kWinWords = 4 * kWinRows * kWinPitch;
constexpr int kSrcWords = 32 * 16;
constexpr int kKeyWords = 212;         // 209 PU slots (internal index), padded
constexpr int kRawPitch = 40;          // words per raw window row (ten 16-byte granules), see fullpel.cu
constexpr int kFxThreads = 256;        // x 1 CTA per SM (255 registers)
template <bool Wide> struct Q11 { static constexpr int kShift = Wide ? 7 : 6, kEntries = 1 << (2 * kShift); };
template <bool Wide> constexpr int fx_smem_bytes() {
    return (kWinWords + kSrcWords + kKeyWords + Q11<Wide>::kEntries) * 4 + Q11<Wide>::kEntries * 2 + kWinRows * kRawPitch * 4;
}

// see fullpel.cu: lanes whose `bit` is clear keep min(a) over {lane, lane ^ bit}, the others the same for b
__device__ __forceinline__ uint32_t pair_min(uint32_t a, uint32_t b, int bit, uint32_t hi01) {
    uint32_t d, keep, send;
    asm("sub.u32 %0, %1, %2;" : "=r"(d) : "r"(b), "r"(a));
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(keep) : "r"(hi01), "r"(d), "r"(a));
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(send) : "r"(0u - hi01), "r"(d), "r"(b));
    return min(keep, __shfl_xor_sync(0xffffffffu, send, bit));
}
__device__ __forceinline__ uint32_t shfl_min(uint32_t v, int bit) { return min(v, __shfl_xor_sync(0xffffffffu, v, bit)); }
__device__ __forceinline__ void red_min(uint32_t slot_addr, uint32_t key) {
    asm volatile("red.shared.min.u32 [%0], %1;" :: "r"(slot_addr), "r"(key) : "memory");
}

struct Keyer {                       // key = sad * 4096 + (row << 6 | col) as one IMAD per position; dead positions give ~0
    uint32_t mul[4], add[4];
    __device__ __forceinline__ uint32_t operator()(uint32_t a, uint32_t b, uint32_t c, uint32_t d) const {
        return min(min(a * mul[0] + add[0], b * mul[1] + add[1]), min(c * mul[2] + add[2], d * mul[3] + add[3]));
    }
};

}  // namespace

namespace {
// per-lane PU slot tables: which PU (internal index) a lane owns after the transpose-reductions of each class
enum { T8A = 0, T8B, T16x8, T32x8, T8x16, T16x16, T32x16, T64x16, T8x32, T16x32, T32x32, T64x32, T16x64, T32x64, kNumTables };

__device__ __forceinline__ int lane_pu(int table, int lane) {
    const int l16 = (lane >> 4) & 1, l8 = (lane >> 3) & 1, l4 = (lane >> 2) & 1, l2 = (lane >> 1) & 1, l1 = lane & 1;
    const int j = l16 + 2 * l8 + 4 * l4, q = l16 + 2 * l8, p = l16;
    switch (table) {
        case T8A:    return kInternal[21 + (l2 + 2 * l1) * 8 + j];
        case T8B:    return kInternal[21 + (4 + l2 + 2 * l1) * 8 + j];
        case T16x8:  return kInternal[95 + (l4 + 2 * l2 + 4 * l1) * 4 + q];
        case T32x8:  return kInternal[169 + (l8 + 2 * l4 + 4 * l2) * 2 + p];
        case T8x16:  return kInternal[137 + (l2 + 2 * l1) * 8 + j];
        case T16x16: return kInternal[5 + (l4 + 2 * l2) * 4 + q];
        case T32x16: return kInternal[87 + (l8 + 2 * l4) * 2 + p];
        case T64x16: return 201 + l16 + 2 * l8;
        case T8x32:  return kInternal[185 + l2 * 8 + j];
        case T16x32: return kInternal[129 + l4 * 4 + q];
        case T32x32: return 1 + l8 * 2 + p;
        case T64x32: return 85 + l16;
        case T16x64: return 205 + q;
        default:     return 127 + p;      // T32x64
    }
}
}  // namespace

template <bool Wide>
__global__ void __launch_bounds__(kFxThreads, 1) fullpel_ext_kernel(const DJob* __restrict__ jobs, int list) {
    constexpr int QS = Q11<Wide>::kShift, QN = Q11<Wide>::kEntries;
    extern __shared__ __align__(16) uint32_t smem[];
    uint32_t* win = smem;
    uint32_t* src = win + kWinWords;
    uint32_t* keys = src + kSrcWords;
    uint32_t* q11_s = keys + kKeyWords;                                  // bottom 64x32 SAD of every position
    uint16_t* q11_t = reinterpret_cast<uint16_t*>(q11_s + QN);           // 32x16 (x 0, y 48) SAD of every position
    uint32_t* raw = reinterpret_cast<uint32_t*>(q11_t + QN);             // raw window rows (byte offset 100048 = 16 * 6253; Wide 173776 = 16 * 10861)
    __shared__ uint32_t lane_slot[kNumTables][32];                       // shared ADDRESS of the PU slot a lane owns
    __shared__ uint32_t merged_v[Wide ? kMaxPu : 1], merged_o[Wide ? kMaxPu : 1];   // Wide: running winner per PU (value, row << 8 | column)
    const DJob& job = jobs[blockIdx.y];
    const int sb = job.sb_begin + blockIdx.x;
    if (sb >= job.sb_end) return;
    const SvtMeParams& p = job.prm;
    if (list >= p.num_lists || p.num_pus != kMaxPu || wide_search(p) != Wide) return;
    const int tid = threadIdx.x;
    const int ox = (sb % job.sb_cols) * 64, oy = (sb / job.sb_cols) * 64;
    SbScratch& scr = job.scratch[sb];
    SbScratchExt& scx = job.scratch_ext[sb];
    const DPlane& rp = job.ref[list].full;
    const int area_x = scr.sa_x[list], area_y = scr.sa_y[list], area_w = scr.sa_w[list], area_h = scr.sa_h[list];
    const uint32_t keys_addr = (uint32_t)__cvta_generic_to_shared(keys);
    if (Wide) for (int t = tid; t < kMaxPu; t += kFxThreads) { merged_v[t] = 0xFFFFFFFFu; merged_o[t] = 0xFFFFFFFFu; }
#pragma unroll 1
    for (int sub = 0; sub < (Wide ? 4 : 1); sub++) {
    const int tx0 = Wide ? (sub & 1) * 64 : 0, ty0 = Wide ? (sub >> 1) * 64 : 0;
    if (Wide && (tx0 >= area_w || ty0 >= area_h)) continue;       // CTA-uniform
    const int sax = area_x + tx0, say = area_y + ty0, saw = Wide ? min(64, area_w - tx0) : area_w, sah = Wide ? min(64, area_h - ty0) : area_h;

    // ---- stage source (even rows), the slot tables and the four byte-shifted window copies (as in fullpel.cu: the window
    // lands raw in shared memory as 16-byte cp.async granules, then the warps shift it into the four copies)
    const int wrows = sah + 62;
    const int nfill = min(32, ((saw + 62 + 3) >> 2) + 1);
    const uint8_t* g0 = rp.at(ox + sax, oy + say);
    const uint32_t m16 = (uint32_t)(uintptr_t)g0 & 15u;
    {
        const int ngran = ((int)(m16 >> 2) + nfill + 2 + 3) >> 2;
        const uint8_t* ga = g0 - m16;
        const uint32_t raw_addr = (uint32_t)__cvta_generic_to_shared(raw);
        for (int t = tid; t < wrows * ngran; t += kFxThreads) {
            const int r = t / ngran, g = t - r * ngran;
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"(raw_addr + (uint32_t)(r * kRawPitch * 4 + g * 16)),
                         "l"(ga + (size_t)r * rp.stride + g * 16) : "memory");
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    }
    if (!Wide || sub == 0) {
        for (int t = tid; t < kSrcWords; t += kFxThreads)
            src[t] = *reinterpret_cast<const uint32_t*>(job.cur.full.at(ox + 4 * (t & 15), oy + 2 * (t >> 4)));
        for (int t = tid; t < kNumTables * 32; t += kFxThreads) lane_slot[t >> 5][t & 31] = keys_addr + 4u * (uint32_t)lane_pu(t >> 5, t & 31);
    }
    for (int t = tid; t < kKeyWords; t += kFxThreads) keys[t] = 0xFFFFFFFFu;
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();
    {
        const int c = tid & 31;
        if (c < nfill) {
            const uint32_t m = m16 & 3u;
            const uint32_t* rw = raw + (m16 >> 2) + c;
            for (int r = tid >> 5; r < wrows; r += kFxThreads / 32) {
                const uint32_t* w = rw + r * kRawPitch;
                const uint32_t w0 = w[0], w1 = w[1], w2 = w[2];
#pragma unroll
                for (int sc = 0; sc < 4; sc++) {
                    const uint32_t o = m + sc;
                    const uint32_t lo = o < 4 ? w0 : w1, hi = o < 4 ? w1 : w2;
                    win[(sc * kWinRows + r) * kWinPitch + c] = __funnelshift_r(lo, hi, (o & 3u) * 8u);
                }
            }
        }
    }
    __syncthreads();

    const int ngp = (saw + 15) >> 4;                   // 16-position groups per row
    const int total = sah * ngp * 4;
    const int s = tid & 3, lane = tid & 31;
    const uint32_t l16 = (lane >> 4) & 1, l8 = (lane >> 3) & 1, l4 = (lane >> 2) & 1, l2 = (lane >> 1) & 1, l1 = lane & 1;
#define SLOT(t) lane_slot[t][lane]

    for (int u0 = 0; u0 < total; u0 += kFxThreads) {
        const int u = u0 + tid;
        const bool live = u < total;
        int y = 0, x0 = s;
        if (live) { const int q = u >> 2; const int G = q % ngp; y = q / ngp; x0 = 16 * G + s; }
        Keyer K;
        bool val[4];
#pragma unroll
        for (int k = 0; k < 4; k++) {
            val[k] = live && (x0 + 4 * k < saw);
            K.mul[k] = val[k] ? 4096u : 0u;
            K.add[k] = val[k] ? (uint32_t)((y << 6) | (x0 + 4 * k)) : 0xFFFFFFFFu;
        }
        uint32_t acc[2][4][8];                          // 8x8 SADs of the current band [b & 1] and of the band before
        uint32_t h16[4][4], v16[4][4], w16[4][4];       // 16x16 (band pair), 16x32 (half), 16x64 column sums
        uint32_t c32[4][8];                             // 8x32 column sums of the current half
        uint32_t s32[4][2], w32[4][2], s64h[4], s64[4]; // 32x32 of the half, 32x64, 64x32 of the half, 64x64
        // carries of the streaming transpose-reductions, one chain per PU class
        uint32_t c8a = 0, c8b = 0, c168a = 0, c168b = 0, c168c = 0, c328a = 0, c328b = 0, c328c = 0;
        uint32_t c816a = 0, c816b = 0, c1616a = 0, c1616b = 0, c3216a = 0, c3216b = 0, c6416a = 0, c6416b = 0;
        uint32_t c832 = 0, c1632 = 0, c3232 = 0, c6432 = 0;
#pragma unroll
        for (int b = 0; b < 8; b++) {
            uint32_t (&A)[4][8] = acc[b & 1];
#pragma unroll
            for (int k = 0; k < 4; k++)
#pragma unroll
                for (int j = 0; j < 8; j++) A[k][j] = 0;
#pragma unroll
            for (int rr = 0; rr < 4; rr++) {
                const int r = 4 * b + rr;
                const uint4* wp = reinterpret_cast<const uint4*>(win + (s * kWinRows + y + 2 * r) * kWinPitch + (x0 >> 2));
                uint32_t ref[20];
#pragma unroll
                for (int i = 0; i < 5; i++) {
                    const uint4 v = wp[i];
                    ref[4 * i] = v.x; ref[4 * i + 1] = v.y; ref[4 * i + 2] = v.z; ref[4 * i + 3] = v.w;
                }
                const uint4* sp = reinterpret_cast<const uint4*>(src + r * 16);
#pragma unroll
                for (int cc = 0; cc < 4; cc++) {
                    const uint4 sv = sp[cc];
                    const uint32_t sw[4] = { sv.x, sv.y, sv.z, sv.w };
#pragma unroll
                    for (int j = 0; j < 4; j++) {
                        const int c = 4 * cc + j;
#pragma unroll
                        for (int k = 0; k < 4; k++) A[k][c >> 1] = sad4(sw[j], ref[c + k], A[k][c >> 1]);
                    }
                }
            }
            // ---- band b (rows 8b .. 8b+7): 8x8, 16x8, 32x8 PUs of this band
            uint32_t h[4][4], g[4][2];
#pragma unroll
            for (int k = 0; k < 4; k++) {
#pragma unroll
                for (int q = 0; q < 4; q++) h[k][q] = A[k][2 * q] + A[k][2 * q + 1];
                g[k][0] = h[k][0] + h[k][1]; g[k][1] = h[k][2] + h[k][3];
            }
            {   // 8x8: lanes end up with column (l16,l8,l4) and band-in-half (l2,l1)
                uint32_t k8[8];
#pragma unroll
                for (int j = 0; j < 8; j++) k8[j] = K(A[0][j], A[1][j], A[2][j], A[3][j]);
                const uint32_t a0 = pair_min(k8[0], k8[1], 16, l16), a1 = pair_min(k8[2], k8[3], 16, l16);
                const uint32_t a2 = pair_min(k8[4], k8[5], 16, l16), a3 = pair_min(k8[6], k8[7], 16, l16);
                const uint32_t t3 = pair_min(pair_min(a0, a1, 8, l8), pair_min(a2, a3, 8, l8), 4, l4);
                if ((b & 1) == 0) c8a = t3;
                else {
                    const uint32_t t4 = pair_min(c8a, t3, 2, l2);
                    if ((b & 3) == 1) c8b = t4;
                    else red_min(b == 3 ? SLOT(T8A) : SLOT(T8B), pair_min(c8b, t4, 1, l1));
                }
            }
            {   // 16x8: column (l16,l8), band (l4,l2,l1)
                uint32_t kq[4];
#pragma unroll
                for (int q = 0; q < 4; q++) kq[q] = K(h[0][q], h[1][q], h[2][q], h[3][q]);
                const uint32_t r2 = pair_min(pair_min(kq[0], kq[1], 16, l16), pair_min(kq[2], kq[3], 16, l16), 8, l8);
                if ((b & 1) == 0) c168a = r2;
                else {
                    const uint32_t r3 = pair_min(c168a, r2, 4, l4);
                    if ((b & 3) == 1) c168b = r3;
                    else {
                        const uint32_t r4 = pair_min(c168b, r3, 2, l2);
                        if (b == 3) c168c = r4; else red_min(SLOT(T16x8), pair_min(c168c, r4, 1, l1));
                    }
                }
            }
            {   // 32x8: column l16, band (l8,l4,l2); bit 1 by shuffle
                const uint32_t r1 = pair_min(K(g[0][0], g[1][0], g[2][0], g[3][0]), K(g[0][1], g[1][1], g[2][1], g[3][1]), 16, l16);
                if ((b & 1) == 0) c328a = r1;
                else {
                    const uint32_t r2 = pair_min(c328a, r1, 8, l8);
                    if ((b & 3) == 1) c328b = r2;
                    else {
                        const uint32_t r3 = pair_min(c328b, r2, 4, l4);
                        if (b == 3) c328c = r3;
                        else { const uint32_t f = shfl_min(pair_min(c328c, r3, 2, l2), 1); if (!l1) red_min(SLOT(T32x8), f); }
                    }
                }
            }
            if ((b & 1) == 0) {
#pragma unroll
                for (int k = 0; k < 4; k++)
#pragma unroll
                    for (int q = 0; q < 4; q++) h16[k][q] = h[k][q];
                continue;
            }
            // ---- band pair (rows 16*pair .. +15): 8x16, 16x16, 32x16, 64x16
            const int pair = b >> 1;
            uint32_t (&E)[4][8] = acc[0];               // the even band of the pair (A is acc[1] here)
            uint32_t m32[4][2];
#pragma unroll
            for (int k = 0; k < 4; k++) {
#pragma unroll
                for (int j = 0; j < 8; j++) E[k][j] += A[k][j];          // E becomes the 8x16 column sums
#pragma unroll
                for (int q = 0; q < 4; q++) h16[k][q] += h[k][q];
                m32[k][0] = h16[k][0] + h16[k][1]; m32[k][1] = h16[k][2] + h16[k][3];
            }
            {   // 8x16: column (l16,l8,l4), pair (l2,l1)
                uint32_t k8[8];
#pragma unroll
                for (int j = 0; j < 8; j++) k8[j] = K(E[0][j], E[1][j], E[2][j], E[3][j]);
                const uint32_t a0 = pair_min(k8[0], k8[1], 16, l16), a1 = pair_min(k8[2], k8[3], 16, l16);
                const uint32_t a2 = pair_min(k8[4], k8[5], 16, l16), a3 = pair_min(k8[6], k8[7], 16, l16);
                const uint32_t t3 = pair_min(pair_min(a0, a1, 8, l8), pair_min(a2, a3, 8, l8), 4, l4);
                if ((pair & 1) == 0) c816a = t3;
                else {
                    const uint32_t t4 = pair_min(c816a, t3, 2, l2);
                    if (pair == 1) c816b = t4; else red_min(SLOT(T8x16), pair_min(c816b, t4, 1, l1));
                }
            }
            {   // 16x16: column (l16,l8), pair (l4,l2); bit 1 by shuffle
                uint32_t kq[4];
#pragma unroll
                for (int q = 0; q < 4; q++) kq[q] = K(h16[0][q], h16[1][q], h16[2][q], h16[3][q]);
                const uint32_t r2 = pair_min(pair_min(kq[0], kq[1], 16, l16), pair_min(kq[2], kq[3], 16, l16), 8, l8);
                if ((pair & 1) == 0) c1616a = r2;
                else {
                    const uint32_t r3 = pair_min(c1616a, r2, 4, l4);
                    if (pair == 1) c1616b = r3;
                    else { const uint32_t f = shfl_min(pair_min(c1616b, r3, 2, l2), 1); if (!l1) red_min(SLOT(T16x16), f); }
                }
            }
            {   // 32x16: column l16, pair (l8,l4); bits 2,1 by shuffle. The slot of PU #5 (x 0, y 48) is overwritten by the
                // replay of quirk Q11 after the search.
                const uint32_t r1 = pair_min(K(m32[0][0], m32[1][0], m32[2][0], m32[3][0]), K(m32[0][1], m32[1][1], m32[2][1], m32[3][1]), 16, l16);
                if ((pair & 1) == 0) c3216a = r1;
                else {
                    const uint32_t r2 = pair_min(c3216a, r1, 8, l8);
                    if (pair == 1) c3216b = r2;
                    else { const uint32_t f = shfl_min(shfl_min(pair_min(c3216b, r2, 4, l4), 2), 1); if ((lane & 3) == 0) red_min(SLOT(T32x16), f); }
                }
            }
            {   // 64x16: pair (l16,l8); bits 4,2,1 by shuffle
                const uint32_t r0 = K(m32[0][0] + m32[0][1], m32[1][0] + m32[1][1], m32[2][0] + m32[2][1], m32[3][0] + m32[3][1]);
                if ((pair & 1) == 0) c6416a = r0;
                else {
                    const uint32_t r1 = pair_min(c6416a, r0, 16, l16);
                    if (pair == 1) c6416b = r1;
                    else { const uint32_t f = shfl_min(shfl_min(shfl_min(pair_min(c6416b, r1, 8, l8), 4), 2), 1); if ((lane & 7) == 0) red_min(SLOT(T64x16), f); }
                }
            }
            if ((pair & 1) == 0) {
#pragma unroll
                for (int k = 0; k < 4; k++) {
#pragma unroll
                    for (int j = 0; j < 8; j++) c32[k][j] = E[k][j];
#pragma unroll
                    for (int q = 0; q < 4; q++) v16[k][q] = h16[k][q];
                }
                continue;
            }
            // ---- half (rows 32*half .. +31): 8x32, 16x32, 32x32, 64x32
            const int half = b >> 2;
#pragma unroll
            for (int k = 0; k < 4; k++) {
#pragma unroll
                for (int j = 0; j < 8; j++) c32[k][j] += E[k][j];
#pragma unroll
                for (int q = 0; q < 4; q++) v16[k][q] += h16[k][q];
                s32[k][0] = v16[k][0] + v16[k][1]; s32[k][1] = v16[k][2] + v16[k][3];
                s64h[k] = s32[k][0] + s32[k][1];
            }
            {   // 8x32: column (l16,l8,l4), half l2; bit 1 by shuffle
                uint32_t k8[8];
#pragma unroll
                for (int j = 0; j < 8; j++) k8[j] = K(c32[0][j], c32[1][j], c32[2][j], c32[3][j]);
                const uint32_t a0 = pair_min(k8[0], k8[1], 16, l16), a1 = pair_min(k8[2], k8[3], 16, l16);
                const uint32_t a2 = pair_min(k8[4], k8[5], 16, l16), a3 = pair_min(k8[6], k8[7], 16, l16);
                const uint32_t t3 = pair_min(pair_min(a0, a1, 8, l8), pair_min(a2, a3, 8, l8), 4, l4);
                if (half == 0) c832 = t3;
                else { const uint32_t f = shfl_min(pair_min(c832, t3, 2, l2), 1); if (!l1) red_min(SLOT(T8x32), f); }
            }
            {   // 16x32: column (l16,l8), half l4; bits 2,1 by shuffle
                uint32_t kq[4];
#pragma unroll
                for (int q = 0; q < 4; q++) kq[q] = K(v16[0][q], v16[1][q], v16[2][q], v16[3][q]);
                const uint32_t r2 = pair_min(pair_min(kq[0], kq[1], 16, l16), pair_min(kq[2], kq[3], 16, l16), 8, l8);
                if (half == 0) c1632 = r2;
                else { const uint32_t f = shfl_min(shfl_min(pair_min(c1632, r2, 4, l4), 2), 1); if ((lane & 3) == 0) red_min(SLOT(T16x32), f); }
            }
            {   // 32x32: column l16, half l8; bits 4,2,1 by shuffle
                const uint32_t r1 = pair_min(K(s32[0][0], s32[1][0], s32[2][0], s32[3][0]), K(s32[0][1], s32[1][1], s32[2][1], s32[3][1]), 16, l16);
                if (half == 0) c3232 = r1;
                else { const uint32_t f = shfl_min(shfl_min(shfl_min(pair_min(c3232, r1, 8, l8), 4), 2), 1); if ((lane & 7) == 0) red_min(SLOT(T32x32), f); }
            }
            {   // 64x32: half l16; the rest by shuffle
                const uint32_t r0 = K(s64h[0], s64h[1], s64h[2], s64h[3]);
                if (half == 0) c6432 = r0;
                else {
                    const uint32_t f = shfl_min(shfl_min(shfl_min(shfl_min(pair_min(c6432, r0, 16, l16), 8), 4), 2), 1);
                    if ((lane & 15) == 0) red_min(SLOT(T64x32), f);
                }
            }
            if (half == 0) {
#pragma unroll
                for (int k = 0; k < 4; k++) {
#pragma unroll
                    for (int q = 0; q < 4; q++) w16[k][q] = v16[k][q];
                    w32[k][0] = s32[k][0]; w32[k][1] = s32[k][1]; s64[k] = s64h[k];
                }
                continue;
            }
            // ---- whole SB: 16x64, 32x64, 64x64, and the per-position S / T tables of quirk Q11
#pragma unroll
            for (int k = 0; k < 4; k++) {
#pragma unroll
                for (int q = 0; q < 4; q++) w16[k][q] += v16[k][q];
                w32[k][0] += s32[k][0]; w32[k][1] += s32[k][1]; s64[k] += s64h[k];
                if (val[k]) { const int pos = ((ty0 + y) << QS) | (tx0 + x0 + 4 * k); q11_s[pos] = s64h[k]; q11_t[pos] = (uint16_t)m32[k][0]; }
            }
            {
                uint32_t kq[4];
#pragma unroll
                for (int q = 0; q < 4; q++) kq[q] = K(w16[0][q], w16[1][q], w16[2][q], w16[3][q]);
                const uint32_t r2 = pair_min(pair_min(kq[0], kq[1], 16, l16), pair_min(kq[2], kq[3], 16, l16), 8, l8);
                const uint32_t f = shfl_min(shfl_min(shfl_min(r2, 4), 2), 1);
                if ((lane & 7) == 0) red_min(SLOT(T16x64), f);
            }
            {
                const uint32_t r1 = pair_min(K(w32[0][0], w32[1][0], w32[2][0], w32[3][0]), K(w32[0][1], w32[1][1], w32[2][1], w32[3][1]), 16, l16);
                const uint32_t f = shfl_min(shfl_min(shfl_min(shfl_min(r1, 8), 4), 2), 1);
                if ((lane & 15) == 0) red_min(SLOT(T32x64), f);
            }
            {
                const uint32_t m = __reduce_min_sync(0xffffffffu, K(s64[0], s64[1], s64[2], s64[3]));
                if (lane == 0) red_min(keys_addr, m);
            }
        }
    }
#undef SLOT
    __syncthreads();

    // ---- decode the winning keys into p_sb_best_sad / p_sb_best_mv form (internal index = slot)
    for (int t = tid; t < kMaxPu; t += kFxThreads) {
        if (t == 87 + 5) continue;                                  // quirk Q11: replayed below
        const uint32_t key = keys[t];
        const uint32_t v = key >> 12; const int y = (key >> 6) & 63, x = key & 63;
        if (Wide) {                                                 // merge: smaller SAD, then earlier row, then earlier column
            const uint32_t o = ((uint32_t)(ty0 + y) << 8) | (uint32_t)(tx0 + x);
            if (key == 0xFFFFFFFFu || !(v < merged_v[t] || (v == merged_v[t] && o < merged_o[t]))) continue;
            merged_v[t] = v; merged_o[t] = o;
        }
        const uint32_t sad = v << 1, mv = make_mv((int16_t)((sax + x) * 4), (int16_t)((say + y) * 4));
        if (t < kNumPu) { scr.best_sad[list][t] = sad; scr.best_mv[list][t] = mv; }
        else { scx.best_sad[list][t - kNumPu] = sad; scx.best_mv[list][t - kNumPu] = mv; }
    }
    if (Wide) __syncthreads();                                      // before the next sub-area's window and key reset
    }   // sub-areas

    // ---- quirk Q11: replay the sequential update rule of 32x16 PU #5 over the positions in raster order.
    // `best` strictly decreases (T <= S), so the rule is a short chain of queries "first position after n whose S is
    // below best": every thread keeps its share of the S table in registers (Wide: reads it from shared memory) and the
    // CTA answers one query per step (a handful of steps: S covers four times the area of T, so it rarely drops below an
    // earlier T).
    {
        constexpr int kPer = QN / kFxThreads;
        auto s_of = [&](int n) { return ((n & ((1 << QS) - 1)) < area_w && (n >> QS) < area_h) ? q11_s[n] : 0xFFFFFFFFu; };
        uint32_t S[Wide ? 1 : kPer];
        if (!Wide) {
#pragma unroll
            for (int i = 0; i < kPer; i++) S[i] = s_of(tid + kFxThreads * i);      // raster index (row << QS | col)
        }
        uint32_t* red = keys + kMaxPu;                                // three spare words of the key table
        uint32_t best = 0xFFFFFFFFu;
        int cur = -1;
        const int lane = tid & 31;
        for (;;) {
            uint32_t cand = 0xFFFFFFFFu;
            if (!Wide) {
#pragma unroll
                for (int i = kPer - 1; i >= 0; i--) {
                    const int n = tid + kFxThreads * i;
                    if (n > cur && S[i] < best) cand = (uint32_t)n;
                }
            } else {
                for (int n = tid + kFxThreads * max(0, (cur + 1 - tid + kFxThreads - 1) / kFxThreads); n < QN && (n >> QS) < area_h; n += kFxThreads)
                    if (n > cur && s_of(n) < best) { cand = (uint32_t)n; break; }
            }
            cand = __reduce_min_sync(0xffffffffu, cand);
            if (tid == 0) red[0] = 0xFFFFFFFFu;
            __syncthreads();
            if (lane == 0 && cand != 0xFFFFFFFFu) atomicMin(&red[0], cand);
            __syncthreads();
            const uint32_t first = red[0];
            __syncthreads();
            if (first == 0xFFFFFFFFu) break;
            cur = (int)first; best = q11_t[first];
        }
        if (tid == 0) {
            scx.best_sad[list][87 + 5 - kNumPu] = best << 1;
            scx.best_mv[list][87 + 5 - kNumPu] = make_mv((int16_t)((area_x + (cur & ((1 << QS) - 1))) * 4), (int16_t)((area_y + (cur >> QS)) * 4));
        }
    }
}

cudaError_t configure_fullpel_ext() {
    cudaError_t rc = cudaFuncSetAttribute(fullpel_ext_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, fx_smem_bytes<false>());
    if (rc != cudaSuccess) return rc;
    return cudaFuncSetAttribute(fullpel_ext_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, fx_smem_bytes<true>());
}

void launch_fullpel_ext(const DJob* d_jobs, int n_jobs, int max_sb, int list, bool wide, cudaStream_t st) {
    if (wide) fullpel_ext_kernel<true><<<dim3(max_sb, n_jobs), kFxThreads, fx_smem_bytes<true>(), st>>>(d_jobs, list);
    else      fullpel_ext_kernel<false><<<dim3(max_sb, n_jobs), kFxThreads, fx_smem_bytes<false>(), st>>>(d_jobs, list);
}

}  // namespace svtme
