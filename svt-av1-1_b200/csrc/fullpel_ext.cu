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
// B200 mapping: as in fullpel.cu -- four byte-shifted copies of the reference window in shared memory, VABSDIFF4.U8.ACC
// on the ALU pipe, keys built with IMAD on the FMA pipe and reduced across the warp by streaming transpose-reductions
// (pair_min) so that at the end of a tile every lane issues at most one RED.SHARED.MIN per PU class. The non-square SADs
// are sums of the 8x8 SADs: horizontal sums inside a band of eight rows, vertical sums carried across bands in registers.
// Those carried sums are what costs registers, so a quad of positions (x, x+4, x+8, x+12 of one row) is shared by TWO
// lanes, one per 32-column half of the SB (lane bit 16): half the sums per thread, 128 registers, two CTAs (16 warps)
// per SM instead of one 255-register CTA. The three 64-wide classes add the two halves with one shuffle per position.
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
constexpr int kKeyWords = 224;         // 209 PU slots (internal index), padded so that the raw rows behind them start on a 128-byte boundary (TMA)
constexpr int kRawPitch = 40;          // words per raw window row (ten 16-byte granules), see fullpel.cu
template <bool Wide> struct FxCfg {    // 256 threads x 2 CTAs per SM; the 128x128 Q11 tables of Wide leave room for one CTA: 512 threads
    static constexpr int kThreads = Wide ? 512 : 256, kCtas = Wide ? 1 : 2;
};
template <bool Wide> struct Q11 { static constexpr int kShift = Wide ? 7 : 6, kEntries = 1 << (2 * kShift); };
// The raw window rows (20160 bytes) are dead once the four copies are made and the Q11 tables are first written by the
// search: the narrow instantiation overlays them (24576 bytes of tables); Wide restages per sub-area while its tables live on.
template <bool Wide> constexpr int fx_smem_bytes() {
    return (kWinWords + kSrcWords + kKeyWords + Q11<Wide>::kEntries) * 4 + Q11<Wide>::kEntries * 2 + (Wide ? kWinRows * kRawPitch * 4 : 0);
}
static_assert(kWinRows * kRawPitch * 4 <= Q11<false>::kEntries * 6, "raw rows must fit inside the narrow Q11 tables");

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

// The reductions transpose over lane bits 8, 4, 2, 1 in that order (lb[i] = lane bit (8 >> i) as 0 / 1): first the
// thread's own columns (cols4 / cols2: register index bit i -> lane bit i of the order), then time.
// cols4: the four per-thread columns of a class -> one key, lane bits (8, 4) = column within the half
template <typename V> __device__ __forceinline__ uint32_t cols4(const Keyer& K, const V& v, const uint32_t (&lb)[4]) {
    uint32_t k[4];
#pragma unroll
    for (int j = 0; j < 4; j++) k[j] = K(v[0][j], v[1][j], v[2][j], v[3][j]);
    return pair_min(pair_min(k[0], k[1], 8, lb[0]), pair_min(k[2], k[3], 8, lb[0]), 4, lb[1]);
}
template <typename V> __device__ __forceinline__ uint32_t cols2(const Keyer& K, const V& v, const uint32_t (&lb)[4]) {
    return pair_min(K(v[0][0], v[1][0], v[2][0], v[3][0]), K(v[0][1], v[1][1], v[2][1], v[3][1]), 8, lb[0]);
}
// Streaming reduction over time: value v arrives at step t (a constant after unrolling) of 2^LT steps; step bit i is
// transposed onto lane bit 8 >> (LC + i) (LC = bits the columns took). True when the reduced key of all steps drops out.
template <int LC, int LT> __device__ __forceinline__ bool chain(uint32_t (&c)[3], int t, uint32_t v, uint32_t& out, const uint32_t (&lb)[4]) {
#pragma unroll
    for (int i = 0; i < LT; i++) {
        if (((t >> i) & 1) == 0) { c[i] = v; return false; }
        v = pair_min(c[i], v, 8 >> (LC + i), lb[LC + i]);
    }
    out = v;
    return true;
}

}  // namespace

namespace {
// per-lane PU slot tables: which PU (internal index) a lane owns after the transpose-reductions of each class
enum { T8A = 0, T8B, T16x8, T32x8, T8x16, T16x16, T32x16, T64x16, T8x32, T16x32, T32x32, T64x32, T16x64, T32x64, kNumTables };

// Lane bit 16 = column half H; lane bits 8, 4, 2, 1 = v[0..3]. A class whose threads hold 2^LC columns and that runs
// over 2^LT time steps ends with column = v[0..LC-1] (+ H above them) and step = v[LC..LC+LT-1]; lanes with any
// later bit set issue nothing (their entry is unused).
__device__ __forceinline__ int lane_pu(int table, int lane) {
    const int H = (lane >> 4) & 1, v0 = (lane >> 3) & 1, v1 = (lane >> 2) & 1, v2 = (lane >> 1) & 1, v3 = lane & 1;
    const int j = v0 + 2 * v1 + 4 * H, q = v0 + 2 * H, p = H;       // 8-, 16-, 32-pixel column
    switch (table) {
        case T8A:    return kInternal[21 + (v2 + 2 * v3) * 8 + j];
        case T8B:    return kInternal[21 + (4 + v2 + 2 * v3) * 8 + j];
        case T16x8:  return kInternal[95 + (v1 + 2 * v2 + 4 * v3) * 4 + q];
        case T32x8:  return kInternal[169 + (v0 + 2 * v1 + 4 * v2) * 2 + p];
        case T8x16:  return kInternal[137 + (v2 + 2 * v3) * 8 + j];
        case T16x16: return kInternal[5 + (v1 + 2 * v2) * 4 + q];
        case T32x16: return kInternal[87 + (v0 + 2 * v1) * 2 + p];
        case T64x16: return 201 + v0 + 2 * v1;
        case T8x32:  return kInternal[185 + v2 * 8 + j];
        case T16x32: return kInternal[129 + v1 * 4 + q];
        case T32x32: return 1 + v0 * 2 + p;
        case T64x32: return 85 + v0;
        case T16x64: return 205 + q;
        default:     return 127 + p;      // T32x64
    }
}
}  // namespace

template <bool Wide>
__global__ void __launch_bounds__(FxCfg<Wide>::kThreads, FxCfg<Wide>::kCtas) fullpel_ext_kernel(const DJob* __restrict__ jobs, int list) {
    constexpr int QS = Q11<Wide>::kShift, QN = Q11<Wide>::kEntries, NT = FxCfg<Wide>::kThreads;
    extern __shared__ __align__(128) uint32_t smem[];
    __shared__ __align__(8) uint64_t tma_bar;
    uint32_t* win = smem;
    uint32_t* src = win + kWinWords;
    uint32_t* keys = src + kSrcWords;
    uint32_t* q11_s = keys + kKeyWords;                                  // bottom 64x32 SAD of every position
    uint16_t* q11_t = reinterpret_cast<uint16_t*>(q11_s + QN);           // 32x16 (x 0, y 48) SAD of every position
    // raw window rows: over the Q11 tables (byte offset 75520 = 128 * 590), Wide behind them (173824 = 128 * 1358)
    uint32_t* raw = Wide ? reinterpret_cast<uint32_t*>(q11_t + QN) : q11_s;
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
    if (Wide) for (int t = tid; t < kMaxPu; t += NT) { merged_v[t] = 0xFFFFFFFFu; merged_o[t] = 0xFFFFFFFFu; }
    const void* tmap = job.ref_tmap[list];                        // CTA-uniform; see fullpel.cu
    const uint32_t bar_addr = (uint32_t)__cvta_generic_to_shared(&tma_bar);
    uint32_t tma_phase = 0;
    if (tmap && tid == 0) mbar_init(bar_addr, 1);
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
    if (tmap) {
        if (tid == 0) {
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // earlier generic-proxy accesses of this memory (Q11 tables / the previous sub-area) come first
            mbar_expect_tx(bar_addr, kTmaBoxBytes * kTmaBoxRows);
            tma_load_2d((uint32_t)__cvta_generic_to_shared(raw), tmap, rp.ox + ox + sax - (int)m16, rp.oy + oy + say, bar_addr);
        }
    } else {
        const int ngran = ((int)(m16 >> 2) + nfill + 2 + 3) >> 2;
        const uint8_t* ga = g0 - m16;
        const uint32_t raw_addr = (uint32_t)__cvta_generic_to_shared(raw);
        for (int t = tid; t < wrows * ngran; t += NT) {
            const int r = t / ngran, g = t - r * ngran;
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"(raw_addr + (uint32_t)(r * kRawPitch * 4 + g * 16)),
                         "l"(ga + (size_t)r * rp.stride + g * 16) : "memory");
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    }
    if (!Wide || sub == 0) {
        for (int t = tid; t < kSrcWords; t += NT)
            src[t] = *reinterpret_cast<const uint32_t*>(job.cur.full.at(ox + 4 * (t & 15), oy + 2 * (t >> 4)));
        for (int t = tid; t < kNumTables * 32; t += NT) lane_slot[t >> 5][t & 31] = keys_addr + 4u * (uint32_t)lane_pu(t >> 5, t & 31);
    }
    for (int t = tid; t < kKeyWords; t += NT) keys[t] = 0xFFFFFFFFu;
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();
    if (tmap) { mbar_wait(bar_addr, tma_phase); tma_phase ^= 1u; }
    {
        const int c = tid & 31;
        if (c < nfill) {
            const uint32_t m = m16 & 3u;
            const uint32_t* rw = raw + (m16 >> 2) + c;
            for (int r = tid >> 5; r < wrows; r += NT / 32) {
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
    const int nquads = sah * ngp;                      // quad = the 4 positions x0, x0 + 4, x0 + 8, x0 + 12 of one (row, group, phase)
    const int lane = tid & 31, s = lane & 3, H = lane >> 4;
    // lane bits 8, 4, 2, 1 as 0 / 1 (the transposition bits of the reductions, in the order the chains use them)
    const uint32_t lb[4] = { (uint32_t)(lane >> 3) & 1u, (uint32_t)(lane >> 2) & 1u, (uint32_t)(lane >> 1) & 1u, (uint32_t)lane & 1u };
#define SLOT(t) lane_slot[t][lane]

    for (int q0 = 0; q0 < nquads; q0 += NT / 8) {
        const int qi = q0 + (tid >> 5) * 4 + ((lane >> 2) & 3);
        const bool live = qi < nquads;
        int y = 0, x0 = s;
        if (live) { const int G = qi % ngp; y = qi / ngp; x0 = 16 * G + s; }
        Keyer K;
        bool val[4];
#pragma unroll
        for (int k = 0; k < 4; k++) {
            val[k] = live && (x0 + 4 * k < saw);
            K.mul[k] = val[k] ? 4096u : 0u;
            K.add[k] = val[k] ? (uint32_t)((y << 6) | (x0 + 4 * k)) : 0xFFFFFFFFu;
        }
        uint32_t acc[2][4][4];                          // 8x8 SADs of the current band [b & 1] and of the band before (this thread's four columns)
        uint32_t h16[4][2], v16[4][2], w16[4][2];       // 16x16 (band pair), 16x32 (half), 16x64 column sums
        uint32_t c32[4][4];                             // 8x32 column sums of the current half
        uint32_t w32[4];                                // 32x64 of this thread's 32 columns
        // carries of the streaming transpose-reductions, one chain per PU class
        uint32_t c8[3], c168[3], c328[3], c816[3], c1616[3], c3216[3], c6416[3], c832[3], c1632[3], c3232[3], c6432[3];
#pragma unroll
        for (int b = 0; b < 8; b++) {
            uint32_t (&A)[4][4] = acc[b & 1];
#pragma unroll
            for (int k = 0; k < 4; k++)
#pragma unroll
                for (int j = 0; j < 4; j++) A[k][j] = 0;
#pragma unroll
            for (int rr = 0; rr < 4; rr++) {
                const int r = 4 * b + rr;
                const uint4* wp = reinterpret_cast<const uint4*>(win + (s * kWinRows + y + 2 * r) * kWinPitch + (x0 >> 2) + 8 * H);
                uint32_t ref[12];
#pragma unroll
                for (int i = 0; i < 3; i++) {
                    const uint4 v = wp[i];
                    ref[4 * i] = v.x; ref[4 * i + 1] = v.y; ref[4 * i + 2] = v.z; ref[4 * i + 3] = v.w;
                }
                const uint4* sp = reinterpret_cast<const uint4*>(src + r * 16 + 8 * H);
#pragma unroll
                for (int cc = 0; cc < 2; cc++) {
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
            uint32_t out;
            // ---- band b (rows 8b .. 8b+7): 8x8, 16x8, 32x8 PUs of this band
            uint32_t h[4][2], g[4];
#pragma unroll
            for (int k = 0; k < 4; k++) {
                h[k][0] = A[k][0] + A[k][1]; h[k][1] = A[k][2] + A[k][3];
                g[k] = h[k][0] + h[k][1];
            }
            if (chain<2, 2>(c8, b & 3, cols4(K, A, lb), out, lb)) red_min(b == 3 ? SLOT(T8A) : SLOT(T8B), out);
            if (chain<1, 3>(c168, b, cols2(K, h, lb), out, lb)) red_min(SLOT(T16x8), out);
            if (chain<0, 3>(c328, b, K(g[0], g[1], g[2], g[3]), out, lb)) { out = shfl_min(out, 1); if (!lb[3]) red_min(SLOT(T32x8), out); }
            if ((b & 1) == 0) {
#pragma unroll
                for (int k = 0; k < 4; k++) { h16[k][0] = h[k][0]; h16[k][1] = h[k][1]; }
                continue;
            }
            // ---- band pair (rows 16*pair .. +15): 8x16, 16x16, 32x16, 64x16
            const int pair = b >> 1;
            uint32_t (&E)[4][4] = acc[0];               // the even band of the pair (A is acc[1] here)
            uint32_t m32[4], m64[4];
#pragma unroll
            for (int k = 0; k < 4; k++) {
#pragma unroll
                for (int j = 0; j < 4; j++) E[k][j] += A[k][j];          // E becomes the 8x16 column sums
                h16[k][0] += h[k][0]; h16[k][1] += h[k][1];
                m32[k] = h16[k][0] + h16[k][1];
                m64[k] = m32[k] + __shfl_xor_sync(0xffffffffu, m32[k], 16);
            }
            if (chain<2, 2>(c816, pair, cols4(K, E, lb), out, lb)) red_min(SLOT(T8x16), out);
            if (chain<1, 2>(c1616, pair, cols2(K, h16, lb), out, lb)) { out = shfl_min(out, 1); if (!lb[3]) red_min(SLOT(T16x16), out); }
            // (the slot of 32x16 PU #5 (x 0, y 48) is overwritten by the replay of quirk Q11 after the search)
            if (chain<0, 2>(c3216, pair, K(m32[0], m32[1], m32[2], m32[3]), out, lb)) { out = shfl_min(shfl_min(out, 2), 1); if (!(lane & 3)) red_min(SLOT(T32x16), out); }
            if (chain<0, 2>(c6416, pair, K(m64[0], m64[1], m64[2], m64[3]), out, lb)) { out = shfl_min(shfl_min(out, 2), 1); if (!(lane & 19)) red_min(SLOT(T64x16), out); }
            if ((pair & 1) == 0) {
#pragma unroll
                for (int k = 0; k < 4; k++) {
#pragma unroll
                    for (int j = 0; j < 4; j++) c32[k][j] = E[k][j];
                    v16[k][0] = h16[k][0]; v16[k][1] = h16[k][1];
                }
                continue;
            }
            // ---- half (rows 32*half .. +31): 8x32, 16x32, 32x32, 64x32
            const int half = b >> 2;
            uint32_t s32[4], s64h[4];
#pragma unroll
            for (int k = 0; k < 4; k++) {
#pragma unroll
                for (int j = 0; j < 4; j++) c32[k][j] += E[k][j];
                v16[k][0] += h16[k][0]; v16[k][1] += h16[k][1];
                s32[k] = v16[k][0] + v16[k][1];
                s64h[k] = s32[k] + __shfl_xor_sync(0xffffffffu, s32[k], 16);
            }
            if (chain<2, 1>(c832, half, cols4(K, c32, lb), out, lb)) { out = shfl_min(out, 1); if (!lb[3]) red_min(SLOT(T8x32), out); }
            if (chain<1, 1>(c1632, half, cols2(K, v16, lb), out, lb)) { out = shfl_min(shfl_min(out, 2), 1); if (!(lane & 3)) red_min(SLOT(T16x32), out); }
            if (chain<0, 1>(c3232, half, K(s32[0], s32[1], s32[2], s32[3]), out, lb)) { out = shfl_min(shfl_min(shfl_min(out, 4), 2), 1); if (!(lane & 7)) red_min(SLOT(T32x32), out); }
            if (chain<0, 1>(c6432, half, K(s64h[0], s64h[1], s64h[2], s64h[3]), out, lb)) { out = shfl_min(shfl_min(shfl_min(out, 4), 2), 1); if (!(lane & 23)) red_min(SLOT(T64x32), out); }
            if (half == 0) {
#pragma unroll
                for (int k = 0; k < 4; k++) { w16[k][0] = v16[k][0]; w16[k][1] = v16[k][1]; w32[k] = s32[k]; }
                continue;
            }
            // ---- whole SB: 16x64, 32x64, 64x64, and the per-position S / T tables of quirk Q11 (S = bottom 64x32, T = the
            // 32x16 PU at x 0, y 48: the left half's 32x16 sum of the last band pair)
            uint32_t s64[4];
#pragma unroll
            for (int k = 0; k < 4; k++) {
                w16[k][0] += v16[k][0]; w16[k][1] += v16[k][1]; w32[k] += s32[k];
                s64[k] = w32[k] + __shfl_xor_sync(0xffffffffu, w32[k], 16);
                if (val[k] && !H) { const int pos = ((ty0 + y) << QS) | (tx0 + x0 + 4 * k); q11_s[pos] = s64h[k]; q11_t[pos] = (uint16_t)m32[k]; }
            }
            out = cols2(K, w16, lb);
            out = shfl_min(shfl_min(shfl_min(out, 4), 2), 1);
            if (!(lane & 7)) red_min(SLOT(T16x64), out);
            out = K(w32[0], w32[1], w32[2], w32[3]);
            out = shfl_min(shfl_min(shfl_min(shfl_min(out, 8), 4), 2), 1);
            if (!(lane & 15)) red_min(SLOT(T32x64), out);
            {
                const uint32_t m = __reduce_min_sync(0xffffffffu, K(s64[0], s64[1], s64[2], s64[3]));
                if (lane == 0) red_min(keys_addr, m);
            }
        }
    }
#undef SLOT
    __syncthreads();

    // ---- decode the winning keys into p_sb_best_sad / p_sb_best_mv form (internal index = slot)
    for (int t = tid; t < kMaxPu; t += NT) {
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
        constexpr int kPer = QN / NT;
        auto s_of = [&](int n) { return ((n & ((1 << QS) - 1)) < area_w && (n >> QS) < area_h) ? q11_s[n] : 0xFFFFFFFFu; };
        uint32_t S[Wide ? 1 : kPer];
        if (!Wide) {
#pragma unroll
            for (int i = 0; i < kPer; i++) S[i] = s_of(tid + NT * i);      // raster index (row << QS | col)
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
                    const int n = tid + NT * i;
                    if (n > cur && S[i] < best) cand = (uint32_t)n;
                }
            } else {
                for (int n = tid + NT * max(0, (cur + 1 - tid + NT - 1) / NT); n < QN && (n >> QS) < area_h; n += NT)
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
    if (wide) fullpel_ext_kernel<true><<<dim3(max_sb, n_jobs), FxCfg<true>::kThreads, fx_smem_bytes<true>(), st>>>(d_jobs, list);
    else      fullpel_ext_kernel<false><<<dim3(max_sb, n_jobs), FxCfg<false>::kThreads, fx_smem_bytes<false>(), st>>>(d_jobs, list);
}

}  // namespace svtme
