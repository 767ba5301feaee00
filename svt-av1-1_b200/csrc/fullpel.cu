// fullpel.cu -- integer full-pel search of one 64x64 SB against one reference list: SADs of all 85
// square PUs (64x64, 4x 32x32, 16x 16x16, 64x 8x8) at every position of the (<= 64x64) search area and
// the per-PU argmin, bit-exact with FullPelSearch_LCU and its AVX2/SSE2 leaves
// (reference Source/Lib/Codec/EbMotionEstimation.c:1237-1551,
//  Source/Lib/ASM_AVX2/EbComputeSAD_Intrinsic_AVX2.c:3414-3932,
//  Source/Lib/ASM_SSE2/EbMeSadCalculation_Intrinsic_SSE2.c:10-127).
//
// B200 mapping (INT-pipe bound: VABSDIFF4.U8.ACC issues at 64 lanes/clk/SM on the ALU pipe, shared
// with SHF/LOP3/VIMNMX, so the hot loop must contain nothing but SADs):
//   * the reference window (<= 126 rows x 128 bytes) is staged once into shared memory as FOUR copies,
//     copy s shifted left by s bytes, so that every search position reads word-aligned data and the
//     inner loop needs no funnel shifts;
//   * one thread = one "tile" of 4 positions x, x+4, x+8, x+12 on one row: they share the same copy,
//     19 reference words feed 64 SADs per source row, source words are broadcast LDS.128;
//   * the reference kernels look at 8 rows of every 8x8 block sub-sampled by two (rows 0,2,4,6) and
//     double the sum; per band of four such rows a tile owns 8 column blocks x 4 positions = 32
//     accumulators; 16x16 / 32x32 / 64x64 SADs are sums of those;
//   * argmin: 32-bit key = sad * 4096 + (row << 6 | col), built with one IMAD (FMA pipe); the keys of a warp are
//     reduced by a streaming transpose (pair_min below) and every lane issues one RED.MIN to its own PU slot,
//     which is exactly "smallest SAD, first in raster order"; the 32x32 PUs use sad * 8192 + (row, group, idx)
//     to carry the reference's gcc-only MV quirk (Q4); the 64x64 PU is one warp-wide minimum.
//   * search areas wider or taller than 64 (MIN(search_area, 127), ME.c:6644-6645; template parameter Wide): the area
//     is cut into sub-areas of at most 64x64 positions on multiples of 64 from its origin (so the reference's
//     8-position groups are not split), each sub-area is staged and searched exactly as above, and the winners are
//     merged per PU in the reference's order: smaller SAD, then earlier row, then earlier group / column. The
//     narrow instantiation compiles to the single-area code.
#include "me_common.cuh"

namespace svtme {

constexpr int kFpThreads = 256;
constexpr int kWinRows = 126;          // 64 search rows + 62
constexpr int kWinPitch = 36;          // words; = 4 (mod 32) keeps the 8 row-distinct LDS.128 of a warp conflict-free
constexpr int kRawPitch = 40;          // words: one raw window row = ten 16-byte granules from the aligned address below its first byte
constexpr int kFpSmemBytes = (4 * kWinRows * kWinPitch + 32 * 16 + 96 + kWinRows * kRawPitch) * 4;

// One step of a streaming transpose-reduction: two keys a, b live in every lane; afterwards lanes whose `bit` is
// clear hold min(a) over {lane, lane ^ bit} and lanes whose bit is set hold the same for b. Five such steps
// (bits 16, 8, 4, 2, 1) turn 32 keys x 32 lanes into one fully reduced key per lane at 4 instructions per key
// (2 SEL + SHFL + VIMNMX), after which every lane issues one RED.MIN to its own PU slot: no per-key CREDUX,
// leader election or branch.
// The keep/send selection is written as two IMADs (hi01 in {0,1}: keep = a + hi01*(b-a), send = b - hi01*(b-a)) so
// that it runs on the FMA pipe: every ALU-pipe instruction costs as much as one VABSDIFF4 (both 64 lanes/clk/SM).
__device__ __forceinline__ uint32_t pair_min(uint32_t a, uint32_t b, int bit, uint32_t hi01) {
    uint32_t d, keep, send;
    asm("sub.u32 %0, %1, %2;" : "=r"(d) : "r"(b), "r"(a));
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(keep) : "r"(hi01), "r"(d), "r"(a));
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(send) : "r"(0u - hi01), "r"(d), "r"(b));
    return min(keep, __shfl_xor_sync(0xffffffffu, send, bit));
}
__device__ __forceinline__ void red_min(uint32_t slot_addr, uint32_t key) {
    asm volatile("red.shared.min.u32 [%0], %1;" :: "r"(slot_addr), "r"(key) : "memory");
}

template <bool Wide>
__global__ void __launch_bounds__(kFpThreads, 2) fullpel_kernel(const DJob* __restrict__ jobs, int list) {
    extern __shared__ __align__(128) uint32_t smem[];
    __shared__ uint32_t merged_v[Wide ? kNumPu : 1], merged_o[Wide ? kNumPu : 1];   // Wide: running winner per PU (value, order key)
    __shared__ __align__(8) uint64_t tma_bar;
    uint32_t* win = smem;
    uint32_t* src = smem + 4 * kWinRows * kWinPitch;
    uint32_t* keys = src + 32 * 16;
    uint32_t* raw = keys + 96;                                   // 128-byte aligned (a TMA destination): (4 * 126 * 36 + 512 + 96) * 4 = 128 * 586
    const DJob& job = jobs[blockIdx.y];
    const int sb = job.sb_begin + blockIdx.x;
    if (sb >= job.sb_end) return;
    const SvtMeParams& p = job.prm;
    if (list >= p.num_lists || p.num_pus != kNumPu) return;     // 209-PU jobs are searched by fullpel_ext_kernel
    if (wide_search(p) != Wide) return;
    const int tid = threadIdx.x;
    const int ox = (sb % job.sb_cols) * 64, oy = (sb / job.sb_cols) * 64;
    SbScratch& scr = job.scratch[sb];
    const DPlane& rp = job.ref[list].full;
    const int area_x = scr.sa_x[list], area_y = scr.sa_y[list], area_w = scr.sa_w[list], area_h = scr.sa_h[list];
    if (Wide && tid < kNumPu) { merged_v[tid] = 0xFFFFFFFFu; merged_o[tid] = 0xFFFFFFFFu; }
    const void* tmap = job.ref_tmap[list];                        // CTA-uniform
    const uint32_t bar_addr = (uint32_t)__cvta_generic_to_shared(&tma_bar);
    uint32_t tma_phase = 0;
    if (tmap && tid == 0) mbar_init(bar_addr, 1);
    for (int sub = 0; sub < (Wide ? 4 : 1); sub++) {
    // sub-area of this pass: the whole area, or (Wide) one of up to 2 x 2 pieces of at most 64 x 64 positions
    const int tx0 = Wide ? (sub & 1) * 64 : 0, ty0 = Wide ? (sub >> 1) * 64 : 0;
    if (Wide && (tx0 >= area_w || ty0 >= area_h)) continue;       // CTA-uniform
    const int sax = area_x + tx0, say = area_y + ty0, saw = Wide ? min(64, area_w - tx0) : area_w, sah = Wide ? min(64, area_h - ty0) : area_h;

    // ---- stage source (even rows) and the four byte-shifted window copies.
    // The window first lands RAW in shared memory as 16-byte cp.async granules (fire and forget: no thread holds a
    // register or waits per row, the whole window costs one L2 round trip), then every warp shifts its rows out of the raw
    // buffer into the four copies (3 LDS + 4 SHF + 4 STS per word, no global latency inside the loop). With ordinary
    // loads the staging loop was sixteen dependent L2 round trips per warp, a fifth of the CTA's lifetime, during which
    // the other CTA of the SM was alone on the SAD pipe. (TMA cannot make the copies: tensor boxes must start on a
    // 16-byte boundary, profiles/microbench/tma_probe_r01.txt.)
    const int wrows = sah + 62;
    const int nfill = min(32, ((saw + 62 + 3) >> 2) + 1);
    const uint8_t* g0 = rp.at(ox + sax, oy + say);
    const uint32_t m16 = (uint32_t)(uintptr_t)g0 & 15u;           // the same for every row: the pitch is a multiple of 64
    if (tmap) {
        // one box for the whole window (rows past the window and bytes past the plane's pitch are fetched or zero-filled and
        // never read). Thread 0 both initialised the mbarrier and issues; every other thread first touches it after the
        // __syncthreads() below.
        if (tid == 0) {
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // the previous sub-area's reads of `raw` (generic proxy) come first
            mbar_expect_tx(bar_addr, kTmaBoxBytes * kTmaBoxRows);
            tma_load_2d((uint32_t)__cvta_generic_to_shared(raw), tmap, rp.ox + ox + sax - (int)m16, rp.oy + oy + say, bar_addr);
        }
    } else {
        const int ngran = ((int)(m16 >> 2) + nfill + 2 + 3) >> 2;     // granules per row that hold the words read below (<= 10)
        const uint8_t* ga = g0 - m16;
        const uint32_t raw_addr = (uint32_t)__cvta_generic_to_shared(raw);
        for (int t = tid; t < wrows * ngran; t += kFpThreads) {
            const int r = t / ngran, g = t - r * ngran;
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"(raw_addr + (uint32_t)(r * kRawPitch * 4 + g * 16)),
                         "l"(ga + (size_t)r * rp.stride + g * 16) : "memory");
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    }
    if (!Wide || sub == 0) {                                      // behind the window's asynchronous copies
        for (int t = tid; t < 32 * 16; t += kFpThreads)
            src[t] = *reinterpret_cast<const uint32_t*>(job.cur.full.at(ox + 4 * (t & 15), oy + 2 * (t >> 4)));
    }
    if (tid < 96) keys[tid] = 0xFFFFFFFFu;
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();
    if (tmap) { mbar_wait(bar_addr, tma_phase); tma_phase ^= 1u; }
    {
        // lane = window column (word), warp = every eighth row: three raw words give the four byte-shifted copies of a word
        const int c = tid & 31;
        if (c < nfill) {
            const uint32_t m = m16 & 3u;
            const uint32_t* rw = raw + (m16 >> 2) + c;
            for (int r = tid >> 5; r < wrows; r += kFpThreads / 32) {
                const uint32_t* w = rw + r * kRawPitch;
                const uint32_t w0 = w[0], w1 = w[1], w2 = w[2];
#pragma unroll
                for (int sc = 0; sc < 4; sc++) {
                    const uint32_t o = m + sc;                       // byte offset of copy sc inside the aligned word
                    const uint32_t lo = o < 4 ? w0 : w1, hi = o < 4 ? w1 : w2;
                    win[(sc * kWinRows + r) * kWinPitch + c] = __funnelshift_r(lo, hi, (o & 3u) * 8u);
                }
            }
        }
    }
    __syncthreads();

    const int ngroups = saw >> 3, mult8 = ngroups << 3, leftover = saw - mult8;
    // Wide: the width & 7 leftover positions (the reference's single-position SSE2 kernels) ride in the tiles of the
    // 16-position pair they fall into -- the empty half of the last pair when the number of groups is odd (a 63-wide
    // sub-area), else one more pair; they differ from group members only in their 32x32 key. The narrow instantiation keeps
    // separate single-position tiles: with 64-wide preset areas there are none, and the per-position test costs it 1 %.
    const int ngpairs = Wide ? (saw + 15) >> 4 : (ngroups + 1) >> 1;
    const int tiles_g = sah * ngpairs * 4;                         // 8-position-group tiles
    const int tiles_s = (!Wide && leftover) ? sah * 4 : 0;         // single-position tiles (width & 7 leftovers)
    const int total = tiles_g + tiles_s;
    const int s = tid & 3;
    const uint32_t keys_addr = (uint32_t)__cvta_generic_to_shared(keys);
    const int lane = tid & 31;
    const uint32_t l16 = (lane >> 4) & 1, l8 = (lane >> 3) & 1, l4 = (lane >> 2) & 1, l2 = (lane >> 1) & 1, l1 = lane & 1;
    // PU slot owned by this lane after the transpose-reductions (key index = bit-reversed lane):
    //   8x8  : column block j = (l16, l8, l4), block row within the half = (l2, l1); internal index = Morton(j, row)
    //   16x16: q = (l16, l8), row = (l4, l2);   32x32: q = l16, half = l8
    const uint32_t addr8 = keys_addr + 4u * (21u + (l16 ? 1u : 0u) + (l8 ? 4u : 0u) + (l4 ? 16u : 0u) + (l2 ? 2u : 0u) + (l1 ? 8u : 0u));
    const uint32_t addr16 = keys_addr + 4u * (5u + (l16 ? 1u : 0u) + (l8 ? 4u : 0u) + (l4 ? 2u : 0u) + (l2 ? 8u : 0u));
    const uint32_t addr32 = keys_addr + 4u * (1u + (l16 ? 1u : 0u) + (l8 ? 2u : 0u));

    for (int u0 = 0; u0 < total; u0 += kFpThreads) {
        const int u = u0 + tid;
        const bool live = u < total;
        const bool single = !Wide && live && u >= tiles_g;
        int y = 0, x0 = 0, G = 0;
        bool val[4] = { false, false, false, false };
        if (live) {
            if (Wide) {
                const int q = u >> 2;
                G = q % ngpairs; y = q / ngpairs; x0 = 16 * G + s;
#pragma unroll
                for (int k = 0; k < 4; k++) val[k] = x0 + 4 * k < saw;
            } else if (!single) {
                const int q = u >> 2;
                G = q % ngpairs; y = q / ngpairs; x0 = 16 * G + s;
                val[0] = val[1] = (2 * G < ngroups); val[2] = val[3] = (2 * G + 1 < ngroups);
            } else {
                // keep the tile 16-byte aligned: with an odd number of groups the leftovers are positions 2,3
                y = (u - tiles_g) >> 2; x0 = ((ngroups & 1) ? mult8 - 8 : mult8) + s;
#pragma unroll
                for (int k = 0; k < 4; k++) val[k] = (x0 + 4 * k >= mult8) && (x0 + 4 * k < saw);
            }
        }
        // key = sad * 4096 + (row << 6 | col) as ONE IMAD (FMA pipe); a dead position gets 0 * sad + ~0
        uint32_t kmul[4], kadd[4];
#pragma unroll
        for (int k = 0; k < 4; k++) { kmul[k] = val[k] ? 4096u : 0u; kadd[k] = val[k] ? (uint32_t)((y << 6) | (x0 + 4 * k)) : 0xFFFFFFFFu; }
        uint32_t h16[4][4], s32[4][2], s64[4];
#pragma unroll
        for (int k = 0; k < 4; k++) { s64[k] = 0; s32[k][0] = s32[k][1] = 0; }
        // carries of the streaming transpose-reductions (see pair_min)
        uint32_t t3e = 0, t4a = 0, u2a = 0, u3a = 0, v32a = 0;
#pragma unroll
        for (int b = 0; b < 8; b++) {
            uint32_t acc[4][8];
#pragma unroll
            for (int k = 0; k < 4; k++)
#pragma unroll
                for (int j = 0; j < 8; j++) acc[k][j] = 0;
            {   // (threads past the last tile run on tile (0,0); their keys are masked by kmul/kadd)
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
                            for (int k = 0; k < 4; k++) acc[k][c >> 1] = sad4(sw[j], ref[c + k], acc[k][c >> 1]);
                        }
                    }
                }
            }
            // ---- 8x8 PUs of block-row b: per-thread min over its 4 positions, then the transpose-reduction
            {
                uint32_t k8[8];
#pragma unroll
                for (int j = 0; j < 8; j++)
                    k8[j] = min(min(acc[0][j] * kmul[0] + kadd[0], acc[1][j] * kmul[1] + kadd[1]),
                                min(acc[2][j] * kmul[2] + kadd[2], acc[3][j] * kmul[3] + kadd[3]));
                const uint32_t a0 = pair_min(k8[0], k8[1], 16, l16), a1 = pair_min(k8[2], k8[3], 16, l16);
                const uint32_t a2 = pair_min(k8[4], k8[5], 16, l16), a3 = pair_min(k8[6], k8[7], 16, l16);
                const uint32_t t3 = pair_min(pair_min(a0, a1, 8, l8), pair_min(a2, a3, 8, l8), 4, l4);
                if ((b & 1) == 0) t3e = t3;
                else {
                    const uint32_t t4 = pair_min(t3e, t3, 2, l2);
                    if ((b & 3) == 1) t4a = t4;
                    else red_min(addr8 + (b >> 2) * (4u * 32u), pair_min(t4a, t4, 1, l1));    // lane holds one (band, column) PU
                }
            }
            // ---- 16x16: sum of two block-rows
            if ((b & 1) == 0) {
#pragma unroll
                for (int k = 0; k < 4; k++)
#pragma unroll
                    for (int q = 0; q < 4; q++) h16[k][q] = acc[k][2 * q] + acc[k][2 * q + 1];
                continue;
            }
#pragma unroll
            for (int k = 0; k < 4; k++)
#pragma unroll
                for (int q = 0; q < 4; q++) {
                    h16[k][q] += acc[k][2 * q] + acc[k][2 * q + 1];
                    s32[k][q >> 1] += h16[k][q];
                }
            {
                uint32_t c16[4];
#pragma unroll
                for (int q = 0; q < 4; q++)
                    c16[q] = min(min(h16[0][q] * kmul[0] + kadd[0], h16[1][q] * kmul[1] + kadd[1]),
                                 min(h16[2][q] * kmul[2] + kadd[2], h16[3][q] * kmul[3] + kadd[3]));
                const uint32_t u2 = pair_min(pair_min(c16[0], c16[1], 16, l16), pair_min(c16[2], c16[3], 16, l16), 8, l8);
                if ((b & 3) == 1) u2a = u2;
                else {
                    const uint32_t u3 = pair_min(u2a, u2, 4, l4);
                    if (b == 3) u3a = u3;
                    else {
                        uint32_t u4 = pair_min(u3a, u3, 2, l2);
                        u4 = min(u4, __shfl_xor_sync(0xffffffffu, u4, 1));
                        if (!l1) red_min(addr16, u4);
                    }
                }
            }
            if ((b & 3) != 3) continue;
            // ---- 32x32: two PUs of this half of the SB
            {
                // quirk Q4 as a plain key: inside an 8-position group the reference takes the first-min of positions
                // 0..3 ("lo") and of 4..7 ("hi"); lo wins only if strictly smaller and is then reported 4 to the right,
                // hi 4 to the left. That is the minimum of (value, hi-before-lo, position), and the low three key bits
                // (lo: 4 + s, hi: s) are exactly the reported index; groups are ordered (value, row, group) like the
                // reference's strict-less scan. Leftover single positions use slots 8.. and a plain index.
                uint32_t ka32[4], km32[4];
#pragma unroll
                for (int k = 0; k < 4; k++) {
                    uint32_t a;
                    if (Wide) {     // from kadd = row << 6 | x: (group << 3 | lo/hi | s) is x with bit 2 flipped; leftovers: slot 8.., plain index
                        const uint32_t x = kadd[k] & 63u, y7 = (kadd[k] & ~63u) << 1;
                        a = (int)x >= mult8 ? (y7 | ((8u + x - (uint32_t)mult8) << 3)) : (y7 | (x ^ 4u));
                    } else {
                        a = single ? (((uint32_t)y << 7) | ((uint32_t)(8 + x0 + 4 * k - mult8) << 3))
                                   : (((uint32_t)y << 7) | ((uint32_t)(2 * G + (k >> 1)) << 3) | ((k & 1) ? 0u : 4u) | (uint32_t)s);
                    }
                    ka32[k] = val[k] ? a : 0xFFFFFFFFu;
                    km32[k] = kmul[k] << 1;
                }
                uint32_t k32[2];
#pragma unroll
                for (int q = 0; q < 2; q++) {
                    k32[q] = min(min(s32[0][q] * km32[0] + ka32[0], s32[1][q] * km32[1] + ka32[1]),
                                 min(s32[2][q] * km32[2] + ka32[2], s32[3][q] * km32[3] + ka32[3]));
#pragma unroll
                    for (int k = 0; k < 4; k++) { s64[k] += s32[k][q]; s32[k][q] = 0; }
                }
                const uint32_t v32 = pair_min(k32[0], k32[1], 16, l16);
                if (b == 3) v32a = v32;
                else {
                    uint32_t w = pair_min(v32a, v32, 8, l8);
                    w = min(w, __shfl_xor_sync(0xffffffffu, w, 4));
                    w = min(w, __shfl_xor_sync(0xffffffffu, w, 2));
                    w = min(w, __shfl_xor_sync(0xffffffffu, w, 1));
                    if ((lane & 7) == 0) red_min(addr32, w);
                }
            }
            if (b != 7) continue;
            {
                const uint32_t key = min(min(s64[0] * kmul[0] + kadd[0], s64[1] * kmul[1] + kadd[1]),
                                         min(s64[2] * kmul[2] + kadd[2], s64[3] * kmul[3] + kadd[3]));
                const uint32_t m = __reduce_min_sync(0xffffffffu, key);
                if (lane == 0) red_min(keys_addr, m);
            }
        }
    }
    __syncthreads();

    // ---- decode the winning keys into p_sb_best_sad / p_sb_best_mv form
    if (tid < kNumPu) {
        const uint32_t key = keys[tid];
        uint32_t v; int x, y;
        if (tid >= 1 && tid <= 4) {
            v = key >> 13; y = (key >> 7) & 63;
            const int slot = (key >> 3) & 15, idx = key & 7;
            x = slot < 8 ? slot * 8 + idx : mult8 + slot - 8;
        } else {
            v = key >> 12; y = (key >> 6) & 63; x = key & 63;
        }
        if (!Wide) {
            scr.best_sad[list][tid] = v << 1;
            scr.best_mv[list][tid] = make_mv((int16_t)((sax + x) * 4), (int16_t)((say + y) * 4));
        } else if (key != 0xFFFFFFFFu) {
            // merge in the reference's visiting order: row, then 8-position group, then (32x32: the quirk's in-group rank =
            // the low key bits; others: the column). Sub-areas start on multiples of 64, so groups keep their identity.
            const uint32_t o = (tid >= 1 && tid <= 4) ? (((uint32_t)(ty0 + y) << 8) | ((uint32_t)((tx0 >> 3) + ((key >> 3) & 15)) << 3) | (key & 7))
                                                      : (((uint32_t)(ty0 + y) << 8) | (uint32_t)(tx0 + x));
            if (v < merged_v[tid] || (v == merged_v[tid] && o < merged_o[tid])) {
                merged_v[tid] = v; merged_o[tid] = o;
                scr.best_sad[list][tid] = v << 1;
                scr.best_mv[list][tid] = make_mv((int16_t)((sax + x) * 4), (int16_t)((say + y) * 4));
            }
        }
    }
    if (Wide) __syncthreads();                                    // before the next sub-area's window and key reset
    }   // sub-areas
}

cudaError_t configure_fullpel() {
    cudaError_t e = cudaFuncSetAttribute(fullpel_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kFpSmemBytes);
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(fullpel_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kFpSmemBytes);
}

void launch_fullpel_ext(const DJob* d_jobs, int n_jobs, int max_sb, int list, bool wide, cudaStream_t st);   // fullpel_ext.cu

void launch_fullpel(const DJob* d_jobs, int n_jobs, int max_sb, int list, bool ext, bool wide, cudaStream_t st) {
    if (ext)       launch_fullpel_ext(d_jobs, n_jobs, max_sb, list, wide, st);
    else if (wide) fullpel_kernel<true><<<dim3(max_sb, n_jobs), kFpThreads, kFpSmemBytes, st>>>(d_jobs, list);
    else           fullpel_kernel<false><<<dim3(max_sb, n_jobs), kFpThreads, kFpSmemBytes, st>>>(d_jobs, list);
}

}  // namespace svtme
