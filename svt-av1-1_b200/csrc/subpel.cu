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
#include "pack.cuh"

namespace svtme {

constexpr int kSpThreads = 256;
// Plane geometry: the window is the bounding box of the 256 (level, block) positions, at most (area - 1) + 56 + 15 rows and
// the same + 6 bytes (+ one spare word for the last B taps) per row.
//   kGeomSmall   88 rows x 28 words: boxes of SBs whose best full-pel positions lie within about 17 rows / 30 columns of each
//                other (coherent motion: all but 0.03-0.3 % of the SBs of the bench sequence) -- 53 KB of shared memory and
//                64 registers, FOUR CTAs per SM. An SB whose box does not fit is appended to a work list instead and
//                searched by the kGeomFull instantiation afterwards.
//   kGeomFull    134 x 36: any box of a search area up to 64 x 64 (two CTAs per SM); persistent CTAs over the work list
//   kGeomWide    198 x 52: search areas up to 127 x 127 (MIN(search_area, 127), ME.c:6644-6645), one CTA per SM
enum { kGeomSmall = 0, kGeomFull = 1, kGeomWide = 2 };
template <int Geom> struct PlaneGeom {
    static constexpr int kRows = Geom == kGeomWide ? 198 : Geom == kGeomFull ? 134 : 88;
    static constexpr int kPitchW = Geom == kGeomWide ? 52 : Geom == kGeomFull ? 36 : 28;          // words
    static constexpr int kWords = kRows * kPitchW + ((kRows - 1) & ~15) + 4;                    // + room for the row skew below
    static constexpr int kRawPitch = kPitchW + 8;                                               // words per raw window row: 16-byte granules (3 + pitch + 1 words at most)
    static constexpr int kCtasPerSm = Geom == kGeomWide ? 1 : Geom == kGeomFull ? 2 : 4;
};
// Word offset of window row r. Rows 8 apart would land on the same bank (8 * 36 = 9 * 32); the (monotonic) skew
// moves each group of eight rows by 1 or 15 banks more than the previous one, so that the 8x4 blocks a warp reads
// together (two words apart along x, eight rows apart along y) touch 32 different banks when they share a motion
// vector: bank offsets of the four block rows are {0, 1, 16, 17} or {0, 15, 16, 31}.
template <int Geom> __device__ __forceinline__ int plane_row(int r) { return r * PlaneGeom<Geom>::kPitchW + ((r >> 3) & 1) + (r & ~15); }
static_assert(PlaneGeom<0>::kRows * PlaneGeom<0>::kRawPitch <= 2 * PlaneGeom<0>::kWords && PlaneGeom<1>::kRows * PlaneGeom<1>::kRawPitch <= 2 * PlaneGeom<1>::kWords &&
              PlaneGeom<2>::kRows * PlaneGeom<2>::kRawPitch <= 2 * PlaneGeom<2>::kWords, "the raw window must fit the storage of planes H and J");
constexpr int kMargin = 3;              // window starts 3 pixels left/above the search-area origin

template <int NPU, int Geom>
struct SpSharedT {
    uint32_t plane[4][PlaneGeom<Geom>::kWords];   // A, B, H, J
    uint32_t src[64 * 16];
    uint32_t sum_ssd[9][NPU];           // per (candidate, PU) distortion sums, accumulated with shared atomics
    uint32_t sum_sad[9][NPU];
    uint32_t pu_sad[NPU], pu_mv[NPU], pu_ssd[NPU];
    int32_t  pu_dir[NPU];
    uint32_t pu_imv[NPU];               // MV at entry of the current stage
    uint32_t pu_win[NPU];               // winning half-pel candidate (1..8) of the PU, 0 = the integer position stays
    uint32_t rowoff[PlaneGeom<Geom>::kRows + 2];  // plane_row() of every window row: one LDS instead of five integer instructions
    int32_t  bbox[4];                   // min X, min Y, max -X... see the kernel: bounding box of the blocks' window positions
};

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
__device__ __forceinline__ uint2 load8(const uint32_t* plane, const uint32_t* rowoff, int r, int c) {
    const uint32_t* w = plane + rowoff[r] + (c >> 2);
    const uint32_t sh = (c & 3) * 8;
    const uint32_t a = w[0], b = w[1], d = w[2];
    return make_uint2(__funnelshift_r(a, b, sh), __funnelshift_r(b, d, sh));
}
// Quirk Q5 squares the per-byte difference (a - b) mod 256 read as int8. |a - b| read as int8 has the same square
// (for a < b the two bytes are negatives of each other mod 256, and -128 squares like 128), so one VABSDIFF4 replaces
// the five-instruction byte-wise subtraction.
__device__ __forceinline__ uint32_t wrap_ssd4(uint32_t a, uint32_t b, uint32_t acc) {
    const uint32_t t = __vabsdiffu4(a, b);
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
// The same filter from tap words already split into their even / odd byte lanes (a vertical pass slides over the rows:
// every row word is split once instead of four times) ...
struct Lanes { uint32_t e, o; };
__device__ __forceinline__ Lanes split_lanes(uint32_t t) {
    Lanes l; l.e = __byte_perm(t, 0, 0x4240); l.o = __byte_perm(t, 0, 0x4341); return l;
}
__device__ __forceinline__ uint32_t filt4_split(const Lanes& a, const Lanes& b, const Lanes& c, const Lanes& d) {
    return __byte_perm(filt4_lanes(a.e, b.e, c.e, d.e), filt4_lanes(a.o, b.o, c.o, d.o), 0x6240);
}
// ... and along a row: with left = bytes x-1..x+2, w1 = x..x+3, w2 = x+4..x+7 the eight lane words of the four byte-shifted
// taps are only five different byte pairs (x-1,x+1) (x,x+2) (x+1,x+3) (x+2,x+4) (x+3,x+5)
__device__ __forceinline__ uint32_t filt4_row(uint32_t left, uint32_t w1, uint32_t w2) {
    const uint32_t f16 = __funnelshift_r(w1, w2, 16);
    const uint32_t pm = __byte_perm(left, 0, 0x4240), p0 = __byte_perm(w1, 0, 0x4240), p1 = __byte_perm(w1, 0, 0x4341);
    const uint32_t p2 = __byte_perm(f16, 0, 0x4240), p3 = __byte_perm(f16, 0, 0x4341);
    return __byte_perm(filt4_lanes(pm, p0, p1, p2), filt4_lanes(p0, p1, p2, p3), 0x6240);
}
// validity of quarter-pel candidate k (order L R T B TL TR BR BL) given the best half-pel direction (ME.c:2859-2881):
// the candidate itself or one of its two neighbours in the 8-cycle TL,T,TR,R,BR,B,BL,L; mirrored when the
// half-pel MV has a fractional part (method != 0)
__device__ __forceinline__ bool quarter_valid(int method, int dir, int k) {
    const int q8 = k == 0 ? D_L : k == 1 ? D_R : k == 2 ? D_T : k == 3 ? D_B : k == 4 ? D_TL : k == 5 ? D_TR : k == 6 ? D_BR : D_BL;
    const int centre = method ? ((q8 + 4) & 7) : q8;
    return dir == centre || dir == ((centre + 1) & 7) || dir == ((centre + 7) & 7);
}
// The same as a bit mask over k (exactly three bits set), from a packed table: byte d of the constant is the mask for
// direction d with method == 0; a fractional half-pel MV (method != 0) mirrors the direction (d + 4).
__device__ __forceinline__ uint32_t quarter_mask(int method, int dir) {
    const unsigned long long table = 0x9189C84A62263415ull;
    return (uint32_t)(table >> (8 * ((dir + (method ? 4 : 0)) & 7))) & 0xFFu;
}

// Lanes of a warp = 8x4 blocks (bx = lane & 7, by = lane >> 3) of one level L (warp-uniform). Sum over the blocks of
// each PU that lie in this warp: all 32 for 64x64, 4x4 for 32x32, 2x2 for 16x16, the block itself for 8x8.
__device__ __forceinline__ uint32_t pu_sum(uint32_t v, int L) {
    if (L == 0) return __reduce_add_sync(0xffffffffu, v);
    if (L == 3) return v;
    v += __shfl_xor_sync(0xffffffffu, v, 1); v += __shfl_xor_sync(0xffffffffu, v, 8);
    if (L == 1) { v += __shfl_xor_sync(0xffffffffu, v, 2); v += __shfl_xor_sync(0xffffffffu, v, 16); }
    return v;
}
__device__ __forceinline__ bool pu_leader(int L, int lane) {
    return L == 0 ? lane == 0 : L == 1 ? (lane & 0x1B) == 0 : L == 2 ? (lane & 0x09) == 0 : true;
}
// The same for any of the 14 PU classes of the 209-PU mode (wb x hb blocks, warp-uniform): a warp holds block rows
// 4w .. 4w+3, so PUs eight blocks tall are shared by the two warps of the level (the caller adds those with atomics).
__device__ __forceinline__ uint32_t pu_sum_any(uint32_t v, int wb, int hb) {
    if (wb > 1) v += __shfl_xor_sync(0xffffffffu, v, 1);
    if (wb > 2) v += __shfl_xor_sync(0xffffffffu, v, 2);
    if (wb > 4) v += __shfl_xor_sync(0xffffffffu, v, 4);
    if (hb > 1) v += __shfl_xor_sync(0xffffffffu, v, 8);
    if (hb > 2) v += __shfl_xor_sync(0xffffffffu, v, 16);
    return v;
}
__device__ __forceinline__ bool pu_leader_any(int wb, int hb, int lane) {
    return (lane & ((wb - 1) | ((min(hb, 4) - 1) << 3))) == 0;
}

// Ext = the 209-PU mode: 14 PU classes ("levels") instead of 4; the non-square PUs live in SbScratchExt. Every level
// is 64 8x8-block tasks, so a warp always works on one level.
#define PROW(r) ((int)sh.rowoff[r])
// One SB. Returns false (and has done nothing else) when the SB's window does not fit this instantiation's planes: only the
// kGeomSmall one can say so, the caller then leaves the SB to the kGeomFull pass.
template <bool Ext, int Geom>
__device__ __forceinline__ bool subpel_sb(const DJob& job, const int sb, const int list, uint8_t* smem_raw) {
    constexpr int NPU = Ext ? kMaxPu : kNumPu, NLEV = Ext ? kMaxLevels : kNumLevels;
    constexpr int kPlRows = PlaneGeom<Geom>::kRows, kRawPitchSp = PlaneGeom<Geom>::kRawPitch;
    using SpShared = SpSharedT<NPU, Geom>;
    SpShared& sh = *reinterpret_cast<SpShared*>(smem_raw);
    const SvtMeParams& p = job.prm;
    auto block_pu = [](int L, int bx, int by) { if constexpr (Ext) return pu_of_block_any(L, bx, by); else return pu_of_block(L, bx, by); };
    // 8x8 half/quarter-pel is switched off by cu8x8_mode 1 (disable8x8CuInMeFlag); the non-square classes never are
    auto pu_enabled = [&](int pu) { return pu < 21 || pu >= kNumPu || p.cu8x8_mode != 1; };
    const int tid = threadIdx.x;
    const int ox = (sb % job.sb_cols) * 64, oy = (sb / job.sb_cols) * 64;
    SbScratch& scr = job.scratch[sb];
    const DPlane& rp = job.ref[list].full;
    // ---- stage: source SB (asynchronous copies: first read in the half-pel phase, three barriers from here), PU state
    for (int t = tid; t < 64 * 16; t += kSpThreads) {
        const uint32_t dst = (uint32_t)__cvta_generic_to_shared(&sh.src[t]);
        asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" :: "r"(dst), "l"(job.cur.full.at(ox + 4 * (t & 15), oy + (t >> 4))) : "memory");
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    if (tid < kNumPu) { sh.pu_sad[tid] = scr.best_sad[list][tid]; sh.pu_mv[tid] = scr.best_mv[list][tid]; sh.pu_imv[tid] = sh.pu_mv[tid]; }
    if constexpr (Ext) {
        const SbScratchExt& scx = job.scratch_ext[sb];
        if (tid >= kNumPu && tid < NPU) { sh.pu_sad[tid] = scx.best_sad[list][tid - kNumPu]; sh.pu_mv[tid] = scx.best_mv[list][tid - kNumPu]; sh.pu_imv[tid] = sh.pu_mv[tid]; }
    }
    if (tid < 4) sh.bbox[tid] = 0x7FFFFFFF;
    if (tid < kPlRows + 2) sh.rowoff[tid] = (uint32_t)plane_row<Geom>(tid);
    __syncthreads();
    // The refinement only reads the window around the 256 (level, 8x8 block) positions the full-pel search chose
    // (every later MV stays within one integer pixel of them), so the window that is staged and interpolated is their
    // bounding box, not the whole search area: with coherent motion that is a fraction of it. The box becomes the new
    // window origin (sax, say) and extent (ww words x wr rows); everything below is unchanged.
    int sax = scr.sa_x[list], say = scr.sa_y[list];
    for (int t = tid; t < NLEV * 64; t += kSpThreads) {
        const int blk = t & 63, L = t >> 6, bx = blk & 7, by = blk >> 3;
        const uint32_t mv = sh.pu_mv[block_pu(L, bx, by)];
        const int X = kMargin + (mv_x(mv) >> 2) - sax + bx * 8, Y = kMargin + (mv_y(mv) >> 2) - say + by * 8;
        const int xmin = __reduce_min_sync(0xffffffffu, X), ymin = __reduce_min_sync(0xffffffffu, Y);
        const int xneg = __reduce_min_sync(0xffffffffu, -X), yneg = __reduce_min_sync(0xffffffffu, -Y);
        if ((tid & 31) == 0) { atomicMin(&sh.bbox[0], xmin); atomicMin(&sh.bbox[1], ymin); atomicMin(&sh.bbox[2], xneg); atomicMin(&sh.bbox[3], yneg); }
    }
    __syncthreads();
    const int dx = (sh.bbox[0] - kMargin) & ~3, dy = sh.bbox[1] - kMargin;   // window origin moves by (dx, dy), dx word-aligned
    sax += dx; say += dy;
    const int ww = ((-sh.bbox[2] - dx + 15) >> 2) + 1;   // window words per row actually used (+1: taps of the last B word)
    const int wr = -sh.bbox[3] - dy + 12;                 // window rows: 8 block rows + one above, three below for the taps
    if (Geom == kGeomSmall && (ww > PlaneGeom<Geom>::kPitchW || wr > kPlRows)) {     // CTA-uniform: the box is wider than these planes
        asm volatile("cp.async.wait_group 0;" ::: "memory");                     // (the source SB's copies target this CTA's memory)
        return false;
    }

    // ---- plane A
    const uint32_t inv_ww = 0xFFFFFFFFu / (uint32_t)ww + 1u;   // t / ww == umulhi(t, inv_ww) for t * ww < 2^32
    // The window lands raw (16-byte cp.async granules from the aligned address below its first byte: one L2 round trip
    // for the whole window instead of one per loop iteration) in the storage of planes H and J, which are written only
    // after A and B are complete; the threads then shift it into plane A's skewed rows from shared memory.
    {
        const uint8_t* g0 = rp.at(ox + sax - kMargin, oy + say - kMargin);
        const uint32_t m16 = (uint32_t)(uintptr_t)g0 & 15u;           // the same for every row: the pitch is a multiple of 64
        const int q = (int)(m16 >> 2), ngran = (q + ww + 1 + 3) >> 2; // words q .. q + ww of a raw row are read below
        const uint8_t* ga = g0 - m16;
        uint32_t* raw = sh.plane[S_H];                                // kRawPitchSp words per row, <= 2 planes in total
        const uint32_t raw_addr = (uint32_t)__cvta_generic_to_shared(raw);
        for (int t = tid; t < wr * ngran; t += kSpThreads) {
            const int r = t / ngran, g = t - r * ngran;
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"(raw_addr + (uint32_t)(r * kRawPitchSp * 4 + g * 16)),
                         "l"(ga + (size_t)r * rp.stride + g * 16) : "memory");
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
        asm volatile("cp.async.wait_group 0;" ::: "memory");
        __syncthreads();
        const uint32_t bs = (m16 & 3u) * 8u;
        for (int t = tid; t < wr * ww; t += kSpThreads) {
            const int r = (int)__umulhi((uint32_t)t, inv_ww), c = t - r * ww;
            const uint32_t* w = raw + r * kRawPitchSp + q + c;
            sh.plane[S_A][PROW(r) + c] = __funnelshift_r(w[0], w[1], bs);
        }
    }
    __syncthreads();
    if (p.use_subpel == 1) {
        // ---- plane B (horizontal filter of A): B(x) = f(A[x-1], A[x], A[x+1], A[x+2]) on every word but the last
        for (int t = tid; t < wr * ww; t += kSpThreads) {
            const int r = (int)__umulhi((uint32_t)t, inv_ww), c = t - r * ww;
            const uint32_t* a = sh.plane[S_A] + PROW(r) + c;
            if (c < ww - 1) {
                const uint32_t w1 = a[0], w2 = a[1];
                const uint32_t left = c ? __funnelshift_r(a[-1], w1, 24) : w1 << 8;
                sh.plane[S_B][PROW(r) + c] = filt4_row(left, w1, w2);
            }
        }
        __syncthreads();
        // ---- planes H (vertical filter of A) and J (vertical filter of B): each thread slides a four-row window down
        // one word column over its share of the rows, so every row is read once: X(y) = f(X[y-1], X[y], X[y+1], X[y+2])
        {
            const int chunk = (int)__umulhi((uint32_t)tid, inv_ww), c = tid - chunk * ww;
            const int n_chunks = kSpThreads / ww, rows = wr - 3;                 // output rows 1 .. wr-3
            const int per = (rows + n_chunks - 1) / n_chunks;
            const int r0 = 1 + chunk * per, r1 = min(r0 + per, wr - 2);
            if (chunk < n_chunks && r0 < r1) {
                const uint32_t* pa = sh.plane[S_A] + c; const uint32_t* pb = sh.plane[S_B] + c;
                Lanes a0 = split_lanes(pa[PROW(r0 - 1)]), a1 = split_lanes(pa[PROW(r0)]), a2 = split_lanes(pa[PROW(r0 + 1)]);
                Lanes b0 = split_lanes(pb[PROW(r0 - 1)]), b1 = split_lanes(pb[PROW(r0)]), b2 = split_lanes(pb[PROW(r0 + 1)]);
#pragma unroll 2
                for (int r = r0; r < r1; r++) {
                    const int o2 = PROW(r + 2), o = PROW(r);
                    const Lanes a3 = split_lanes(pa[o2]), b3 = split_lanes(pb[o2]);
                    sh.plane[S_H][o + c] = filt4_split(a0, a1, a2, a3);
                    sh.plane[S_J][o + c] = filt4_split(b0, b1, b2, b3);
                    a0 = a1; a1 = a2; a2 = a3; b0 = b1; b1 = b2; b2 = b3;
                }
            }
        }
        __syncthreads();

        for (int t = tid; t < 9 * NPU; t += kSpThreads) { (&sh.sum_ssd[0][0])[t] = 0; (&sh.sum_sad[0][0])[t] = 0; }
        asm volatile("cp.async.wait_group 0;" ::: "memory");             // the source SB has landed
        __syncthreads();
        // ---- half-pel: one thread per (level, 8x8 block) walks the ten plane rows its nine candidates share
        // (L/R read the same three B words per row, T/B the same H rows, the four diagonals the same J rows and words);
        // the wrapping SSD is summed per (candidate, PU).
        // Only the SAD of the WINNING candidate of a PU is ever used (ME.c:1934-1950), so the candidates' SADs
        // are not computed here at all: after the decision every block recomputes the SAD of its PU's winner (eight rows of
        // one plane) and that is summed over the PU.
        for (int t = tid; t < NLEV * 64; t += kSpThreads) {
            const int blk = t & 63, L = t >> 6, bx = blk & 7, by = blk >> 3;
            const int pu = block_pu(L, bx, by);
            const uint32_t mv = sh.pu_imv[pu];
            const int X = kMargin + (mv_x(mv) >> 2) - sax + bx * 8, Y = kMargin + (mv_y(mv) >> 2) - say + by * 8;
            const int wa = X >> 2, wb = (X - 1) >> 2;
            const uint32_t sa = (X & 3) * 8, sb_ = ((X - 1) & 3) * 8;
            uint2 src[8];
#pragma unroll
            for (int r = 0; r < 8; r++) src[r] = *reinterpret_cast<const uint2*>(&sh.src[(by * 8 + r) * 16 + bx * 2]);
            uint32_t ssd[9];
#pragma unroll
            for (int k = 0; k < 9; k++) ssd[k] = 0;
#define HP_ACC(k, lo, hi, s)                                                                           \
            do { ssd[k] = wrap_ssd4((s).x, lo, ssd[k]); ssd[k] = wrap_ssd4((s).y, hi, ssd[k]); } while (0)
#pragma unroll
            for (int rr = -1; rr < 8; rr++) {
                const int row = PROW(Y + rr);
                const uint32_t* ph = sh.plane[S_H] + row + wa;
                const uint32_t h0 = ph[0], h1 = ph[1], h2 = ph[2];
                const uint32_t hl = __funnelshift_r(h0, h1, sa), hh = __funnelshift_r(h1, h2, sa);
                const uint32_t* pj = sh.plane[S_J] + row + wb;
                const uint32_t j0 = pj[0], j1 = pj[1], j2 = pj[2];
                const uint32_t jl0 = __funnelshift_r(j0, j1, sb_), jl1 = __funnelshift_r(j1, j2, sb_);     // from X - 1
                const uint32_t jr0 = __funnelshift_r(jl0, jl1, 8), jr1 = __funnelshift_r(jl1, j2 >> sb_, 8); // from X
                if (rr < 7) {                                  // candidates one row up: T, TL, TR against source row rr + 1
                    HP_ACC(3, hl, hh, src[rr + 1]); HP_ACC(5, jl0, jl1, src[rr + 1]); HP_ACC(6, jr0, jr1, src[rr + 1]);
                }
                if (rr >= 0) {                                 // B, BR, BL, the integer position and L, R against source row rr
                    HP_ACC(4, hl, hh, src[rr]); HP_ACC(7, jr0, jr1, src[rr]); HP_ACC(8, jl0, jl1, src[rr]);
                    const uint32_t* pa = sh.plane[S_A] + row + wa;
                    const uint32_t a0 = pa[0], a1 = pa[1], a2 = pa[2];
                    const uint32_t al = __funnelshift_r(a0, a1, sa), ah = __funnelshift_r(a1, a2, sa);
                    ssd[0] = wrap_ssd4(src[rr].x, al, ssd[0]); ssd[0] = wrap_ssd4(src[rr].y, ah, ssd[0]);
                    const uint32_t* pb = sh.plane[S_B] + row + wb;
                    const uint32_t b0 = pb[0], b1 = pb[1], b2 = pb[2];
                    const uint32_t bl0 = __funnelshift_r(b0, b1, sb_), bl1 = __funnelshift_r(b1, b2, sb_);
                    const uint32_t br0 = __funnelshift_r(bl0, bl1, 8), br1 = __funnelshift_r(bl1, b2 >> sb_, 8);
                    HP_ACC(1, bl0, bl1, src[rr]); HP_ACC(2, br0, br1, src[rr]);
                }
            }
#undef HP_ACC
            // a warp holds 8x4 blocks of ONE level: sum the blocks of each PU inside the warp, one lane per PU writes
            if constexpr (Ext) {
                const int pwb = kLevelWb[L], phb = kLevelHb[L];
                // quirk Q10: for 8-wide PUs taller than 8 the reference ranks the candidates on the top eight rows only
                const bool ranked = !(pwb == 1 && phb > 1 && (by % phb) != 0);
                const bool leader = pu_leader_any(pwb, phb, tid & 31);
#pragma unroll
                for (int k = 0; k < 9; k++) {
                    const uint32_t a = pu_sum_any(ranked ? ssd[k] : 0u, pwb, phb);
                    if (leader) {
                        if (phb == 8) atomicAdd(&sh.sum_ssd[k][pu], a);   // two warps share the PU
                        else sh.sum_ssd[k][pu] = a;
                    }
                }
            } else {
                const bool leader = pu_leader(L, tid & 31);
#pragma unroll
                for (int k = 0; k < 9; k++) {
                    const uint32_t a = pu_sum(ssd[k], L);
                    if (leader) {
                        if (L == 0) atomicAdd(&sh.sum_ssd[k][0], a);      // two warps share PU 0
                        else sh.sum_ssd[k][pu] = a;
                    }
                }
            }
        }
        __syncthreads();
        // ---- per PU: nine candidates in the reference's order, strict '<' on the SSD  (ME.c:1909-2239)
        if (tid < NPU) sh.pu_win[tid] = 0;
        if (tid < NPU && pu_enabled(tid)) {
            uint32_t dist[9];
#pragma unroll
            for (int k = 0; k < 9; k++) dist[k] = sh.sum_ssd[k][tid];
            const uint32_t imv = sh.pu_imv[tid];
            uint32_t bssd = dist[0], bmv = imv;
            int kw = 0;
            for (int k = 1; k < 9; k++)
                if (dist[k] < bssd) { bssd = dist[k]; kw = k; bmv = make_mv((int16_t)(mv_x(imv) + kHalf[k][3]), (int16_t)(mv_y(imv) + kHalf[k][4])); }
            uint32_t best = dist[1];
            for (int k = 2; k < 9; k++) best = min(best, dist[k]);
            // tie order L R T B TL TR BL BR
            const int dir = best == dist[1] ? D_L : best == dist[2] ? D_R : best == dist[3] ? D_T : best == dist[4] ? D_B :
                            best == dist[5] ? D_TL : best == dist[6] ? D_TR : best == dist[8] ? D_BL : D_BR;
            sh.pu_ssd[tid] = bssd; sh.pu_mv[tid] = bmv; sh.pu_dir[tid] = dir;
            sh.pu_win[tid] = (uint32_t)kw;
            if (kw) sh.pu_sad[tid] = 0;                       // re-summed below from the blocks' SADs of candidate kw
        }
        __syncthreads();
        // ---- SAD of the winning half-pel candidate, summed over the PU's blocks
        for (int t = tid; t < NLEV * 64; t += kSpThreads) {
            const int blk = t & 63, L = t >> 6, bx = blk & 7, by = blk >> 3;
            const int pu = block_pu(L, bx, by);
            const int kw = (int)sh.pu_win[pu];
            uint32_t v = 0;
            if (kw) {                                          // the block's SAD at the winning candidate's plane position
                const uint32_t mv = sh.pu_imv[pu];
                const int x = kMargin + (mv_x(mv) >> 2) - sax + bx * 8 + kHalf[kw][1], y = kMargin + (mv_y(mv) >> 2) - say + by * 8 + kHalf[kw][2];
                const uint32_t* pl = sh.plane[kHalf[kw][0]];
#pragma unroll
                for (int r = 0; r < 8; r++) {
                    const uint2 c = load8(pl, sh.rowoff, y + r, x);
                    const uint2 sv = *reinterpret_cast<const uint2*>(&sh.src[(by * 8 + r) * 16 + bx * 2]);
                    v = sad4(sv.x, c.x, v); v = sad4(sv.y, c.y, v);
                }
            }
            if constexpr (Ext) {
                const int pwb = kLevelWb[L], phb = kLevelHb[L];
                v = pu_sum_any(v, pwb, phb);
                if (kw && pu_leader_any(pwb, phb, tid & 31)) { if (phb == 8) atomicAdd(&sh.pu_sad[pu], v); else sh.pu_sad[pu] = v; }
            } else {
                v = pu_sum(v, L);
                if (kw && pu_leader(L, tid & 31)) { if (L == 0) atomicAdd(&sh.pu_sad[0], v); else sh.pu_sad[pu] = v; }
            }
        }
        __syncthreads();                                       // the loop above reads pu_imv
        if (tid < NPU) sh.pu_imv[tid] = sh.pu_mv[tid];
        for (int t = tid; t < 9 * NPU; t += kSpThreads) { (&sh.sum_ssd[0][0])[t] = 0; (&sh.sum_sad[0][0])[t] = 0; }
        __syncthreads();

        // ---- quarter-pel: one thread per (level, 8x8 block); exactly three of the eight positions are valid for a
        // given half-pel direction -> partial true-SSD and SAD of the averaged plane pair, source rows kept in registers
        for (int t = tid; t < NLEV * 64; t += kSpThreads) {
            const int blk = t & 63, L = t >> 6, bx = blk & 7, by = blk >> 3;
            const int pu = block_pu(L, bx, by);
            const bool in_area = (L != 0) || (bx < 4 && by < 4);          // quirk Q6: 64x64 refined on its top-left 32x32
            const bool active = in_area && (L != 3 || p.cu8x8_mode != 1);
            const uint32_t mv = sh.pu_imv[pu];
            const int xm = mv_x(mv), ym = mv_y(mv), dir = sh.pu_dir[pu];
            const int method = (ym & 2) + ((xm & 2) >> 1);
            const int X = kMargin + ((xm + 2) >> 2) - sax + bx * 8, Y = kMargin + ((ym + 2) >> 2) - say + by * 8;
            uint2 src[8];
#pragma unroll
            for (int r = 0; r < 8; r++) src[r] = *reinterpret_cast<const uint2*>(&sh.src[(by * 8 + r) * 16 + bx * 2]);
            int pwb = 1, phb = 1;
            if constexpr (Ext) { pwb = kLevelWb[L]; phb = kLevelHb[L]; }
            // lane 0 of the 64x64 level is inside the area
            const bool leader = active && (Ext ? pu_leader_any(pwb, phb, tid & 31) : pu_leader(L, tid & 31));
            uint32_t kmask = quarter_mask(method, dir);                    // inactive lanes may hold a stale dir: still three bits
#pragma unroll 1
            for (int j = 0; j < 3; j++) {
                const int kk = __ffs(kmask) - 1;
                kmask &= kmask - 1;
                uint32_t ssd = 0, sad = 0;
                if (active) {
                    const int8_t* q = kQuarter[method][kk];
                    const uint32_t* p1 = sh.plane[q[0]]; const uint32_t* p2 = sh.plane[q[3]];
                    const int x1 = X + q[1], y1 = Y + q[2], x2 = X + q[4], y2 = Y + q[5];
#pragma unroll
                    for (int r = 0; r < 8; r++) {
                        const uint2 c1 = load8(p1, sh.rowoff, y1 + r, x1), c2 = load8(p2, sh.rowoff, y2 + r, x2);
                        const uint32_t a0 = __vavgu4(c1.x, c2.x), a1 = __vavgu4(c1.y, c2.y);
                        const uint32_t d0 = __vabsdiffu4(src[r].x, a0), d1 = __vabsdiffu4(src[r].y, a1);   // SSD and SAD from one word
                        ssd = __dp4a(d0, d0, ssd); ssd = __dp4a(d1, d1, ssd);
                        sad = __dp4a(d0, 0x01010101u, sad); sad = __dp4a(d1, 0x01010101u, sad);
                    }
                }
                if constexpr (Ext) {
                    ssd = pu_sum_any(ssd, pwb, phb); sad = pu_sum_any(sad, pwb, phb);
                    if (leader) {
                        if (L != 0 && phb == 8) { atomicAdd(&sh.sum_ssd[kk][pu], ssd); atomicAdd(&sh.sum_sad[kk][pu], sad); }   // both warps hold blocks
                        else { sh.sum_ssd[kk][pu] = ssd; sh.sum_sad[kk][pu] = sad; }
                    }
                } else {
                    ssd = pu_sum(ssd, L); sad = pu_sum(sad, L);
                    if (leader) { sh.sum_ssd[kk][pu] = ssd; sh.sum_sad[kk][pu] = sad; }   // 64x64: only the first warp has active blocks
                }
            }
        }
        __syncthreads();
        if (tid < NPU && pu_enabled(tid)) {
            const uint32_t imv = sh.pu_imv[tid];
            const int method = (mv_y(imv) & 2) + ((mv_x(imv) & 2) >> 1), dir = sh.pu_dir[tid];
            uint32_t bssd = sh.pu_ssd[tid], bsad = sh.pu_sad[tid], bmv = imv;
            const uint32_t kmask = quarter_mask(method, dir);
#pragma unroll
            for (int k = 0; k < 8; k++) {
                if (!((kmask >> k) & 1u)) continue;
                const uint32_t a = sh.sum_ssd[k][tid], b = sh.sum_sad[k][tid];
                if (a < bssd) { bssd = a; bsad = b; bmv = make_mv((int16_t)(mv_x(imv) + kQmv[k][0]), (int16_t)(mv_y(imv) + kQmv[k][1])); }
            }
            sh.pu_sad[tid] = bsad; sh.pu_mv[tid] = bmv;
        }
        __syncthreads();
        if (tid < kNumPu) { scr.best_sad[list][tid] = sh.pu_sad[tid]; scr.best_mv[list][tid] = sh.pu_mv[tid]; }
        if constexpr (Ext) {
            SbScratchExt& scx = job.scratch_ext[sb];
            if (tid >= kNumPu && tid < NPU) { scx.best_sad[list][tid - kNumPu] = sh.pu_sad[tid]; scx.best_mv[list][tid - kNumPu] = sh.pu_mv[tid]; }
        }
    } else {
        // planes B/H/J are not needed: integer MVs only use plane A below
        asm volatile("cp.async.wait_group 0;" ::: "memory");
    }

    // ---- final uni-prediction blocks of this list for every PU level. List 0 stores them (one 64x64 byte image per level);
    // list 1 meets them again right here: bi-prediction SAD(src, (p0 + p1 + 1) >> 1) per 8x8 block while its own blocks are in
    // registers, then the PU rows are ordered and packed (pack.cuh) -- the list-1 blocks never travel through memory and the
    // separate bi-prediction kernel only remains for pictures with one list (BiPredictionSearch, ME.c:4933-5342, 6973-7146).
    if (p.num_lists == 2) {
        uint32_t* part = &sh.sum_ssd[0][0];                     // [NLEV][64] bi-prediction SADs; the candidate sums are dead by now
        static_assert(9 * NPU >= NLEV * 64, "part[] fits the candidate sums");
        if (list == 1) __syncthreads();                         // every thread is done with sum_ssd; the source SB has landed for all
        for (int t = tid; t < NLEV * 64; t += kSpThreads) {
            const int blk = t & 63, L = t >> 6, bx = blk & 7, by = blk >> 3;
            const uint32_t mv = sh.pu_mv[block_pu(L, bx, by)];
            const int xm = mv_x(mv), ym = mv_y(mv);
            const int8_t* f = kFrac[(xm & 3) + ((ym & 3) << 2)];
            const int X = kMargin + (xm >> 2) - sax + bx * 8, Y = kMargin + (ym >> 2) - say + by * 8;
            uint8_t* pred = scr.pred[L < kNumLevels ? L : 0];
            if constexpr (Ext) { if (L >= kNumLevels) pred = job.scratch_ext[sb].pred[L - kNumLevels]; }
            uint32_t* dst = reinterpret_cast<uint32_t*>(pred) + (by * 8) * 16 + bx * 2;      // list 0: written; list 1: list 0's block, read
            uint32_t p0[16];
            if (list == 1) {
#pragma unroll
                for (int r = 0; r < 8; r++) { p0[2 * r] = dst[r * 16]; p0[2 * r + 1] = dst[r * 16 + 1]; }
            }
            uint32_t sad = 0;
#pragma unroll
            for (int r = 0; r < 8; r++) {
                uint2 c = load8(sh.plane[f[0]], sh.rowoff, Y + f[2] + r, X + f[1]);
                if (f[3] >= 0) {
                    const uint2 d = load8(sh.plane[f[3]], sh.rowoff, Y + f[5] + r, X + f[4]);
                    c.x = __vavgu4(c.x, d.x); c.y = __vavgu4(c.y, d.y);
                }
                if (list == 0) { dst[r * 16] = c.x; dst[r * 16 + 1] = c.y; }
                else {
                    const uint2 sv = *reinterpret_cast<const uint2*>(&sh.src[(by * 8 + r) * 16 + bx * 2]);
                    sad = sad4(sv.x, __vavgu4(p0[2 * r], c.x), sad);
                    sad = sad4(sv.y, __vavgu4(p0[2 * r + 1], c.y), sad);
                }
            }
            if (list == 1) part[t] = sad;
        }
        if (list == 1) {
            __syncthreads();
            if (tid < NPU) {
                // list 0 from the scratch the earlier kernels wrote, list 1 from this CTA's shared state (the same values it just stored)
                auto best_sad = [&](int l, int n) -> uint32_t {
                    if (l == 1) return sh.pu_sad[n];
                    if constexpr (Ext) { if (n >= kNumPu) return job.scratch_ext[sb].best_sad[0][n - kNumPu]; }
                    return scr.best_sad[0][n];
                };
                auto best_mv = [&](int l, int n) -> uint32_t {
                    if (l == 1) return sh.pu_mv[n];
                    if constexpr (Ext) { if (n >= kNumPu) return job.scratch_ext[sb].best_mv[0][n - kNumPu]; }
                    return scr.best_mv[0][n];
                };
                pack_pu_result<Ext>(job, sb, tid, part, best_sad, best_mv);
            }
        }
    }
    return true;
}

// Direct grid (one CTA per SB of the batch). The kGeomSmall instantiation appends the SBs it cannot hold to `worklist`
// (entry = job index << 16 | SB offset inside the job's range; worklist[-1] is the counter, zeroed by the host before).
template <bool Ext, int Geom>
__global__ void __launch_bounds__(kSpThreads, PlaneGeom<Geom>::kCtasPerSm) subpel_kernel(const DJob* __restrict__ jobs, int list, uint32_t* __restrict__ worklist) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    const DJob& job = jobs[blockIdx.y];
    const int sb = job.sb_begin + blockIdx.x;
    if (sb >= job.sb_end) return;
    const SvtMeParams& p = job.prm;
    if (list >= p.num_lists || (p.num_pus == kMaxPu) != Ext || wide_search(p) != (Geom == kGeomWide)) return;
    if (!subpel_sb<Ext, Geom>(job, sb, list, smem_raw) && threadIdx.x == 0)
        worklist[atomicAdd(worklist - 1, 1u)] = ((uint32_t)blockIdx.y << 16) | (uint32_t)blockIdx.x;
}

// Persistent CTAs over the work list the kGeomSmall pass left (normally a few SBs, often none).
template <bool Ext>
__global__ void __launch_bounds__(kSpThreads, PlaneGeom<kGeomFull>::kCtasPerSm) subpel_deferred_kernel(const DJob* __restrict__ jobs, int list, const uint32_t* __restrict__ worklist) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    const uint32_t n = worklist[-1];
    for (uint32_t i = blockIdx.x; i < n; i += gridDim.x) {
        const uint32_t item = worklist[i];
        const DJob& job = jobs[item >> 16];
        subpel_sb<Ext, kGeomFull>(job, job.sb_begin + (int)(item & 0xFFFFu), list, smem_raw);
        __syncthreads();                                   // the next SB reuses the shared planes
    }
}

template <bool Ext, int Geom> static constexpr size_t sp_smem() { return sizeof(SpSharedT<Ext ? kMaxPu : kNumPu, Geom>); }
template <bool Ext> static cudaError_t configure_ext() {
    cudaError_t e = cudaFuncSetAttribute(subpel_kernel<Ext, kGeomSmall>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sp_smem<Ext, kGeomSmall>());
    if (e == cudaSuccess) e = cudaFuncSetAttribute(subpel_kernel<Ext, kGeomWide>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sp_smem<Ext, kGeomWide>());
    if (e == cudaSuccess) e = cudaFuncSetAttribute(subpel_deferred_kernel<Ext>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sp_smem<Ext, kGeomFull>());
    return e;
}
cudaError_t configure_subpel() {
    const cudaError_t e = configure_ext<false>();
    return e != cudaSuccess ? e : configure_ext<true>();
}

// `worklist` points just behind a 32-bit counter; both are private to this launch sequence (stream) until it has run.
template <bool Ext> static void launch_ext(const DJob* d_jobs, int n_jobs, int max_sb, int list, bool wide, uint32_t* worklist, cudaStream_t st) {
    if (wide) { subpel_kernel<Ext, kGeomWide><<<dim3(max_sb, n_jobs), kSpThreads, sp_smem<Ext, kGeomWide>(), st>>>(d_jobs, list, nullptr); return; }
    cudaMemsetAsync(worklist - 1, 0, sizeof(uint32_t), st);
    subpel_kernel<Ext, kGeomSmall><<<dim3(max_sb, n_jobs), kSpThreads, sp_smem<Ext, kGeomSmall>(), st>>>(d_jobs, list, worklist);
    subpel_deferred_kernel<Ext><<<2 * 148, kSpThreads, sp_smem<Ext, kGeomFull>(), st>>>(d_jobs, list, worklist);
}
void launch_subpel(const DJob* d_jobs, int n_jobs, int max_sb, int list, bool ext, bool wide, uint32_t* worklist, cudaStream_t st) {
    if (ext) launch_ext<true>(d_jobs, n_jobs, max_sb, list, wide, worklist, st);
    else     launch_ext<false>(d_jobs, n_jobs, max_sb, list, wide, worklist, st);
}

}  // namespace svtme
