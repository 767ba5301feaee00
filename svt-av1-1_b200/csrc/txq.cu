// txq.cu -- AV1 forward 2-D transform + quantisation for a batch of equally sized transform blocks.
// Reference: Av1EstimateTransform (Source/Lib/Codec/EbTransforms.c:4410-4728) -> Av1TranformTwoDCore_c (:3701-3775)
// with the 1-D kernels av1_fdct*_new / av1_fadst*_new / av1_fidentity*_c (:1314-3660), the 64-point zero-out /
// repack / three_quad_energy (:3895-3926, 4072-4224), and aom_highbd_quantize_b{,_32x32,_64x64}_avx2
// (Source/Lib/ASM_AVX2/highbd_quantize_intrin_avx2.c:82-330) behind aom_quantize_b*.
//
// B200 mapping: integer butterflies in registers, no tensor cores (not a contraction), HBM-bound (2 B in,
// 12 B + 2 B out per pixel). One thread owns one column for the vertical pass and one row for the horizontal
// pass (straight-line generated code, txfm1d_gen.cuh, cospi folded at compile time); the transpose goes through
// shared memory with an odd pitch (conflict-free both ways). The packed coefficients return to shared memory and
// are quantised by all threads of the block with coalesced 128-byte stores; eob comes from a warp ballot (blocks of up to 32
// coefficients, visited in scan order) or a per-lane running maximum of (iscan + 1) over the non-zero dequantised
// coefficients reduced once per block (REDUX). Several blocks share a CTA so that every CTA has 256 threads even for 4x4.
#include "me_common.cuh"
#include "txfm1d_gen.cuh"
#include "txq.cuh"
#include <mutex>

namespace svtme {

__device__ __forceinline__ int32_t rshift_round(int32_t v, int bit) { return (v + (1 << (bit - 1))) >> bit; }
__device__ __forceinline__ int32_t rshift_round64(int64_t v, int bit) { return (int32_t)((v + ((int64_t)1 << (bit - 1))) >> bit); }
__device__ __forceinline__ int32_t mul_sqrt2(int32_t v, int mult = 1) { return rshift_round64((int64_t)v * (mult * 5793), 12); }

// av1_fadst4_new (EbTransforms.c:2764-2855)
template <int BIT> __device__ __forceinline__ void tx_fadst4(int32_t (&x)[4]) {
    constexpr int32_t S[5] = { 0, kTxSinpi[BIT - 10][1], kTxSinpi[BIT - 10][2], kTxSinpi[BIT - 10][3], kTxSinpi[BIT - 10][4] };
    const int32_t x0 = x[0], x1 = x[1], x2 = x[2], x3 = x[3];
    if (!(x0 | x1 | x2 | x3)) return;
    const int32_t s0 = S[1] * x0, s1 = S[4] * x0, s2 = S[2] * x1, s3 = S[1] * x1, s4 = S[3] * x2, s5 = S[4] * x3, s6 = S[2] * x3;
    const int32_t s7 = x0 + x1 - x3;
    const int32_t a0 = s0 + s2 + s5, a1 = S[3] * s7, a2 = s1 - s3 + s6, a3 = s4;
    x[0] = rshift_round64(a0 + a3, BIT); x[1] = rshift_round64(a1, BIT);
    x[2] = rshift_round64(a2 - a3, BIT); x[3] = rshift_round64(a2 - a0 + a3, BIT);
}

// one N-point forward transform in registers; kind is uniform across the CTA
template <int N, int BIT> __device__ __forceinline__ void txfm1d(int kind, int32_t (&x)[N]) {
    if (kind == K_IDTX) {                                        // av1_fidentity*_c
#pragma unroll
        for (int i = 0; i < N; i++)
            x[i] = N == 4 ? mul_sqrt2(x[i]) : N == 8 ? x[i] * 2 : N == 16 ? mul_sqrt2(x[i], 2) : N == 32 ? x[i] * 4 : mul_sqrt2(x[i], 4);
        return;
    }
    if (kind == K_DCT) {
        if constexpr (N == 4) tx_fdct4<BIT>(x);
        else if constexpr (N == 8) tx_fdct8<BIT>(x);
        else if constexpr (N == 16) tx_fdct16<BIT>(x);
        else if constexpr (N == 32) tx_fdct32<BIT>(x);
        else tx_fdct64<BIT>(x);
        return;
    }
    if constexpr (N == 4) tx_fadst4<BIT>(x);
    else if constexpr (N == 8) tx_fadst8<BIT>(x);
    else if constexpr (N == 16) tx_fadst16<BIT>(x);
}

__host__ __device__ constexpr int ilog2(int v) { return v <= 1 ? 0 : 1 + ilog2(v >> 1); }
// fwd_cos_bit_col / fwd_cos_bit_row and fwd_txfm_shift_ls, EbTransforms.h:119-151
constexpr int kCosCol[5][5] = { { 13, 13, 13, 0, 0 }, { 13, 13, 13, 12, 0 }, { 13, 13, 13, 12, 13 }, { 0, 13, 13, 12, 13 }, { 0, 0, 13, 12, 13 } };
constexpr int kCosRow[5][5] = { { 13, 13, 12, 0, 0 }, { 13, 13, 13, 12, 0 }, { 13, 13, 12, 13, 12 }, { 0, 12, 13, 12, 11 }, { 0, 0, 12, 11, 10 } };
__host__ __device__ constexpr int cmax(int a, int b) { return a > b ? a : b; }
__host__ __device__ constexpr int cmin(int a, int b) { return a < b ? a : b; }
__host__ __device__ constexpr int shift0(int w, int h) { return ((w == 64 && h == 64) || (w == 32 && h == 64) || (w == 16 && h == 64)) ? 0 : 2; }
__host__ __device__ constexpr int shift1(int w, int h) {
    return (w == 64 && h == 64) ? 2 : (w == 32 && h == 64) ? 2 : (w == 64 && h == 32) ? 4 : (w == 16 && h == 64) ? 2 : (w == 64 && h == 16) ? 4 :
           (cmax(w, h) == 32 && cmin(w, h) == 8) ? 2 : cmax(w, h) == 32 ? 4 : (cmax(w, h) == 16 && cmin(w, h) == 4) ? 1 :
           cmax(w, h) == 16 ? 2 : cmax(w, h) == 8 ? 1 : 0;
}
__host__ __device__ constexpr int shift2(int w, int h) { return ((w == 64 && h == 64) || (w == 32 && h == 64) || (w == 64 && h == 32)) ? 2 : 0; }

// Geometry of one CTA: B = 256 / min(TXW, TXH) transform blocks, consecutive in raster order. The column pass has
// B * TXW work items (one TXH-point transform each), the row pass B * TXH items (one TXW-point transform each); the
// 256 threads loop over them, so rectangular sizes keep every thread busy in both passes.
// ENERGY = false (64-point sizes whose caller does not want three_quad_energy, e.g. the EncDec call sites): the rows and
// columns that are only squared into the energy are never produced -- the column pass keeps ROWS = 32 rows, the row
// pass reads 32 outputs, and the compiler prunes the straight-line 64-point graphs down to what feeds them.
template <int TXW, int TXH, bool ENERGY> struct TxqGeom {
    static constexpr int PW = cmin(TXW, 32), PH = cmin(TXH, 32), N = PW * PH;
    static constexpr int ROWS = ENERGY ? TXH : PH;  // rows of the column-pass output that are kept
    static constexpr int B = 256 / cmin(TXW, ROWS);
    static constexpr int PITCH = TXW + 1;          // odd: row r of slot s starts at word PITCH * (s * ROWS + r)
    static constexpr int SLOT = ROWS * PITCH;
    static constexpr int SMEM = B * SLOT * (int)sizeof(int32_t);
    static constexpr bool BIG = TXW == 64 || TXH == 64;
    // resident CTAs per SM the register allocation is held to (256 threads each)
    static constexpr int MIN_CTAS = BIG ? 2 : TXW * TXH <= 32 ? 8 : TXW * TXH <= 128 ? 6 : TXW * TXH <= 256 ? 4 : 3;
};

__device__ __forceinline__ int32_t mul_shift(int32_t x, int32_t y, int sh) {   // (int64(x) * y) >> sh, low 32 bits
    int64_t p;
    asm("mul.wide.s32 %0, %1, %2;" : "=l"(p) : "r"(x), "r"(y));
    return (int32_t)(p >> sh);
}
__device__ __forceinline__ int32_t top_bit_plus1(unsigned int bits) {            // 0 for 0, else 1 + index of the highest set bit
    int32_t r;
    asm("bfind.u32 %0, %1;" : "=r"(r) : "r"(bits));
    return r + 1;
}

// Quantisation of the CTA's packed coefficients (aom_highbd_quantize_b*_avx2). The B * N coefficients of a CTA are one
// contiguous run of each output array; warp w owns the w-th eighth of it and walks it 32 coefficients at a time, so
// every warp store is one full 128-byte line and (for B >= 8) a transform block never leaves its warp: eob is a
// per-lane running maximum of (iscan + 1) over non-zero dequantised values, reduced once per block.
template <int TXW, int TXH, bool ENERGY, bool FULL>
__device__ __forceinline__ void quant_phase(const TxqArgs& a, const int32_t* smem, unsigned int* s_eob, int blk0, int n_here) {
    using G = TxqGeom<TXW, TXH, ENERGY>;
    constexpr int B = G::B, PITCH = G::PITCH, SLOT = G::SLOT, PW = G::PW, N = G::N;
    constexpr int LS = (TXW * TXH > 256) + (TXW * TXH > 1024);               // av1_get_tx_scale
    constexpr int WARP_COEFS = B * N / 8, ITER = WARP_COEFS / 32;
    constexpr bool OWNED = WARP_COEFS >= N;                                   // whole blocks per warp
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int slot_base = (warp * WARP_COEFS) / N, rc_base = (warp * WARP_COEFS) % N;
    // blocks of up to 32 coefficients: lane l takes the l-th coefficient in SCAN order (the store still fills one
    // line), so eob falls out of a ballot; larger blocks: raster order and a running maximum of iscan + 1
    constexpr bool BALLOT = N <= 32;
    const int rc_l = BALLOT ? (int)__ldg(a.scan + lane % N) : lane;
    const int soff0 = (slot_base + (N >= 32 ? 0 : lane / N)) * SLOT + (rc_base / PW + rc_l / PW) * PITCH + rc_l % PW;
    const int16_t* isc_ptr = a.iscan + rc_base + lane;
    const TxqLane ac = a.ac;
    TxqLane first = ac;                                                       // lane used where the block's scan restarts
    if (rc_base == 0 && rc_l == 0) first = a.dc;
    const size_t out0 = (size_t)blk0 * N + warp * WARP_COEFS + (BALLOT ? (lane / N) * N + rc_l : lane);
    const int n_out = n_here * N - warp * WARP_COEFS - (BALLOT ? (lane / N) * N : lane);   // coefficients from this lane's first one on
    // outer loop: one block (or 32 / N blocks of a 4x4 size, or this warp's share of a 64x64 block) per trip
    constexpr int SUBS = N < 32 ? 1 : OWNED ? N / 32 : WARP_COEFS / 32, OUTER = ITER / SUBS;
    constexpr int SLOTS_PER_TRIP = N >= 32 ? 1 : 32 / N;
    constexpr int UNROLL_OUTER = SUBS >= 8 ? 1 : (8 / SUBS < OUTER ? 8 / SUBS : OUTER);
#pragma unroll(UNROLL_OUTER)
    for (int s = 0; s < OUTER; s++) {
        const int32_t* src = smem + soff0 + s * SLOTS_PER_TRIP * SLOT;
        const size_t o_s = out0 + (size_t)s * 32 * SUBS;
        unsigned int m = 0;
#pragma unroll(SUBS < 8 ? SUBS : 8)
        for (int u = 0; u < SUBS; u++) {
            if (FULL || s * 32 * SUBS + 32 * u < n_out) {
                const TxqLane& L = u == 0 ? first : ac;
                const int32_t v = src[((32 * u) / PW) * PITCH];
                const int32_t isc = BALLOT ? 1 : (int32_t)__ldg(isc_ptr + 32 * u) + 1;
                const int32_t av = v < 0 ? -v : v;
                int32_t sgn = (v >> 31) | 1;
                sgn = av >= L.zbin ? sgn : 0;
                const int32_t t0 = av + L.round;
                const int32_t t1 = __mulhi(t0, L.quant_hi) + t0;
                const int32_t q = mul_shift(t1, L.qshift, 16 - LS);
                const int32_t dq = (int32_t)((uint32_t)(q * L.dequant) >> LS);
                const int32_t dqs = dq * sgn;
                const size_t o = o_s + 32 * u;
                if (FULL || a.coeff) a.coeff[o] = v;
                if (FULL || a.qcoeff) a.qcoeff[o] = q * sgn;
                if (FULL || a.dqcoeff) a.dqcoeff[o] = dqs;
                m = max(m, dqs != 0 ? (unsigned int)isc : 0u);
            }
        }
        const int slot = slot_base + s * SLOTS_PER_TRIP;
        if (BALLOT) {
            const unsigned int nz = __ballot_sync(0xffffffffu, m != 0);
            if (lane < SLOTS_PER_TRIP) {
                const unsigned int bits = N == 32 ? nz : (nz >> (lane * (N & 31))) & ((1u << (N & 31)) - 1u);
                s_eob[slot + lane] = top_bit_plus1(bits);
            }
        } else {
            m = __reduce_max_sync(0xffffffffu, m);
            if (lane == 0) { if (OWNED) s_eob[slot] = m; else if (m) atomicMax(&s_eob[slot], m); }
        }
    }
}

template <int TXW, int TXH, bool ENERGY>
__global__ void __launch_bounds__(256, TxqGeom<TXW, TXH, ENERGY>::MIN_CTAS) txq_kernel(const TxqArgs a) {
    using G = TxqGeom<TXW, TXH, ENERGY>;
    constexpr int B = G::B, PITCH = G::PITCH, SLOT = G::SLOT, PW = G::PW, PH = G::PH, ROWS = G::ROWS;
    constexpr int CC = kCosCol[ilog2(TXW) - 2][ilog2(TXH) - 2], CR = kCosRow[ilog2(TXW) - 2][ilog2(TXH) - 2];
    constexpr bool RECT = (ilog2(TXW) - ilog2(TXH) == 1) || (ilog2(TXH) - ilog2(TXW) == 1);
    extern __shared__ __align__(16) int32_t smem_txq[];
    __shared__ unsigned int s_eob[B];
    __shared__ unsigned long long s_energy[B];
    const int tid = threadIdx.x;
    const int blk0 = blockIdx.x * B;
    const int n_here = min(B, a.n_blocks - blk0);            // blocks of this CTA that exist
    if (tid < B) { s_eob[tid] = 0; s_energy[tid] = 0; }

    const int by0 = blk0 / a.blocks_x, bx0 = blk0 - by0 * a.blocks_x;
    // ---- vertical pass: one item = one column of one block (Av1TranformTwoDCore_c, "Columns")
#pragma unroll
    for (int i = tid; i < B * TXW; i += 256) {
        const int slot = i / TXW, c = i % TXW;
        if (slot >= n_here) continue;
        int bx = bx0 + slot, by = by0;                        // block (blk0 + slot) of the raster
        while (bx >= a.blocks_x) { bx -= a.blocks_x; by++; }
        const int16_t* in = a.residual + (size_t)by * TXH * a.stride + (size_t)bx * TXW + c;
        int32_t x[TXH];
        const bool ud = a.vk == K_FLIPADST;
#pragma unroll
        for (int r = 0; r < TXH; r++) x[r] = (int32_t)__ldg(in + (size_t)(ud ? TXH - 1 - r : r) * a.stride) * (1 << shift0(TXW, TXH));
        txfm1d<TXH, CC>(ud ? K_ADST : a.vk, x);
        int32_t* col = smem_txq + slot * SLOT + (a.hk == K_FLIPADST ? TXW - 1 - c : c);
#pragma unroll
        for (int r = 0; r < ROWS; r++) col[r * PITCH] = shift1(TXW, TXH) ? rshift_round(x[r], shift1(TXW, TXH)) : x[r];
    }
    __syncthreads();
    // ---- horizontal pass: one item = one row of one block ("Rows"); the kept PW columns go back in place
#pragma unroll
    for (int j = tid; j < B * ROWS; j += 256) {
        const int slot = j / ROWS, r = j % ROWS;
        unsigned long long dropped = 0;
        if (slot < n_here) {
            int32_t* row = smem_txq + j * PITCH;                  // == slot * SLOT + r * PITCH
            int32_t y[TXW];
#pragma unroll
            for (int c = 0; c < TXW; c++) y[c] = row[c];
            txfm1d<TXW, CR>(a.hk == K_FLIPADST ? K_ADST : a.hk, y);
#pragma unroll
            for (int c = 0; c < (ENERGY ? TXW : PW); c++) {
                if (shift2(TXW, TXH)) y[c] = rshift_round(y[c], shift2(TXW, TXH));
                if (RECT) y[c] = mul_sqrt2(y[c]);
            }
            if (G::BIG && ENERGY) {                                     // HandleTransform64x64_c & co: energy of what is dropped
#pragma unroll
                for (int c = 0; c < TXW; c++)
                    if (c >= PW || r >= PH) dropped += (unsigned long long)((long long)y[c] * (long long)y[c]);
            }
            if (r < PH) {
#pragma unroll
                for (int c = 0; c < PW; c++) row[c] = y[c];
            }
        }
        if (G::BIG && ENERGY) {
            // the rows of a block sit on consecutive lanes (B * ROWS is a multiple of 256, so every lane of the warp is here):
            // one shared 64-bit atomic per warp and block instead of one per row (64 rows contending for the same word)
            constexpr int RW = ROWS < 32 ? ROWS : 32;
#pragma unroll
            for (int o = RW / 2; o > 0; o >>= 1) dropped += __shfl_xor_sync(0xffffffffu, dropped, o);
            if ((tid & (RW - 1)) == 0 && dropped) atomicAdd(&s_energy[slot], dropped);
        }
    }
    __syncthreads();

    // ---- quantisation: fast path when every block of the CTA exists and all three coefficient arrays are wanted
    if (n_here == B && a.coeff && a.qcoeff && a.dqcoeff) quant_phase<TXW, TXH, ENERGY, true>(a, smem_txq, s_eob, blk0, n_here);
    else quant_phase<TXW, TXH, ENERGY, false>(a, smem_txq, s_eob, blk0, n_here);
    __syncthreads();
    if (tid < n_here) {
        if (a.eob) a.eob[blk0 + tid] = (uint16_t)s_eob[tid];
        if (a.energy) a.energy[blk0 + tid] = s_energy[tid];
    }
}

template <int TXW, int TXH, bool ENERGY> static cudaError_t launch_variant(const TxqArgs& a, cudaStream_t st) {
    constexpr int B = TxqGeom<TXW, TXH, ENERGY>::B;
    const int smem = TxqGeom<TXW, TXH, ENERGY>::SMEM;
    // the attribute is per device and launches may come from any thread: one bit per device ordinal under a mutex
    static std::mutex mu;
    static uint64_t configured = 0;
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    {
        std::lock_guard<std::mutex> lock(mu);
        if (dev < 0 || dev > 63 || !((configured >> dev) & 1u)) {
            e = cudaFuncSetAttribute(txq_kernel<TXW, TXH, ENERGY>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
            if (e != cudaSuccess) return e;
            if (dev >= 0 && dev <= 63) configured |= (uint64_t)1 << dev;
        }
    }
    txq_kernel<TXW, TXH, ENERGY><<<(a.n_blocks + B - 1) / B, 256, smem, st>>>(a);
    return cudaGetLastError();
}
template <int TXW, int TXH> static cudaError_t launch_one(const TxqArgs& a, cudaStream_t st) {
    if (TXW == 64 || TXH == 64) {                             // only the 64-point sizes drop coefficients
        if (a.energy) return launch_variant<TXW, TXH, true>(a, st);
        return launch_variant<TXW, TXH, (TXW != 64 && TXH != 64)>(a, st);
    }
    return launch_variant<TXW, TXH, true>(a, st);
}

cudaError_t launch_txq(int txw, int txh, const TxqArgs& a, cudaStream_t st) {
#define CASE(W, H) if (txw == W && txh == H) return launch_one<W, H>(a, st);
    CASE(4, 4) CASE(8, 8) CASE(16, 16) CASE(32, 32) CASE(64, 64)
    CASE(4, 8) CASE(8, 4) CASE(8, 16) CASE(16, 8) CASE(16, 32) CASE(32, 16) CASE(32, 64) CASE(64, 32)
    CASE(4, 16) CASE(16, 4) CASE(8, 32) CASE(32, 8) CASE(16, 64) CASE(64, 16)
#undef CASE
    return cudaErrorInvalidValue;
}

}  // namespace svtme
