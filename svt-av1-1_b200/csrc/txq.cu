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
// are quantised by all threads of the block with coalesced 128-byte stores; eob is a shared-memory atomicMax over
// (iscan + 1) of the non-zero dequantised coefficients. Several blocks share a CTA so that every CTA has 256
// threads even for 4x4.
#include "me_common.cuh"
#include "txfm1d_gen.cuh"
#include "txq.cuh"

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
template <int TXW, int TXH> struct TxqGeom {
    static constexpr int B = 256 / cmin(TXW, TXH);
    static constexpr int PITCH = TXW + 1;          // odd: row r of slot s starts at word PITCH * (s * TXH + r)
    static constexpr int SLOT = TXH * PITCH;
    static constexpr int PW = cmin(TXW, 32), PH = cmin(TXH, 32), N = PW * PH;
    static constexpr int SMEM = B * SLOT * (int)sizeof(int32_t);
    static constexpr bool BIG = TXW == 64 || TXH == 64;
};

struct QuantLane { int32_t zbin, round, quant_hi, qshift, dequant; };    // one lane (DC or AC) of the quantiser

template <int LS> __device__ __forceinline__ QuantLane quant_lane(const SvtTxqQuant& q, int ac) {
    constexpr int RND = LS ? (1 << (LS - 1)) : 0;
    QuantLane l;
    l.zbin = (int32_t)(int16_t)((uint16_t)((ac ? q.zbin[1] : q.zbin[0]) + RND) >> LS);
    l.round = (int32_t)(int16_t)((uint16_t)((ac ? q.round[1] : q.round[0]) + RND) >> LS);
    l.quant_hi = (int32_t)(ac ? q.quant[1] : q.quant[0]) << 16;           // (t * quant) >> 16 == mulhi(t, quant << 16)
    l.qshift = ac ? q.quant_shift[1] : q.quant_shift[0];
    l.dequant = ac ? q.dequant[1] : q.dequant[0];
    return l;
}

template <int TXW, int TXH>
__global__ void __launch_bounds__(256, TxqGeom<TXW, TXH>::BIG ? 2 : 1) txq_kernel(const TxqArgs a) {
    using G = TxqGeom<TXW, TXH>;
    constexpr int B = G::B, PITCH = G::PITCH, SLOT = G::SLOT, PW = G::PW, PH = G::PH, N = G::N;
    constexpr int CC = kCosCol[ilog2(TXW) - 2][ilog2(TXH) - 2], CR = kCosRow[ilog2(TXW) - 2][ilog2(TXH) - 2];
    constexpr bool RECT = (ilog2(TXW) - ilog2(TXH) == 1) || (ilog2(TXH) - ilog2(TXW) == 1);
    extern __shared__ __align__(16) int32_t smem_txq[];
    __shared__ unsigned int s_eob[B];
    __shared__ unsigned long long s_energy[B];
    const int tid = threadIdx.x;
    const int blk0 = blockIdx.x * B;
    const int n_here = min(B, a.n_blocks - blk0);            // blocks of this CTA that exist
    if (tid < B) { s_eob[tid] = 0; s_energy[tid] = 0; }

    // ---- vertical pass: one item = one column of one block (Av1TranformTwoDCore_c, "Columns")
#pragma unroll 1
    for (int i = tid; i < B * TXW; i += 256) {
        const int slot = i / TXW, c = i % TXW;
        if (slot >= n_here) break;
        const int blk = blk0 + slot, by = blk / a.blocks_x, bx = blk - by * a.blocks_x;
        const int16_t* in = a.residual + (size_t)by * TXH * a.stride + (size_t)bx * TXW + c;
        int32_t x[TXH];
        const bool ud = a.vk == K_FLIPADST;
#pragma unroll
        for (int r = 0; r < TXH; r++) x[r] = (int32_t)in[(size_t)(ud ? TXH - 1 - r : r) * a.stride] * (1 << shift0(TXW, TXH));
        txfm1d<TXH, CC>(ud ? K_ADST : a.vk, x);
        int32_t* col = smem_txq + slot * SLOT + (a.hk == K_FLIPADST ? TXW - 1 - c : c);
#pragma unroll
        for (int r = 0; r < TXH; r++) col[r * PITCH] = shift1(TXW, TXH) ? rshift_round(x[r], shift1(TXW, TXH)) : x[r];
    }
    __syncthreads();
    // ---- horizontal pass: one item = one row of one block ("Rows"); the kept PW columns go back in place
#pragma unroll 1
    for (int j = tid; j < B * TXH; j += 256) {
        const int slot = j / TXH, r = j % TXH;
        if (slot >= n_here) break;
        int32_t* row = smem_txq + j * PITCH;                  // == slot * SLOT + r * PITCH
        int32_t y[TXW];
#pragma unroll
        for (int c = 0; c < TXW; c++) y[c] = row[c];
        txfm1d<TXW, CR>(a.hk == K_FLIPADST ? K_ADST : a.hk, y);
#pragma unroll
        for (int c = 0; c < TXW; c++) {
            if (shift2(TXW, TXH)) y[c] = rshift_round(y[c], shift2(TXW, TXH));
            if (RECT) y[c] = mul_sqrt2(y[c]);
        }
        if (G::BIG) {                                         // HandleTransform64x64_c & co: energy of what is dropped
            unsigned long long dropped = 0;
#pragma unroll
            for (int c = 0; c < TXW; c++)
                if (c >= PW || r >= PH) dropped += (unsigned long long)((long long)y[c] * (long long)y[c]);
            if (dropped) atomicAdd(&s_energy[slot], dropped);
        }
        if (r < PH) {
#pragma unroll
            for (int c = 0; c < PW; c++) row[c] = y[c];
        }
    }
    __syncthreads();

    // ---- quantisation (aom_highbd_quantize_b*_avx2): the CTA's B * N coefficients are one contiguous run of each
    // output array; thread tid takes elements tid, tid + 256, ... so every warp store is one full 128-byte line.
    constexpr int LS = (TXW * TXH > 256) + (TXW * TXH > 1024);               // av1_get_tx_scale
    constexpr int SUB = N >= 256 ? N / 256 : 1;                               // iterations per block
    constexpr int SPI = N >= 256 ? 1 : 256 / N;                               // blocks per iteration
    constexpr int ITER = B * N / 256;
    const int slot0 = N >= 256 ? 0 : tid / N, rc0 = N >= 256 ? tid : tid % N;
    const int off0 = slot0 * SLOT + (rc0 / PW) * PITCH + rc0 % PW;
    const QuantLane ac = quant_lane<LS>(a.q, 1);
    QuantLane first = ac;                                                     // lane of this thread's rc0 (DC for rc0 == 0)
    if (rc0 == 0) first = quant_lane<LS>(a.q, 0);
    int32_t isc[SUB];
#pragma unroll
    for (int u = 0; u < SUB; u++) isc[u] = (int32_t)__ldg(&a.iscan[rc0 + 256 * u]) + 1;
    const size_t out0 = (size_t)blk0 * N + tid;
    const int n_out = n_here * N;
    unsigned int m = 0;
#pragma unroll
    for (int k = 0; k < ITER; k++) {
        const int u = k % SUB;
        const int slot = slot0 + (N >= 256 ? k / SUB : k * SPI);
        const int off = off0 + (N >= 256 ? (k / SUB) * SLOT + u * (256 / PW) * PITCH : k * SPI * SLOT);
        const bool live = tid + 256 * k < n_out;
        if (u == 0) m = 0;
        if (live) {
            const QuantLane& L = u == 0 ? first : ac;
            const int32_t v = smem_txq[off];
            const int32_t av = v < 0 ? -v : v;
            int32_t sgn = (v >> 31) | 1;
            sgn = av >= L.zbin ? sgn : 0;
            const int32_t t0 = av + L.round;
            const int32_t t1 = __mulhi(t0, L.quant_hi) + t0;
            const int32_t q = (int32_t)(((int64_t)t1 * L.qshift) >> (16 - LS));
            const int32_t dq = (int32_t)((uint32_t)(q * L.dequant) >> LS);
            const size_t o = out0 + 256 * k;
            if (a.coeff) a.coeff[o] = v;
            if (a.qcoeff) a.qcoeff[o] = q * sgn;
            if (a.dqcoeff) a.dqcoeff[o] = dq * sgn;
            if (dq * sgn != 0) m = max(m, (unsigned int)isc[u]);
        }
        if (u == SUB - 1) {                                   // all of this thread's coefficients of block `slot` are done
            if (N >= 32) {
                m = __reduce_max_sync(0xffffffffu, m);
                if ((tid & 31) == 0 && m) atomicMax(&s_eob[slot], m);
            } else {
                if (m) atomicMax(&s_eob[slot], m);
            }
        }
    }
    __syncthreads();
    if (tid < n_here) {
        if (a.eob) a.eob[blk0 + tid] = (uint16_t)s_eob[tid];
        if (a.energy) a.energy[blk0 + tid] = s_energy[tid];
    }
}

template <int TXW, int TXH> static cudaError_t launch_one(const TxqArgs& a, cudaStream_t st) {
    constexpr int B = TxqGeom<TXW, TXH>::B;
    const int smem = TxqGeom<TXW, TXH>::SMEM;
    static bool configured = false;
    if (!configured) {
        cudaError_t e = cudaFuncSetAttribute(txq_kernel<TXW, TXH>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) return e;
        configured = true;
    }
    txq_kernel<TXW, TXH><<<(a.n_blocks + B - 1) / B, 256, smem, st>>>(a);
    return cudaGetLastError();
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
