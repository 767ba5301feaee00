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

template <int TXW, int TXH>
__global__ void __launch_bounds__(256) txq_kernel(const TxqArgs a) {
    constexpr int T = cmax(TXW, TXH);              // threads per transform block
    constexpr int B = 256 / T;                     // transform blocks per CTA
    constexpr int PITCH = TXW + 1;                 // odd pitch: column writes and row reads are conflict-free
    constexpr int PW = cmin(TXW, 32), PH = cmin(TXH, 32);
    constexpr int CC = kCosCol[ilog2(TXW) - 2][ilog2(TXH) - 2], CR = kCosRow[ilog2(TXW) - 2][ilog2(TXH) - 2];
    constexpr bool RECT = (ilog2(TXW) - ilog2(TXH) == 1) || (ilog2(TXH) - ilog2(TXW) == 1);
    extern __shared__ __align__(16) int32_t smem_txq[];
    __shared__ unsigned int s_eob[B];
    __shared__ unsigned long long s_energy[B];
    const int slot = threadIdx.x / T, t = threadIdx.x % T;
    const int blk = blockIdx.x * B + slot;
    int32_t* buf = smem_txq + slot * (TXH * PITCH);
    if (t == 0) { s_eob[slot] = 0; s_energy[slot] = 0; }
    const bool on = blk < a.n_blocks;
    const int bx = on ? blk % a.blocks_x : 0, by = on ? blk / a.blocks_x : 0;
    const int16_t* in = a.residual + (size_t)by * TXH * a.stride + (size_t)bx * TXW;

    // ---- vertical pass: thread t owns column t (Av1TranformTwoDCore_c, "Columns")
    if (on && t < TXW) {
        int32_t x[TXH];
        const bool ud = a.vk == K_FLIPADST;
#pragma unroll
        for (int r = 0; r < TXH; r++) x[r] = (int32_t)in[(size_t)(ud ? TXH - 1 - r : r) * a.stride + t] * (1 << shift0(TXW, TXH));
        txfm1d<TXH, CC>(a.vk == K_FLIPADST ? K_ADST : a.vk, x);
        const int c = a.hk == K_FLIPADST ? TXW - 1 - t : t;
#pragma unroll
        for (int r = 0; r < TXH; r++) buf[r * PITCH + c] = shift1(TXW, TXH) ? rshift_round(x[r], shift1(TXW, TXH)) : x[r];
    }
    __syncthreads();
    // ---- horizontal pass: thread t owns row t ("Rows"), then the packed block goes back to shared memory
    int32_t y[TXW];
    unsigned long long dropped = 0;
    if (on && t < TXH) {
#pragma unroll
        for (int c = 0; c < TXW; c++) y[c] = buf[t * PITCH + c];
        txfm1d<TXW, CR>(a.hk == K_FLIPADST ? K_ADST : a.hk, y);
#pragma unroll
        for (int c = 0; c < TXW; c++) {
            if (shift2(TXW, TXH)) y[c] = rshift_round(y[c], shift2(TXW, TXH));
            if (RECT) y[c] = mul_sqrt2(y[c]);
        }
        if (TXW > 32 || TXH > 32) {                // HandleTransform64x64_c & co: energy of what is dropped
#pragma unroll
            for (int c = 0; c < TXW; c++)
                if (c >= PW || t >= PH) dropped += (unsigned long long)((long long)y[c] * (long long)y[c]);
        }
    }
    __syncthreads();
    if (on && t < PH) {
#pragma unroll
        for (int c = 0; c < PW; c++) buf[t * PITCH + c] = y[c];
    }
    if ((TXW > 32 || TXH > 32) && dropped) atomicAdd(&s_energy[slot], dropped);
    __syncthreads();

    // ---- quantisation, all T threads of the block, coalesced (aom_highbd_quantize_b*_avx2)
    if (on) {
        constexpr int LS = (TXW * TXH > 256) + (TXW * TXH > 1024);           // av1_get_tx_scale
        constexpr int RND = LS ? (1 << (LS - 1)) : 0;
        const size_t base = (size_t)blk * (PW * PH);
        unsigned int my_eob = 0;
        for (int rc = t; rc < PW * PH; rc += T) {
            const int r = rc / PW, c = rc - r * PW;
            const int32_t v = buf[r * PITCH + c];
            if (a.coeff) a.coeff[base + rc] = v;
            const int ac = rc != 0;
            const int32_t zbin = (int32_t)(int16_t)((uint16_t)(a.q.zbin[ac] + RND) >> LS);
            const int32_t round = (int32_t)(int16_t)((uint16_t)(a.q.round[ac] + RND) >> LS);
            const int32_t av = v < 0 ? -v : v;
            int32_t q = 0, dq = 0;
            if (av >= zbin) {
                int32_t tq = av + round;
                tq += (int32_t)(((int64_t)tq * a.q.quant[ac]) >> 16);
                q = (int32_t)(((int64_t)tq * a.q.quant_shift[ac]) >> (16 - LS));
                dq = (int32_t)((uint32_t)(q * (int32_t)a.q.dequant[ac]) >> LS);
                if (v < 0) { q = -q; dq = -dq; }
                if (dq != 0) my_eob = max(my_eob, (unsigned int)a.iscan[rc] + 1u);
            }
            if (a.qcoeff) a.qcoeff[base + rc] = q;
            if (a.dqcoeff) a.dqcoeff[base + rc] = dq;
        }
        if (my_eob) atomicMax(&s_eob[slot], my_eob);
    }
    __syncthreads();
    if (on && t == 0) {
        if (a.eob) a.eob[blk] = (uint16_t)s_eob[slot];
        if (a.energy) a.energy[blk] = s_energy[slot];
    }
}

template <int TXW, int TXH> static cudaError_t launch_one(const TxqArgs& a, cudaStream_t st) {
    constexpr int T = cmax(TXW, TXH), B = 256 / T;
    const int smem = B * TXH * (TXW + 1) * (int)sizeof(int32_t);
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
