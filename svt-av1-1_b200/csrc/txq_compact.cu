// txq_compact.cu -- eob-bounded return path of the forward transform + quantisation family (include/svttxq.h,
// svttxq_fwd_txfm_quant_compact). After quantisation the consumer of a transform block needs its eob and the `eob`
// leading quantised levels in scan order (the entropy coder walks the scan backwards from eob - 1, and
// dqcoeff = (level * dequant) >> log_scale can be rebuilt where it is needed: EbFullLoop.c:596-662,
// ASM_AVX2/highbd_quantize_intrin_avx2.c:82-200), not three int32 planes. These kernels turn the dense qcoeff plane the
// transform kernel left in HBM into
//     offsets[b] = sum of eob[0 .. b)              (exclusive prefix sum over the blocks of the call, deterministic)
//     levels[offsets[b] + i] = (int16) qcoeff[b][scan[i]],  i < eob[b]
// so that 2 bytes per block + 2 bytes per kept level cross the host link instead of 12 bytes per coefficient.
#include "me_common.cuh"
#include "txq.cuh"

namespace svtme {

constexpr int kScanThreads = 256, kScanPer = 8, kScanTile = kScanThreads * kScanPer;    // blocks per CTA of the prefix sum

__device__ __forceinline__ uint32_t cta_exclusive_scan(uint32_t v, uint32_t* warp_sums, uint32_t& total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += t; }
    if (lane == 31) warp_sums[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        uint32_t w = lane < kScanThreads / 32 ? warp_sums[lane] : 0u, winc = w;
#pragma unroll
        for (int o = 1; o < 8; o <<= 1) { const uint32_t t = __shfl_up_sync(0xffffffffu, winc, o); if (lane >= o) winc += t; }
        if (lane < kScanThreads / 32) warp_sums[lane] = winc - w;
        if (lane == kScanThreads / 32 - 1) warp_sums[kScanThreads / 32] = winc;
    }
    __syncthreads();
    total = warp_sums[kScanThreads / 32];
    return warp_sums[warp] + inc - v;
}

// pass 1: sum of the eobs of each tile of kScanTile blocks
__global__ void __launch_bounds__(kScanThreads) eob_tile_sums_kernel(const uint16_t* __restrict__ eob, int n_blocks, uint32_t* __restrict__ tile_sums) {
    __shared__ uint32_t ws[kScanThreads / 32 + 1];
    const int base = blockIdx.x * kScanTile + threadIdx.x * kScanPer;
    uint32_t s = 0;
#pragma unroll
    for (int i = 0; i < kScanPer; i++) if (base + i < n_blocks) s += eob[base + i];
    uint32_t total;
    cta_exclusive_scan(s, ws, total);
    if (threadIdx.x == 0) tile_sums[blockIdx.x] = total;
}
// pass 2: exclusive scan of the tile sums by one CTA (any number of tiles: chunks of 256 with a running carry);
// tile_sums[n_tiles] receives the grand total, also written to *total_out and the overflow flag is passed through
__global__ void __launch_bounds__(kScanThreads) tile_scan_kernel(uint32_t* __restrict__ tile_sums, int n_tiles, uint32_t* __restrict__ total_out,
                                                                 const uint32_t* __restrict__ carry_in) {
    __shared__ uint32_t ws[kScanThreads / 32 + 1];
    uint32_t carry = carry_in ? carry_in[0] : 0u;                       // levels of the chunks before this one
    for (int t0 = 0; t0 < n_tiles; t0 += kScanThreads) {
        const int t = t0 + threadIdx.x;
        const uint32_t v = t < n_tiles ? tile_sums[t] : 0u;
        uint32_t total;
        const uint32_t ex = cta_exclusive_scan(v, ws, total);
        if (t < n_tiles) tile_sums[t] = carry + ex;
        carry += total;
        __syncthreads();
    }
    if (threadIdx.x == 0) { tile_sums[n_tiles] = carry; total_out[0] = carry; }
}
// pass 3: offsets of every block (offsets[n_blocks] = total)
__global__ void __launch_bounds__(kScanThreads) eob_offsets_kernel(const uint16_t* __restrict__ eob, int n_blocks, const uint32_t* __restrict__ tile_sums,
                                                                   uint32_t* __restrict__ offsets) {
    __shared__ uint32_t ws[kScanThreads / 32 + 1];
    const int base = blockIdx.x * kScanTile + threadIdx.x * kScanPer;
    uint32_t e[kScanPer], s = 0;
#pragma unroll
    for (int i = 0; i < kScanPer; i++) { e[i] = base + i < n_blocks ? eob[base + i] : 0u; s += e[i]; }
    uint32_t total;
    uint32_t o = tile_sums[blockIdx.x] + cta_exclusive_scan(s, ws, total);
#pragma unroll
    for (int i = 0; i < kScanPer; i++) { if (base + i < n_blocks) offsets[base + i] = o; o += e[i]; }
    if (blockIdx.x == gridDim.x - 1 && threadIdx.x == kScanThreads - 1) offsets[n_blocks] = o;
}
// gather: T lanes per block walk its scan up to eob. flags[1] is set when a level does not fit int16.
template <int T>
__global__ void __launch_bounds__(256) gather_levels_kernel(const int32_t* __restrict__ qcoeff, const uint16_t* __restrict__ eob,
                                                            const uint32_t* __restrict__ offsets, const int16_t* __restrict__ scan, int n_coef,
                                                            int n_blocks, int16_t* __restrict__ levels, uint32_t* __restrict__ flags) {
    const int b = (blockIdx.x * 256 + threadIdx.x) / T, l = threadIdx.x % T;
    if (b >= n_blocks) return;
    const int n = eob[b];
    const int32_t* q = qcoeff + (size_t)b * n_coef;
    int16_t* out = levels + offsets[b];
    bool wide = false;
    for (int i = l; i < n; i += T) {
        const int32_t v = q[__ldg(scan + i)];
        wide |= v != (int32_t)(int16_t)v;
        out[i] = (int16_t)v;
    }
    if (wide) atomicOr(&flags[1], 1u);
}

// scratch: (n_tiles + 1) words of tile sums. flags[0] = running total of levels (carry + this call's), flags[1] = overflow.
void launch_txq_compact(const int32_t* d_qcoeff, const uint16_t* d_eob, int n_blocks, int n_coef, const int16_t* d_scan,
                        uint32_t* d_tile_sums, uint32_t* d_offsets, int16_t* d_levels, uint32_t* d_flags, const uint32_t* d_carry, cudaStream_t st) {
    const int n_tiles = (n_blocks + kScanTile - 1) / kScanTile;
    eob_tile_sums_kernel<<<n_tiles, kScanThreads, 0, st>>>(d_eob, n_blocks, d_tile_sums);
    tile_scan_kernel<<<1, kScanThreads, 0, st>>>(d_tile_sums, n_tiles, d_flags, d_carry);
    eob_offsets_kernel<<<n_tiles, kScanThreads, 0, st>>>(d_eob, n_blocks, d_tile_sums, d_offsets);
    if (n_coef <= 16) gather_levels_kernel<16><<<(n_blocks * 16 + 255) / 256, 256, 0, st>>>(d_qcoeff, d_eob, d_offsets, d_scan, n_coef, n_blocks, d_levels, d_flags);
    else              gather_levels_kernel<32><<<((size_t)n_blocks * 32 + 255) / 256, 256, 0, st>>>(d_qcoeff, d_eob, d_offsets, d_scan, n_coef, n_blocks, d_levels, d_flags);
}

}  // namespace svtme
