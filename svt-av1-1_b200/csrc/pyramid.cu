// pyramid.cu -- padded full / quarter / sixteenth luma planes of one picture in a single launch.
//
// Replaces, for the ME path, PadPictureToMultipleOfLcuDimensions + DecimateInputPicture
// (reference Source/Lib/Codec/EbPictureAnalysisProcess.c:4866-4937): generate_padding
// (EbMcp.c:173-216) is edge replication and Decimation2D (:100-125) is point sampling, so every
// output byte of every padded plane is  luma[clamp(y) * step][clamp(x) * step]  -- one fused,
// HBM-bound gather-and-store kernel: reads W*H bytes once, writes (W+136)(H+136) + quarter +
// sixteenth planes once (algorithmic bytes per picture: see DESIGN.md).
#include "me_common.cuh"

namespace svtme {

// Four consecutive output bytes x0 .. x0+3 of a padded plane row: source bytes clamp(x) << shift of `srow`.
// Interior words of an aligned source row come from ONE vector load (4, 8 or 16 source bytes) and a byte permute;
// the words that touch the replicated border take the byte path.
__device__ __forceinline__ uint32_t gather_word(const uint8_t* srow, int x0, int w, int shift, bool aligned) {
    if (aligned && x0 >= 0 && x0 + 3 < w) {
        if (shift == 0) return __ldg(reinterpret_cast<const uint32_t*>(srow + x0));
        if (shift == 1) {
            const uint2 v = __ldg(reinterpret_cast<const uint2*>(srow + 2 * x0));
            return __byte_perm(v.x, v.y, 0x6420);
        }
        const uint4 v = __ldg(reinterpret_cast<const uint4*>(srow + 4 * x0));
        return __byte_perm(__byte_perm(v.x, v.y, 0x0040), __byte_perm(v.z, v.w, 0x0040), 0x5410);
    }
    uint32_t v = 0;
#pragma unroll
    for (int b = 0; b < 4; b++) v |= (uint32_t)__ldg(srow + (min(max(x0 + b, 0), w - 1) << shift)) << (8 * b);
    return v;
}

__global__ void __launch_bounds__(256) pyramid_kernel(const uint8_t* __restrict__ luma, int luma_stride,
                                                      DPyramid pyr) {
    const DPlane pl = blockIdx.z == 0 ? pyr.full : blockIdx.z == 1 ? pyr.quarter : pyr.sixteenth;
    const int shift = blockIdx.z;                       // 0, 1, 2  -> step 1, 2, 4
    const int rows = pl.h + 2 * pl.oy;
    const int words = (pl.w + 2 * pl.ox + 3) >> 2;      // row pitch has slack for the last partial word
    const int wx = blockIdx.x * blockDim.x + threadIdx.x;
    if (wx >= words) return;
    // pad (68 / 32 / 16) and the picture width are multiples of 4, so x0 << shift is 4 / 8 / 16-byte aligned inside a row
    const bool aligned = (((uintptr_t)luma | (uintptr_t)luma_stride) & 15) == 0 && (pl.ox & 3) == 0;
    for (int y = blockIdx.y; y < rows; y += gridDim.y) {
        const int sy = min(max(y - pl.oy, 0), pl.h - 1) << shift;
        *reinterpret_cast<uint32_t*>(pl.buf + (size_t)y * pl.stride + wx * 4) = gather_word(luma + (size_t)sy * luma_stride, wx * 4 - pl.ox, pl.w, shift, aligned);
    }
}

// The same for a batch of pictures of one context in ONE launch (blockIdx.z = 3 * picture + level): a single picture is
// 5.4 MB at 1080p, i.e. about one microsecond of HBM time -- per-picture launches are launch-latency bound.
// All slots of a context share the plane geometry; only the base pointers differ.
__global__ void __launch_bounds__(256) pyramid_batch_kernel(PyramidBatch batch, int luma_stride, DPyramid geom, const uint8_t* geom_base) {
    const int pic = blockIdx.z / 3, shift = blockIdx.z % 3;
    DPlane pl = shift == 0 ? geom.full : shift == 1 ? geom.quarter : geom.sixteenth;
    pl.buf = batch.base[pic] + (pl.buf - geom_base);
    const uint8_t* luma = batch.luma[pic];
    const int rows = pl.h + 2 * pl.oy;
    const int words = (pl.w + 2 * pl.ox + 3) >> 2;
    const int wx = blockIdx.x * blockDim.x + threadIdx.x;
    if (wx >= words) return;
    // pad (68 / 32 / 16) and the picture width are multiples of 4, so x0 << shift is 4 / 8 / 16-byte aligned inside a row
    const bool aligned = (((uintptr_t)luma | (uintptr_t)luma_stride) & 15) == 0 && (pl.ox & 3) == 0;
    for (int y = blockIdx.y; y < rows; y += gridDim.y) {
        const int sy = min(max(y - pl.oy, 0), pl.h - 1) << shift;
        *reinterpret_cast<uint32_t*>(pl.buf + (size_t)y * pl.stride + wx * 4) = gather_word(luma + (size_t)sy * luma_stride, wx * 4 - pl.ox, pl.w, shift, aligned);
    }
}

// The same with 16-byte accesses (source rows 16-byte aligned: any luma plane whose base and stride are multiples of 16).
// One thread = one 16-byte chunk column of eight consecutive padded output rows; a picture's items of all three planes form one flat index space
// (no idle threads on short rows), consecutive lanes write consecutive chunks of a row: 512-byte warp stores. Interior
// chunks read 16 / 32 / 64 source bytes with 32-bit (full plane: the 68-byte padding leaves the source 12 bytes off a
// 16-byte boundary) or 128-bit loads and a byte permute; chunks that touch the replicated border take the byte path.
struct PyramidFlat {
    uint32_t chunks[3];       // 16-byte chunks per padded row of each plane
    uint32_t inv[3];          // ceil(2^32 / chunks): item / chunks == umulhi(item, inv) for the item counts of a picture
    uint32_t first[4];        // first item of each plane, first[3] = items per picture
};
constexpr int kFlatRows = 8;   // padded rows per thread (same chunk column): amortises the index arithmetic, eight independent loads in flight
__device__ __forceinline__ uint4 pyramid_chunk(const uint8_t* srow, int x0, int w, int shift) {
    if (x0 >= 0 && x0 + 15 < w) {
        if (shift == 0) {
            // x0 = 12 (mod 16): four 32-bit loads (two aligned 128-bit loads + a word select measured 0.32 ms against 0.18 ms per
            // 121-picture launch: twice the L1 traffic)
            const uint32_t* s = reinterpret_cast<const uint32_t*>(srow + x0);
            return make_uint4(__ldg(s), __ldg(s + 1), __ldg(s + 2), __ldg(s + 3));
        }
        if (shift == 1) {
            const uint4* s = reinterpret_cast<const uint4*>(srow + 2 * x0);
            const uint4 a = __ldg(s), b = __ldg(s + 1);
            return make_uint4(__byte_perm(a.x, a.y, 0x6420), __byte_perm(a.z, a.w, 0x6420), __byte_perm(b.x, b.y, 0x6420), __byte_perm(b.z, b.w, 0x6420));
        }
        const uint4* s = reinterpret_cast<const uint4*>(srow + 4 * x0);
        uint32_t o[4];
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const uint4 v = __ldg(s + j);
            o[j] = __byte_perm(__byte_perm(v.x, v.y, 0x0040), __byte_perm(v.z, v.w, 0x0040), 0x5410);
        }
        return make_uint4(o[0], o[1], o[2], o[3]);
    }
    uint32_t o[4];
#pragma unroll
    for (int j = 0; j < 4; j++) {
        uint32_t v = 0;
#pragma unroll
        for (int b = 0; b < 4; b++) v |= (uint32_t)__ldg(srow + (min(max(x0 + 4 * j + b, 0), w - 1) << shift)) << (8 * b);
        o[j] = v;
    }
    return make_uint4(o[0], o[1], o[2], o[3]);
}
__global__ void __launch_bounds__(256) pyramid_flat_kernel(PyramidBatch batch, int luma_stride, DPyramid geom, const uint8_t* geom_base, PyramidFlat fl) {
    const uint32_t item = blockIdx.x * 256u + threadIdx.x;
    if (item >= fl.first[3]) return;
    const int shift = item >= fl.first[2] ? 2 : item >= fl.first[1] ? 1 : 0;
    DPlane pl = shift == 0 ? geom.full : shift == 1 ? geom.quarter : geom.sixteenth;
    const uint32_t i = item - fl.first[shift];
    const int g = (int)__umulhi(i, fl.inv[shift]), c = (int)(i - (uint32_t)g * fl.chunks[shift]);
    const int rows = pl.h + 2 * pl.oy, x0 = 16 * c - pl.ox;
    const uint8_t* luma = batch.luma[blockIdx.y];
    uint8_t* dst = batch.base[blockIdx.y] + (pl.buf - geom_base) + 16 * c;
    uint4 out[kFlatRows];
#pragma unroll
    for (int r = 0; r < kFlatRows; r++) {
        const int y = min(g * kFlatRows + r, rows - 1);
        out[r] = pyramid_chunk(luma + (size_t)(min(max(y - pl.oy, 0), pl.h - 1) << shift) * luma_stride, x0, pl.w, shift);
    }
#pragma unroll
    for (int r = 0; r < kFlatRows; r++) {
        const int y = g * kFlatRows + r;
        if (y < rows) *reinterpret_cast<uint4*>(dst + (size_t)y * pl.stride) = out[r];
    }
}

void launch_pyramid_batch(const PyramidBatch& batch, int n, int luma_stride, const DPyramid& geom, const uint8_t* geom_base, cudaStream_t st) {
    bool aligned = (luma_stride & 15) == 0;
    for (int i = 0; i < n; i++) aligned &= ((uintptr_t)batch.luma[i] & 15) == 0;
    if (aligned) {
        PyramidFlat fl;
        const DPlane* pls[3] = { &geom.full, &geom.quarter, &geom.sixteenth };
        uint32_t items = 0;
        for (int l = 0; l < 3; l++) {
            fl.chunks[l] = (uint32_t)(pls[l]->w + 2 * pls[l]->ox + 15) / 16;          // the row pitch (a multiple of 64) holds the last partial chunk
            fl.inv[l] = (uint32_t)((((uint64_t)1 << 32) + fl.chunks[l] - 1) / fl.chunks[l]);
            fl.first[l] = items;
            items += fl.chunks[l] * (uint32_t)((pls[l]->h + 2 * pls[l]->oy + kFlatRows - 1) / kFlatRows);     // items = (group of kFlatRows rows, chunk)
        }
        fl.first[3] = items;
        pyramid_flat_kernel<<<dim3((items + 255) / 256, n), 256, 0, st>>>(batch, luma_stride, geom, geom_base, fl);
        return;
    }
    const int words = (geom.full.w + 2 * geom.full.ox + 3) >> 2;
    dim3 grid((words + 255) / 256, 37, 3 * n);
    pyramid_batch_kernel<<<grid, 256, 0, st>>>(batch, luma_stride, geom, geom_base);
}

// 10-bit input split (UnPack2D -> EB_ENC_msbUnPack2D, C_DEFAULT/EbPackUnPack_C.c:127-151): msb = sample >> 2 (the plane ME
// searches), rem = (uint8_t)(sample << 6). One thread = four samples: two 64-bit loads, one or two 32-bit stores.
__global__ void __launch_bounds__(256) unpack16_kernel(const uint16_t* __restrict__ in, int in_stride, uint8_t* __restrict__ msb,
                                                       uint8_t* __restrict__ rem, int width, int height) {
    const int x = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (x >= width) return;
    for (int y = blockIdx.y; y < height; y += gridDim.y) {
        const uint2 v = *reinterpret_cast<const uint2*>(in + (size_t)y * in_stride + x);      // width % 8 == 0, rows 8-byte aligned
        const uint32_t s0 = v.x & 0xFFFFu, s1 = v.x >> 16, s2 = v.y & 0xFFFFu, s3 = v.y >> 16;
        const uint32_t m = ((s0 >> 2) & 0xFF) | (((s1 >> 2) & 0xFF) << 8) | (((s2 >> 2) & 0xFF) << 16) | (((s3 >> 2) & 0xFF) << 24);
        *reinterpret_cast<uint32_t*>(msb + (size_t)y * width + x) = m;
        if (rem) {
            // quirk Q9: the reference's width == 32 / width == 64 special cases leave the last 16 samples unshifted
            const bool q9 = (width == 64 && x >= 48) || (width == 32 && (y & 1) && x >= 16);
            const uint32_t lo = (s0 & 3) | ((s1 & 3) << 8) | ((s2 & 3) << 16) | ((s3 & 3) << 24);
            const uint32_t r = q9 ? lo : lo << 6;
            *reinterpret_cast<uint32_t*>(rem + (size_t)y * width + x) = r;
        }
    }
}

void launch_unpack16(const uint16_t* d_in, int in_stride, uint8_t* d_msb, uint8_t* d_rem, int width, int height, cudaStream_t st) {
    dim3 grid((width / 4 + 255) / 256, height < 296 ? height : 296);
    unpack16_kernel<<<grid, 256, 0, st>>>(d_in, in_stride, d_msb, d_rem, width, height);
}

void launch_pyramid(const uint8_t* d_luma, int luma_stride, const DPyramid& pyr, cudaStream_t st) {
    if ((((uintptr_t)d_luma | (uintptr_t)luma_stride) & 15) == 0) {          // 16-byte path: a batch of one, offsets relative to the full plane
        PyramidBatch one;
        one.luma[0] = d_luma; one.base[0] = pyr.full.buf;
        launch_pyramid_batch(one, 1, luma_stride, pyr, pyr.full.buf, st);
        return;
    }
    const int words = (pyr.full.w + 2 * pyr.full.ox + 3) >> 2;
    dim3 grid((words + 255) / 256, 148, 3);
    pyramid_kernel<<<grid, 256, 0, st>>>(d_luma, luma_stride, pyr);
}

}  // namespace svtme
