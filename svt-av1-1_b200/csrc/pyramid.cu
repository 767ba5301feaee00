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

__global__ void __launch_bounds__(256) pyramid_kernel(const uint8_t* __restrict__ luma, int luma_stride,
                                                      DPyramid pyr) {
    const DPlane pl = blockIdx.z == 0 ? pyr.full : blockIdx.z == 1 ? pyr.quarter : pyr.sixteenth;
    const int shift = blockIdx.z;                       // 0, 1, 2  -> step 1, 2, 4
    const int rows = pl.h + 2 * pl.oy;
    const int words = (pl.w + 2 * pl.ox + 3) >> 2;      // row pitch has slack for the last partial word
    const int wx = blockIdx.x * blockDim.x + threadIdx.x;
    if (wx >= words) return;
    for (int y = blockIdx.y; y < rows; y += gridDim.y) {
        const int sy = min(max(y - pl.oy, 0), pl.h - 1) << shift;
        const uint8_t* srow = luma + (size_t)sy * luma_stride;
        uint32_t v = 0;
#pragma unroll
        for (int b = 0; b < 4; b++) {
            const int sx = min(max(wx * 4 + b - pl.ox, 0), pl.w - 1) << shift;
            v |= (uint32_t)__ldg(srow + sx) << (8 * b);
        }
        *reinterpret_cast<uint32_t*>(pl.buf + (size_t)y * pl.stride + wx * 4) = v;
    }
}

void launch_pyramid(const uint8_t* d_luma, int luma_stride, const DPyramid& pyr, cudaStream_t st) {
    const int words = (pyr.full.w + 2 * pyr.full.ox + 3) >> 2;
    dim3 grid((words + 255) / 256, 148, 3);
    pyramid_kernel<<<grid, 256, 0, st>>>(d_luma, luma_stride, pyr);
}

}  // namespace svtme
