// zz_sad.cu -- zero-motion SADs of the ME process (ComputeDecimatedZzSad, reference
// Source/Lib/Codec/EbMotionEstimationProcess.c:219-349): per complete 64x64 SB the plain 16x16 SAD
// (Compute16xMSad_AVX2_INTRIN through NxMSadKernel_funcPtrArray[asm][2], EbComputeSAD.h:143) between the current
// picture's sixteenth plane and the previous picture decimated by 4 (Decimation2D, step 4 = sample (4x, 4y) -- exactly what
// the previous picture's own sixteenth plane holds, so both operands are pyramid planes already resident in HBM), then
// the BEA classification (EbDefinitions.h:2770-2790).
// One warp per SB: lane = (row, 8-pixel half), two 8-byte loads, two VABSDIFF4, one warp reduction. 512 bytes per SB: the
// kernel is a handful of microseconds per picture and bound by launch latency, so jobs are batched into one launch.
#include "me_common.cuh"

namespace svtme {

__global__ void __launch_bounds__(256) zz_sad_kernel(const ZzJob* __restrict__ jobs, int width, int height, int sb_cols, int n_sb) {
    const ZzJob& job = jobs[blockIdx.y];
    const int sb = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (sb >= n_sb) return;
    const int ox = (sb % sb_cols) * 64, oy = (sb / sb_cols) * 64;
    const int sb_w = min(64, width - ox), sb_h = min(64, height - oy);
    uint32_t sad = 0xFFFFFFFFu;
    uint8_t cost = 255;                                                   // INVALID_ZZ_COST
    if (sb_w == 64 && sb_h == 64) {                                       // warp-uniform
        const int y = lane >> 1, x = (lane & 1) * 8;
        const uint8_t* a = job.cur16.at((ox >> 2) + x, (oy >> 2) + y);    // 8-byte aligned: padding 16, pitch a multiple of 64, x in {0, 8} + 16 * column
        const uint8_t* b = job.prev16.at((ox >> 2) + x, (oy >> 2) + y);
        const uint2 va = *reinterpret_cast<const uint2*>(a), vb = *reinterpret_cast<const uint2*>(b);
        sad = __reduce_add_sync(0xffffffffu, sad4(va.y, vb.y, sad4(va.x, vb.x, 0u)));
        cost = sad < 256 ? 0 : sad < 512 ? 3 : sad < 1024 ? 10 : sad < 2048 ? 20 : 30;
    }
    if (lane == 0) {
        const uint32_t area = (uint32_t)((sb_w >> 2) * (sb_h >> 2));
        job.out[sb] = cost;
        job.out[n_sb + sb] = sad < area * 2 ? 0 : sad < area * 4 ? 10 : sad < area * 8 ? 20 : 30;
    }
}

void launch_zz_sad(const ZzJob* d_jobs, int n_jobs, int width, int height, int sb_cols, int n_sb, cudaStream_t st) {
    zz_sad_kernel<<<dim3((n_sb + 7) / 8, n_jobs), 256, 0, st>>>(d_jobs, width, height, sb_cols, n_sb);
}

}  // namespace svtme
