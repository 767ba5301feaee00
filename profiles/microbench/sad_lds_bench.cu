// How much do the shared-memory loads of the full-pel inner loop cost the SAD pipe? One "row" = 64 VABSDIFF4.ACC fed by
// NW window LDS.128 (every lane its own 16 bytes) and NS source LDS.128 (one address per warp = broadcast).
//   (5,4) = the shipped kernel: 4 positions x 16 source words;  (4,2) = 8 positions x 8 source words (half an SB row per thread);
//   (0,0) = SADs alone. Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

template <int NW, int NS>
__global__ void __launch_bounds__(256, 2) rows(uint32_t* out, int iters) {
    extern __shared__ __align__(16) uint32_t sm[];
    for (int t = threadIdx.x; t < 20480; t += 256) sm[t] = t * 2654435761u;
    __syncthreads();
    uint32_t acc[32];
#pragma unroll
    for (int i = 0; i < 32; i++) acc[i] = 0;
    const uint4* win = reinterpret_cast<const uint4*>(sm) + (threadIdx.x & 31) * 9 + (threadIdx.x >> 5) * 36;
    const uint4* src = reinterpret_cast<const uint4*>(sm + 18432);
    for (int it = 0; it < iters; it++) {
        uint32_t ref[24], sw[16];
#pragma unroll
        for (int i = 0; i < 6; i++) {
            uint4 v = i < NW ? win[(it & 7) * 288 + i] : make_uint4(it, i, it + i, 7);
            ref[4 * i] = v.x; ref[4 * i + 1] = v.y; ref[4 * i + 2] = v.z; ref[4 * i + 3] = v.w;
        }
#pragma unroll
        for (int i = 0; i < 4; i++) {
            uint4 v = i < NS ? src[(it & 31) * 4 + i] : make_uint4(it ^ 5, i, it - i, 9);
            sw[4 * i] = v.x; sw[4 * i + 1] = v.y; sw[4 * i + 2] = v.z; sw[4 * i + 3] = v.w;
        }
#pragma unroll
        for (int c = 0; c < 16; c++)
#pragma unroll
            for (int k = 0; k < 4; k++)
                asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(acc[(k * 8 + (c >> 1)) & 31]) : "r"(sw[c]), "r"(ref[c + k]));
    }
    uint32_t s = 0;
#pragma unroll
    for (int i = 0; i < 32; i++) s += acc[i];
    out[blockIdx.x * 256 + threadIdx.x] = s;
}

template <int NW, int NS>
static void run(uint32_t* d_out) {
    const int iters = 4096, blocks = 148 * 2 * 4;
    cudaFuncSetAttribute(rows<NW, NS>, cudaFuncAttributeMaxDynamicSharedMemorySize, 20480 * 4);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    rows<NW, NS><<<blocks, 256, 20480 * 4>>>(d_out, iters); cudaDeviceSynchronize();
    float best = 1e30f;
    for (int r = 0; r < 5; r++) {
        cudaEventRecord(e0); rows<NW, NS><<<blocks, 256, 20480 * 4>>>(d_out, iters); cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
    }
    const double lane_sads = (double)blocks * 256 * iters * 64;
    printf("{\"window_lds128\": %d, \"source_lds128_broadcast\": %d, \"ms\": %.4f, \"sad_lane_ops_per_ns_per_sm\": %.2f, \"err\": \"%s\"}\n", NW, NS, best,
           lane_sads / best / 1e6 / 148, cudaGetErrorString(cudaGetLastError()));
}

int main() {
    uint32_t* d_out; cudaMalloc(&d_out, 148 * 8 * 256 * 4);
    run<0, 0>(d_out); run<5, 4>(d_out); run<4, 2>(d_out); run<5, 0>(d_out); run<0, 4>(d_out); run<3, 2>(d_out);
    return 0;
}
