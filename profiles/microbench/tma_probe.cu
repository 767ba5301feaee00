// Probe of the TMA path on this box: argv[1] = mode
//   0: 1-D bulk copy (cp.async.bulk) + mbarrier          1: one 2-D tensor box 16 x 8
//   2: one box 144 x 128                                  3: four byte-shifted boxes 144 x 128 (the full-pel window)
//   4: as 3, descriptor passed as __grid_constant__ kernel parameter instead of global memory
// Every mode checks the bytes against ordinary loads. Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <vector>
#include <cuda.h>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t sptr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(sptr(bar)), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(sptr(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile("{\n\t.reg .pred p;\n\tWAIT_%=:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra DONE_%=;\n\tbra WAIT_%=;\n\tDONE_%=:\n\t}"
                 :: "r"(sptr(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void tma_load_2d(void* dst, const void* tmap, int x, int y, uint64_t* bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 :: "r"(sptr(dst)), "l"(tmap), "r"(sptr(bar)), "r"(x), "r"(y) : "memory");
}
__device__ __forceinline__ void bulk_load_1d(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :: "r"(sptr(dst)), "l"(src), "r"(bytes), "r"(sptr(bar)) : "memory");
}

__global__ void probe(const __grid_constant__ CUtensorMap pm, int use_param, const void* gmap, int mode, int boxb, int boxr, int ncopy,
                      const uint8_t* plane, int stride, int x0, int y0, int* bad) {
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ __align__(8) uint64_t bar;
    const void* tmap = use_param ? (const void*)&pm : gmap;
    const int bytes = boxb * boxr;
    if (threadIdx.x == 0) {
        mbar_init(&bar, 1);
        mbar_expect_tx(&bar, (uint32_t)(ncopy * bytes));
        if (mode == 0) bulk_load_1d(smem, plane + (size_t)y0 * stride, (uint32_t)bytes, &bar);
        else for (int sc = 0; sc < ncopy; sc++) tma_load_2d(smem + sc * bytes, tmap, x0 + sc, y0, &bar);
    }
    __syncthreads();
    mbar_wait(&bar, 0);
    int n = 0;
    for (int t = threadIdx.x; t < ncopy * bytes; t += blockDim.x) {
        const int sc = t / bytes, r = (t / boxb) % boxr, c = t % boxb;
        const uint8_t want = mode == 0 ? plane[(size_t)y0 * stride + t] : plane[(size_t)(y0 + r) * stride + x0 + sc + c];
        if (smem[t] != want) n++;
    }
    if (n) atomicAdd(bad, n);
}

int main(int argc, char** argv) {
    const int mode = argc > 1 ? atoi(argv[1]) : 3;
    const int stride = 2112, rows = 1216;
    std::vector<uint8_t> h((size_t)stride * rows + 256);
    for (size_t i = 0; i < h.size(); i++) h[i] = (uint8_t)((i * 2654435761u) >> 13);
    uint8_t* d; cudaMalloc(&d, h.size()); cudaMemcpy(d, h.data(), h.size(), cudaMemcpyHostToDevice);
    const int boxb = mode == 0 ? 4096 : mode == 1 ? 16 : 144, boxr = mode == 0 ? 1 : mode == 1 ? 8 : 128, ncopy = mode >= 3 ? 4 : 1;
    typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                 const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    void* fn = nullptr; cudaDriverEntryPointQueryResult q;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
    CUtensorMap m; memset(&m, 0, sizeof(m));
    const cuuint64_t dims[2] = { (cuuint64_t)stride, (cuuint64_t)rows }, strides[1] = { (cuuint64_t)stride };
    const cuuint32_t box[2] = { (cuuint32_t)(mode == 0 ? 16 : boxb), (cuuint32_t)(mode == 0 ? 8 : boxr) }, estr[2] = { 1, 1 };
    CUresult r = ((EncodeFn)fn)(&m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, d, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("mode %d: entry point %s q=%d, encode %d\n", mode, cudaGetErrorString(e), (int)q, (int)r);
    CUtensorMap* dm; cudaMalloc(&dm, sizeof(m)); cudaMemcpy(dm, &m, sizeof(m), cudaMemcpyHostToDevice);
    int* bad; cudaMalloc(&bad, 4); cudaMemset(bad, 0, 4);
    const int smem = ncopy * boxb * boxr;
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    for (int x0 : { 64, 68, 69, 70, 71, 131, 1000 }) {
        probe<<<1, 256, smem>>>(m, mode == 4, dm, mode, boxb, boxr, ncopy, d, stride, x0, 37, bad);
        e = cudaDeviceSynchronize();
        int hb = -1; cudaMemcpy(&hb, bad, 4, cudaMemcpyDeviceToHost);
        printf("  x0 %d: %s, mismatching bytes %d\n", x0, cudaGetErrorString(e), hb);
        if (e != cudaSuccess) return 1;
    }
    return 0;
}
