// Micro-benchmark: issue rates of the integer instructions the ME kernels are built from,
// on the device it runs on (B200 / sm_100a). Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3
// Output: one JSON line per instruction mix with Gop/s (warp-lane ops) and ops/clk/SM.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define ITERS 4096
template <int MODE>
__global__ void __launch_bounds__(256) bench(uint32_t* out, uint32_t seed) {
    uint32_t a0 = threadIdx.x * 0x01010101u + seed, a1 = a0 * 3u, a2 = a0 * 5u, a3 = a0 * 7u;
    uint32_t b0 = seed ^ 0x55aa55aau, b1 = b0 + 1u, b2 = b0 + 2u, b3 = b0 + 3u;
    uint32_t c0 = 0, c1 = 0, c2 = 0, c3 = 0, c4 = 0, c5 = 0, c6 = 0, c7 = 0;
    __shared__ uint32_t sm[1024];
    sm[threadIdx.x] = a0; sm[threadIdx.x + 256] = a1; sm[threadIdx.x + 512] = a2; sm[threadIdx.x + 768] = a3;
    __syncthreads();
#pragma unroll 1
    for (int i = 0; i < ITERS; i++) {
        if (MODE == 0) {  // 8 x VABSDIFF4.ACC, independent chains
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c0) : "r"(a0), "r"(b0));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c1) : "r"(a1), "r"(b1));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c2) : "r"(a2), "r"(b2));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c3) : "r"(a3), "r"(b3));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c4) : "r"(a0), "r"(b1));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c5) : "r"(a1), "r"(b2));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c6) : "r"(a2), "r"(b3));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c7) : "r"(a3), "r"(b0));
        } else if (MODE == 1) {  // 8 x VABSDIFF4 + 4 x SHF (funnel shift)
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c0) : "r"(a0), "r"(b0));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c1) : "r"(a1), "r"(b1));
            asm volatile("shf.r.wrap.b32 %0, %1, %2, 8;" : "=r"(b0) : "r"(b0), "r"(b1));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c2) : "r"(a2), "r"(b2));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c3) : "r"(a3), "r"(b3));
            asm volatile("shf.r.wrap.b32 %0, %1, %2, 8;" : "=r"(b1) : "r"(b1), "r"(b2));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c4) : "r"(a0), "r"(b1));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c5) : "r"(a1), "r"(b2));
            asm volatile("shf.r.wrap.b32 %0, %1, %2, 8;" : "=r"(b2) : "r"(b2), "r"(b3));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c6) : "r"(a2), "r"(b3));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c7) : "r"(a3), "r"(b0));
            asm volatile("shf.r.wrap.b32 %0, %1, %2, 8;" : "=r"(b3) : "r"(b3), "r"(b0));
        } else if (MODE == 2) {  // 8 x VABSDIFF4 + 4 x IMAD (mad.lo)
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c0) : "r"(a0), "r"(b0));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c1) : "r"(a1), "r"(b1));
            asm volatile("mad.lo.u32 %0, %1, 4099, %0;" : "+r"(b0) : "r"(c0));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c2) : "r"(a2), "r"(b2));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c3) : "r"(a3), "r"(b3));
            asm volatile("mad.lo.u32 %0, %1, 4099, %0;" : "+r"(b1) : "r"(c1));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c4) : "r"(a0), "r"(b1));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c5) : "r"(a1), "r"(b2));
            asm volatile("mad.lo.u32 %0, %1, 4099, %0;" : "+r"(b2) : "r"(c2));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c6) : "r"(a2), "r"(b3));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c7) : "r"(a3), "r"(b0));
            asm volatile("mad.lo.u32 %0, %1, 4099, %0;" : "+r"(b3) : "r"(c3));
        } else if (MODE == 3) {  // 8 x VABSDIFF4 + 2 x LDS.128
            uint4 v = *reinterpret_cast<uint4*>(&sm[(threadIdx.x * 4 + i) & 1020]);
            uint4 w = *reinterpret_cast<uint4*>(&sm[(i * 4) & 1020]);
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c0) : "r"(v.x), "r"(w.x));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c1) : "r"(v.y), "r"(w.y));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c2) : "r"(v.z), "r"(w.z));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c3) : "r"(v.w), "r"(w.w));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c4) : "r"(v.x), "r"(w.y));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c5) : "r"(v.y), "r"(w.z));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c6) : "r"(v.z), "r"(w.w));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c7) : "r"(v.w), "r"(w.x));
        } else if (MODE == 4) {  // 8 x IADD3-class adds (baseline ALU rate)
            asm volatile("add.u32 %0, %0, %1;" : "+r"(c0) : "r"(a0)); asm volatile("add.u32 %0, %0, %1;" : "+r"(c1) : "r"(a1));
            asm volatile("add.u32 %0, %0, %1;" : "+r"(c2) : "r"(a2)); asm volatile("add.u32 %0, %0, %1;" : "+r"(c3) : "r"(a3));
            asm volatile("add.u32 %0, %0, %1;" : "+r"(c4) : "r"(b0)); asm volatile("add.u32 %0, %0, %1;" : "+r"(c5) : "r"(b1));
            asm volatile("add.u32 %0, %0, %1;" : "+r"(c6) : "r"(b2)); asm volatile("add.u32 %0, %0, %1;" : "+r"(c7) : "r"(b3));
        } else if (MODE == 5) {  // 8 x min.u32 (VIMNMX)
            asm volatile("min.u32 %0, %0, %1;" : "+r"(c0) : "r"(a0)); asm volatile("min.u32 %0, %0, %1;" : "+r"(c1) : "r"(a1));
            asm volatile("min.u32 %0, %0, %1;" : "+r"(c2) : "r"(a2)); asm volatile("min.u32 %0, %0, %1;" : "+r"(c3) : "r"(a3));
            asm volatile("min.u32 %0, %0, %1;" : "+r"(c4) : "r"(b0)); asm volatile("min.u32 %0, %0, %1;" : "+r"(c5) : "r"(b1));
            asm volatile("min.u32 %0, %0, %1;" : "+r"(c6) : "r"(b2)); asm volatile("min.u32 %0, %0, %1;" : "+r"(c7) : "r"(b3));
            a0 += i; a1 ^= c0;
        } else if (MODE == 6) {  // 4 x (vabsdiff4 no-acc + dp4a) : the two-instruction SAD alternative
            uint32_t d0, d1, d2, d3;
            asm volatile("vabsdiff4.u32.u32.u32 %0, %1, %2, %3;" : "=r"(d0) : "r"(a0), "r"(b0), "r"(0));
            asm volatile("vabsdiff4.u32.u32.u32 %0, %1, %2, %3;" : "=r"(d1) : "r"(a1), "r"(b1), "r"(0));
            asm volatile("vabsdiff4.u32.u32.u32 %0, %1, %2, %3;" : "=r"(d2) : "r"(a2), "r"(b2), "r"(0));
            asm volatile("vabsdiff4.u32.u32.u32 %0, %1, %2, %3;" : "=r"(d3) : "r"(a3), "r"(b3), "r"(0));
            c0 = __dp4a(d0, 0x01010101u, c0); c1 = __dp4a(d1, 0x01010101u, c1);
            c2 = __dp4a(d2, 0x01010101u, c2); c3 = __dp4a(d3, 0x01010101u, c3);
            b0 += c0; b1 += c1;
        } else if (MODE == 7) {  // 8 x dp4a alone
            c0 = __dp4a(a0, b0, c0); c1 = __dp4a(a1, b1, c1); c2 = __dp4a(a2, b2, c2); c3 = __dp4a(a3, b3, c3);
            c4 = __dp4a(a0, b1, c4); c5 = __dp4a(a1, b2, c5); c6 = __dp4a(a2, b3, c6); c7 = __dp4a(a3, b0, c7);
        } else if (MODE == 8) {  // 8 x VABSDIFF4 + 4 x min.u32
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c0) : "r"(a0), "r"(b0));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c1) : "r"(a1), "r"(b1));
            asm volatile("min.u32 %0, %0, %1;" : "+r"(b0) : "r"(c0));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c2) : "r"(a2), "r"(b2));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c3) : "r"(a3), "r"(b3));
            asm volatile("min.u32 %0, %0, %1;" : "+r"(b1) : "r"(c1));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c4) : "r"(a0), "r"(b1));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c5) : "r"(a1), "r"(b2));
            asm volatile("min.u32 %0, %0, %1;" : "+r"(b2) : "r"(c2));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c6) : "r"(a2), "r"(b3));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c7) : "r"(a3), "r"(b0));
            asm volatile("min.u32 %0, %0, %1;" : "+r"(b3) : "r"(c3));
        } else if (MODE == 9) {  // 8 x VABSDIFF4 + 4 x min.f32
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c0) : "r"(a0), "r"(b0));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c1) : "r"(a1), "r"(b1));
            asm volatile("min.f32 %0, %0, %1;" : "+f"(*(float*)&b0) : "f"(*(float*)&c0));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c2) : "r"(a2), "r"(b2));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c3) : "r"(a3), "r"(b3));
            asm volatile("min.f32 %0, %0, %1;" : "+f"(*(float*)&b1) : "f"(*(float*)&c1));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c4) : "r"(a0), "r"(b1));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c5) : "r"(a1), "r"(b2));
            asm volatile("min.f32 %0, %0, %1;" : "+f"(*(float*)&b2) : "f"(*(float*)&c2));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c6) : "r"(a2), "r"(b3));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c7) : "r"(a3), "r"(b0));
            asm volatile("min.f32 %0, %0, %1;" : "+f"(*(float*)&b3) : "f"(*(float*)&c3));
        } else if (MODE == 10) {  // 8 x VABSDIFF4 + 4 x min.f16x2
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c0) : "r"(a0), "r"(b0));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c1) : "r"(a1), "r"(b1));
            asm volatile("min.f16x2 %0, %0, %1;" : "+r"(b0) : "r"(c0));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c2) : "r"(a2), "r"(b2));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c3) : "r"(a3), "r"(b3));
            asm volatile("min.f16x2 %0, %0, %1;" : "+r"(b1) : "r"(c1));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c4) : "r"(a0), "r"(b1));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c5) : "r"(a1), "r"(b2));
            asm volatile("min.f16x2 %0, %0, %1;" : "+r"(b2) : "r"(c2));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c6) : "r"(a2), "r"(b3));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c7) : "r"(a3), "r"(b0));
            asm volatile("min.f16x2 %0, %0, %1;" : "+r"(b3) : "r"(c3));
        } else if (MODE == 11) {  // 8 x VABSDIFF4 + 4 x min.u16x2
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c0) : "r"(a0), "r"(b0));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c1) : "r"(a1), "r"(b1));
            asm volatile("min.u16x2 %0, %0, %1;" : "+r"(b0) : "r"(c0));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c2) : "r"(a2), "r"(b2));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c3) : "r"(a3), "r"(b3));
            asm volatile("min.u16x2 %0, %0, %1;" : "+r"(b1) : "r"(c1));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c4) : "r"(a0), "r"(b1));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c5) : "r"(a1), "r"(b2));
            asm volatile("min.u16x2 %0, %0, %1;" : "+r"(b2) : "r"(c2));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c6) : "r"(a2), "r"(b3));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c7) : "r"(a3), "r"(b0));
            asm volatile("min.u16x2 %0, %0, %1;" : "+r"(b3) : "r"(c3));
        } else if (MODE == 12) {  // 8 x VABSDIFF4 + 4 x min3.u32 (min(min))
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c0) : "r"(a0), "r"(b0));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c1) : "r"(a1), "r"(b1));
            asm volatile("{.reg .u32 t; min.u32 t, %0, %1; min.u32 %0, t, %2;}" : "+r"(b0) : "r"(c0), "r"(a3));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c2) : "r"(a2), "r"(b2));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c3) : "r"(a3), "r"(b3));
            asm volatile("{.reg .u32 t; min.u32 t, %0, %1; min.u32 %0, t, %2;}" : "+r"(b1) : "r"(c1), "r"(a3));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c4) : "r"(a0), "r"(b1));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c5) : "r"(a1), "r"(b2));
            asm volatile("{.reg .u32 t; min.u32 t, %0, %1; min.u32 %0, t, %2;}" : "+r"(b2) : "r"(c2), "r"(a3));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c6) : "r"(a2), "r"(b3));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c7) : "r"(a3), "r"(b0));
            asm volatile("{.reg .u32 t; min.u32 t, %0, %1; min.u32 %0, t, %2;}" : "+r"(b3) : "r"(c3), "r"(a3));
        } else if (MODE == 13) {  // 8 x VABSDIFF4 + 4 x fma.rn.f32
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c0) : "r"(a0), "r"(b0));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c1) : "r"(a1), "r"(b1));
            asm volatile("fma.rn.f32 %0, %0, %1, %1;" : "+f"(*(float*)&b0) : "f"(*(float*)&c0));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c2) : "r"(a2), "r"(b2));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c3) : "r"(a3), "r"(b3));
            asm volatile("fma.rn.f32 %0, %0, %1, %1;" : "+f"(*(float*)&b1) : "f"(*(float*)&c1));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c4) : "r"(a0), "r"(b1));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c5) : "r"(a1), "r"(b2));
            asm volatile("fma.rn.f32 %0, %0, %1, %1;" : "+f"(*(float*)&b2) : "f"(*(float*)&c2));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c6) : "r"(a2), "r"(b3));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c7) : "r"(a3), "r"(b0));
            asm volatile("fma.rn.f32 %0, %0, %1, %1;" : "+f"(*(float*)&b3) : "f"(*(float*)&c3));
        } else if (MODE == 14) {  // 8 x VABSDIFF4 + 4 x lop3 xor
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c0) : "r"(a0), "r"(b0));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c1) : "r"(a1), "r"(b1));
            asm volatile("xor.b32 %0, %0, %1;" : "+r"(b0) : "r"(c0));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c2) : "r"(a2), "r"(b2));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c3) : "r"(a3), "r"(b3));
            asm volatile("xor.b32 %0, %0, %1;" : "+r"(b1) : "r"(c1));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c4) : "r"(a0), "r"(b1));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c5) : "r"(a1), "r"(b2));
            asm volatile("xor.b32 %0, %0, %1;" : "+r"(b2) : "r"(c2));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c6) : "r"(a2), "r"(b3));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c7) : "r"(a3), "r"(b0));
            asm volatile("xor.b32 %0, %0, %1;" : "+r"(b3) : "r"(c3));
        } else if (MODE == 15) {  // 8 x VABSDIFF4 + 4 x prmt
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c0) : "r"(a0), "r"(b0));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c1) : "r"(a1), "r"(b1));
            asm volatile("prmt.b32 %0, %0, %1, 0x5410;" : "+r"(b0) : "r"(c0));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c2) : "r"(a2), "r"(b2));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c3) : "r"(a3), "r"(b3));
            asm volatile("prmt.b32 %0, %0, %1, 0x5410;" : "+r"(b1) : "r"(c1));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c4) : "r"(a0), "r"(b1));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c5) : "r"(a1), "r"(b2));
            asm volatile("prmt.b32 %0, %0, %1, 0x5410;" : "+r"(b2) : "r"(c2));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c6) : "r"(a2), "r"(b3));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c7) : "r"(a3), "r"(b0));
            asm volatile("prmt.b32 %0, %0, %1, 0x5410;" : "+r"(b3) : "r"(c3));
        } else if (MODE == 16) {  // 8 x VABSDIFF4 + 4 x sel (setp+selp)
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c0) : "r"(a0), "r"(b0));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c1) : "r"(a1), "r"(b1));
            asm volatile("{.reg .pred p; setp.lt.u32 p, %0, %1; selp.u32 %0, %0, %1, p;}" : "+r"(b0) : "r"(c0));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c2) : "r"(a2), "r"(b2));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c3) : "r"(a3), "r"(b3));
            asm volatile("{.reg .pred p; setp.lt.u32 p, %0, %1; selp.u32 %0, %0, %1, p;}" : "+r"(b1) : "r"(c1));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c4) : "r"(a0), "r"(b1));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c5) : "r"(a1), "r"(b2));
            asm volatile("{.reg .pred p; setp.lt.u32 p, %0, %1; selp.u32 %0, %0, %1, p;}" : "+r"(b2) : "r"(c2));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c6) : "r"(a2), "r"(b3));
            asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(c7) : "r"(a3), "r"(b0));
            asm volatile("{.reg .pred p; setp.lt.u32 p, %0, %1; selp.u32 %0, %0, %1, p;}" : "+r"(b3) : "r"(c3));
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = c0 + c1 + c2 + c3 + c4 + c5 + c6 + c7 + b0 + b1 + b2 + b3;
}

template <int MODE>
static void run(const char* name, double ops_per_iter, uint32_t* d_out) {
    int dev; cudaGetDevice(&dev); cudaDeviceProp p; cudaGetDeviceProperties(&p, dev);
    const int blocks = p.multiProcessorCount * 8;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    bench<MODE><<<blocks, 256>>>(d_out, 1); cudaDeviceSynchronize();
    float best = 1e30f;
    for (int r = 0; r < 5; r++) {
        cudaEventRecord(e0); bench<MODE><<<blocks, 256>>>(d_out, r); cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
    }
    double lane_ops = (double)blocks * 256 * ITERS * ops_per_iter;
    int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, dev);
    printf("{\"mix\": \"%s\", \"ms\": %.4f, \"G_lane_ops_per_s\": %.1f, \"sms\": %d, \"lane_ops_per_ns_per_sm\": %.2f, \"max_clock_khz\": %d}\n",
           name, best, lane_ops / best / 1e6, p.multiProcessorCount, lane_ops / best / 1e6 / p.multiProcessorCount, clk);
}

int main() {
    uint32_t* d_out; cudaMalloc(&d_out, 148 * 16 * 256 * 4 * 4);
    run<0>("vabsdiff4.acc x8 (counted: 8)", 8, d_out);
    run<1>("vabsdiff4.acc x8 + shf x4 (counted: 8 vabsdiff4)", 8, d_out);
    run<2>("vabsdiff4.acc x8 + imad x4 (counted: 8 vabsdiff4)", 8, d_out);
    run<3>("vabsdiff4.acc x8 + lds.128 x2 (counted: 8 vabsdiff4)", 8, d_out);
    run<4>("add.u32 x8", 8, d_out);
    run<5>("min.u32 x8", 8, d_out);
    run<6>("vabsdiff4 + dp4a x4 (counted: 4 word-sads)", 4, d_out);
    run<7>("dp4a x8", 8, d_out);
    run<8>("vabsdiff4.acc x8 + min.u32 x4 (counted: 8 vabsdiff4)", 8, d_out);
    run<9>("vabsdiff4.acc x8 + min.f32 x4 (counted: 8 vabsdiff4)", 8, d_out);
    run<10>("vabsdiff4.acc x8 + min.f16x2 x4 (counted: 8 vabsdiff4)", 8, d_out);
    run<11>("vabsdiff4.acc x8 + min.u16x2 x4 (counted: 8 vabsdiff4)", 8, d_out);
    run<12>("vabsdiff4.acc x8 + min3.u32 (min(min)) x4 (counted: 8 vabsdiff4)", 8, d_out);
    run<13>("vabsdiff4.acc x8 + fma.rn.f32 x4 (counted: 8 vabsdiff4)", 8, d_out);
    run<14>("vabsdiff4.acc x8 + lop3 xor x4 (counted: 8 vabsdiff4)", 8, d_out);
    run<15>("vabsdiff4.acc x8 + prmt x4 (counted: 8 vabsdiff4)", 8, d_out);
    run<16>("vabsdiff4.acc x8 + sel (setp+selp) x4 (counted: 8 vabsdiff4)", 8, d_out);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { printf("CUDA error %s\n", cudaGetErrorString(e)); return 1; }
    return 0;
}
