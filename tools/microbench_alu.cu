// Developer aid (run on the GPU): ALU-pipe cost of u16x2 min/max when every operand is a distinct live register (no operand
// reuse), as in the FAST strength network.  Prints clocks per warp-instruction per SM sub-partition.
#include <cstdio>
#include <cstdint>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
__device__ __forceinline__ uint32_t hmin(uint32_t a, uint32_t b) { __half2 r = __hmin2(*reinterpret_cast<__half2*>(&a), *reinterpret_cast<__half2*>(&b)); return *reinterpret_cast<uint32_t*>(&r); }
__device__ __forceinline__ uint32_t hmax(uint32_t a, uint32_t b) { __half2 r = __hmax2(*reinterpret_cast<__half2*>(&a), *reinterpret_cast<__half2*>(&b)); return *reinterpret_cast<uint32_t*>(&r); }
constexpr int ITERS = 2048, CH = 8;
template <int OP> __global__ void __launch_bounds__(256) bench(uint32_t* out, uint32_t seed) {
    uint32_t v[CH], b[CH], c[CH];
#pragma unroll
    for (int i = 0; i < CH; ++i) { v[i] = seed + i * 77 + threadIdx.x; b[i] = seed * (i + 3) + threadIdx.x * 13; c[i] = seed * (i + 11) ^ threadIdx.x; }
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < CH; ++i) {
            if (OP == 0) v[i] = __vminu2(v[i], b[i]);                                  // 2 distinct register operands
            if (OP == 1) v[i] = __vimin3_u16x2(v[i], b[i], c[i]);                      // 3 distinct register operands
            if (OP == 2) { v[i] = __vminu2(v[i], b[i]); b[i] = __vmaxu2(b[i], c[i]); } // two independent 2-input ops
            if (OP == 3) v[i] = __vminu2(v[i], b[0]);                                  // shared second operand (reuse cache)
            if (OP == 4) { v[i] = __vimin3_u16x2(v[i], b[i], c[i]); b[i] = __vmaxu2(b[i], c[(i + 1) % CH]); }   // 3-input + 2-input
            if (OP == 5) v[i] = (v[i] & b[i]) ^ c[i];                                  // LOP3, 3 operands
            if (OP == 6) v[i] = v[i] + b[i];                                           // IADD
            if (OP == 7) v[i] = hmin(v[i], b[i]);                                      // HMNMX2
            if (OP == 8) { v[i] = hmin(v[i], b[i]); b[i] = hmax(b[i], c[i]); }         // two HMNMX2
            if (OP == 9) { v[i] = __vminu2(v[i], b[i]); b[i] = hmax(b[i], c[i]); }     // VIMNMX + HMNMX2: same pipe or not?
            if (OP == 10) { v[i] = __vimin3_u16x2(v[i], b[i], c[i]); b[i] = hmax(b[i], c[(i + 1) % CH]); }
            if (OP == 11) { v[i] = v[i] * b[i] + c[i]; }                                 // IMAD (FMA pipe)
            if (OP == 12) { v[i] = __vminu2(v[i], b[i]); b[i] = b[i] * c[i] + v[(i + 1) % CH]; }   // VIMNMX + IMAD: different pipes
        }
#pragma unroll
        for (int i = 0; i < CH; ++i) { b[i] += 0x00010001u; c[i] ^= v[(i + 1) % CH] & 0x00010001u; }   // keep operands changing (2 extra ops per chain)
    }
    uint32_t s = 0;
#pragma unroll
    for (int i = 0; i < CH; ++i) s ^= v[i] ^ b[i] ^ c[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int OP> void run(const char* name, int perChain) {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    const int blocks = p.multiProcessorCount * 8;
    uint32_t* out; cudaMalloc(&out, (size_t)blocks * 256 * 4);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    bench<OP><<<blocks, 256>>>(out, 12345u); cudaDeviceSynchronize();
    cudaEventRecord(e0);
    for (int r = 0; r < 5; ++r) bench<OP><<<blocks, 256>>>(out, 12345u + r);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    int khz; cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    const double warpIters = 5.0 * blocks * 8 * ITERS * CH;                 // warp-level executions of one chain step
    const double clkPerSmsp = ms * 1e-3 * khz * 1e3 * (p.multiProcessorCount * 4);
    printf("%-40s %.3f ms  %.2f clk per chain step per SMSP (%d measured + 2 bookkeeping ALU ops per step)\n", name, ms / 5, clkPerSmsp / warpIters, perChain);
    cudaFree(out);
}
int main() {
    run<6>("iadd (2 operands) baseline", 1);
    run<5>("lop3 (3 operands)", 1);
    run<0>("vminu2, distinct operands", 1);
    run<3>("vminu2, shared second operand", 1);
    run<1>("vimin3_u16x2, distinct operands", 1);
    run<2>("vminu2 + vmaxu2", 2);
    run<4>("vimin3_u16x2 + vmaxu2", 2);
    run<7>("hmin2", 1);
    run<8>("hmin2 + hmax2", 2);
    run<9>("vminu2 + hmax2", 2);
    run<10>("vimin3_u16x2 + hmax2", 2);
    run<11>("imad", 1);
    run<12>("vminu2 + imad", 2);
    return 0;
}
