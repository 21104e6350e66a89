// Integer-pipe microbenchmarks for sm_100a (B200): register-resident dependent chains, 8 independent chains per thread,
// full occupancy.  Prints giga thread-ops/s and ops/clk/SM (at the clock measured with clock64 + wall time).
// Used to set the INT roofline for Hamming matching (popc) and to choose the FAST formulation (SIMD min/max).
// build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tools/microbench tools/microbench.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

constexpr int ITERS = 4096, CH = 8;

template <int OP> __device__ __forceinline__ uint32_t op(uint32_t a, uint32_t b, uint32_t c) {
    if (OP == 0) return __popc(a ^ b) + c;                    // LOP3 + POPC + IADD
    if (OP == 1) return (a ^ b) + c;                          // LOP3 + IADD (baseline for 0)
    if (OP == 2) return __popc(a) ;                           // pure POPC chain
    if (OP == 3) return min(min(a, b), c);                    // VIMNMX3 u32 ?
    if (OP == 4) return __vimin3_u16x2(a, b, c);              // DPX u16x2 3-input min
    if (OP == 5) return __vminu2(a, b);                       // u16x2 2-input min
    if (OP == 6) return __byte_perm(a, b, c & 0x7777);        // PRMT
    if (OP == 7) return a * b + c;                            // IMAD
    if (OP == 8) return __vminu4(a, b);                       // u8x4 min (emulated?)
    if (OP == 9) return __vimax3_s16x2(a, b, c);
    if (OP == 10) return __vsubus4(a, b);                     // saturating u8x4 sub
    if (OP == 11) return __vcmpgtu4(a, b);
    if (OP == 12) return __vmaxu4(a, b);
    if (OP == 13) return a + b + c;                           // IADD3
    if (OP == 14) return __vabsdiffu4(a, b);
    if (OP == 15) return __vadd2(a, b);
    if (OP == 16) return __dp4a(a, b, c);                     // IDP.4A u8
    if (OP == 17) return __dp2a_lo(a, b, c);                  // IDP.2A u16 x u8
    if (OP == 18) return __funnelshift_r(a, b, c);            // SHF
    if (OP == 19) return (a & b) ^ c;                         // LOP3
    if (OP == 20) return __shfl_xor_sync(0xffffffffu, a, 1) + c;   // SHFL + IADD
    if (OP == 21) return __vavgu4(a, b);
    if (OP == 22) return __vadd4(a, b);
    return 0;
}

template <int OP> __global__ void __launch_bounds__(256) bench(uint32_t* out, uint32_t seed, long long* cyc) {
    uint32_t v[CH], b = seed * (threadIdx.x + 1), c = seed + blockIdx.x;
#pragma unroll
    for (int i = 0; i < CH; ++i) v[i] = seed + i * 77 + threadIdx.x;
    long long t0 = clock64();
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < CH; ++i) v[i] = op<OP>(v[i], b, c);
        b += 0x01010101u;
    }
    long long t1 = clock64();
    uint32_t s = 0;
#pragma unroll
    for (int i = 0; i < CH; ++i) s ^= v[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

template <int OP> void run(const char* name, int opsPerCall) {
    int dev = 0; cudaDeviceProp p; cudaGetDeviceProperties(&p, dev);
    const int blocks = p.multiProcessorCount * 8;
    uint32_t* out; long long* cyc;
    cudaMalloc(&out, (size_t)blocks * 256 * 4); cudaMalloc(&cyc, 8);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    bench<OP><<<blocks, 256>>>(out, 12345u, cyc);
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    for (int r = 0; r < 5; ++r) bench<OP><<<blocks, 256>>>(out, 12345u + r, cyc);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    long long hc; cudaMemcpy(&hc, cyc, 8, cudaMemcpyDeviceToHost);
    const double calls = 5.0 * blocks * 256.0 * ITERS * CH;
    const double gops = calls / (ms * 1e-3) / 1e9;
    // per-SM per-clock from the in-kernel cycle count of one block (8 blocks/SM resident => divide accordingly)
    const double perClkSM = (double)8 * 256 * ITERS * CH / (double)hc;
    printf("%-28s %9.1f Gcalls/s  (%d hw-op/call)  %6.1f calls/clk/SM (clock64)  %.3f ms\n", name, gops, opsPerCall, perClkSM, ms / 5);
    cudaFree(out); cudaFree(cyc);
}

int main() {
    run<0>("xor+popc+add", 3);
    run<1>("xor+add", 2);
    run<2>("popc chain", 1);
    run<3>("min3 u32", 1);
    run<4>("vimin3_u16x2", 1);
    run<5>("vminu2 (u16x2)", 1);
    run<6>("prmt", 1);
    run<7>("imad", 1);
    run<8>("vminu4 (u8x4)", 1);
    run<9>("vimax3_s16x2", 1);
    run<10>("vsubus4", 1);
    run<11>("vcmpgtu4", 1);
    run<12>("vmaxu4", 1);
    run<13>("iadd3", 1);
    run<14>("vabsdiffu4", 1);
    run<15>("vadd2", 1);
    run<16>("dp4a u8", 1);
    run<17>("dp2a_lo u16xu8", 1);
    run<18>("funnelshift", 1);
    run<19>("lop3", 1);
    run<20>("shfl_xor+add", 2);
    run<21>("vavgu4", 1);
    run<22>("vadd4", 1);
    return 0;
}
