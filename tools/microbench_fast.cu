// Developer aid (run on the GPU): which pipes can carry the FAST-9/16 min/max network on sm_100a.
// Pixel values 0..255 in u16 lanes are ALSO valid fp16 bit patterns (subnormals, ordered like the integers), so half2 min/max
// (HMNMX2) computes the same lattice operations as VIMNMX.U16x2 - on a different execution pipe.
// build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o tools/microbench_fast tools/microbench_fast.cu
#include <cstdio>
#include <cstdint>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <vector>

__device__ __forceinline__ uint32_t hmin(uint32_t a, uint32_t b) {
    __half2 r = __hmin2(*reinterpret_cast<__half2*>(&a), *reinterpret_cast<__half2*>(&b));
    return *reinterpret_cast<uint32_t*>(&r);
}
__device__ __forceinline__ uint32_t hmax(uint32_t a, uint32_t b) {
    __half2 r = __hmax2(*reinterpret_cast<__half2*>(&a), *reinterpret_cast<__half2*>(&b));
    return *reinterpret_cast<uint32_t*>(&r);
}
struct IntOps {
    static __device__ __forceinline__ uint32_t mn(uint32_t a, uint32_t b) { return __vminu2(a, b); }
    static __device__ __forceinline__ uint32_t mx(uint32_t a, uint32_t b) { return __vmaxu2(a, b); }
    static __device__ __forceinline__ uint32_t mn3(uint32_t a, uint32_t b, uint32_t c) { return __vimin3_u16x2(a, b, c); }
    static __device__ __forceinline__ uint32_t mx3(uint32_t a, uint32_t b, uint32_t c) { return __vimax3_u16x2(a, b, c); }
};
struct HalfOps {
    static __device__ __forceinline__ uint32_t mn(uint32_t a, uint32_t b) { return hmin(a, b); }
    static __device__ __forceinline__ uint32_t mx(uint32_t a, uint32_t b) { return hmax(a, b); }
    static __device__ __forceinline__ uint32_t mn3(uint32_t a, uint32_t b, uint32_t c) { return hmin(hmin(a, b), c); }
    static __device__ __forceinline__ uint32_t mx3(uint32_t a, uint32_t b, uint32_t c) { return hmax(hmax(a, b), c); }
};
struct Mix2Ops {            // 2-input on the half pipe, 3-input on the integer pipe
    static __device__ __forceinline__ uint32_t mn(uint32_t a, uint32_t b) { return hmin(a, b); }
    static __device__ __forceinline__ uint32_t mx(uint32_t a, uint32_t b) { return hmax(a, b); }
    static __device__ __forceinline__ uint32_t mn3(uint32_t a, uint32_t b, uint32_t c) { return __vimin3_u16x2(a, b, c); }
    static __device__ __forceinline__ uint32_t mx3(uint32_t a, uint32_t b, uint32_t c) { return __vimax3_u16x2(a, b, c); }
};

// bright side: max over the 16 arcs of min(arc); dark side: min over arcs of max(arc)  (see fast_strength_x2)
template <class B, class D>
__device__ __forceinline__ uint32_t strength(uint32_t v2, const uint32_t (&r)[16]) {
    uint32_t q[8], Q[8], qq[8], QQ[8], a[8], b[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) { q[k] = B::mn(r[2 * k + 1], r[(2 * k + 2) & 15]); Q[k] = D::mx(r[2 * k + 1], r[(2 * k + 2) & 15]); }
#pragma unroll
    for (int k = 0; k < 8; ++k) { qq[k] = B::mn(q[k], q[(k + 1) & 7]); QQ[k] = D::mx(Q[k], Q[(k + 1) & 7]); }
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const uint32_t e = B::mx(r[2 * k], r[(2 * k + 9) & 15]), E = D::mn(r[2 * k], r[(2 * k + 9) & 15]);
        a[k] = B::mn3(qq[k], qq[(k + 2) & 7], e);
        b[k] = D::mx3(QQ[k], QQ[(k + 2) & 7], E);
    }
    uint32_t A = B::mx3(a[0], a[1], a[2]), Bm = D::mn3(b[0], b[1], b[2]);
    A = B::mx3(A, a[3], a[4]); Bm = D::mn3(Bm, b[3], b[4]);
    A = B::mx3(A, a[5], a[6]); Bm = D::mn3(Bm, b[5], b[6]);
    A = B::mx(A, a[7]); Bm = D::mn(Bm, b[7]);
    return __vimax3_u16x2(A + (0x01000100u - v2), (v2 + 0x01000100u) - Bm, 0x01000100u);
}

constexpr int ITERS = 512;
template <class B, class D>
__global__ void __launch_bounds__(128) bench(uint32_t* out, uint32_t seed, long long* cyc) {
    uint32_t r[16], v = ((seed * (threadIdx.x + 3)) >> 3) & 0x00ff00ffu;
#pragma unroll
    for (int i = 0; i < 16; ++i) r[i] = ((seed + i * 7919u) * (threadIdx.x * 2654435761u + blockIdx.x + 1) >> 5) & 0x00ff00ffu;
    uint32_t acc = 0;
    const long long t0 = clock64();
    for (int it = 0; it < ITERS; ++it) {
        const uint32_t s = strength<B, D>(v, r);
        acc += s;
        // rotate the ring and perturb it with the result so that nothing is loop-invariant
        const uint32_t t = r[0];
#pragma unroll
        for (int i = 0; i < 15; ++i) r[i] = r[i + 1];
        r[15] = (t + s) & 0x00ff00ffu;
        v = (v + 0x00010001u) & 0x00ff00ffu;
    }
    const long long t1 = clock64();
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

template <class B, class D> uint64_t run(const char* name) {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    const int blocks = p.multiProcessorCount * 8;
    uint32_t* out; long long* cyc;
    cudaMalloc(&out, (size_t)blocks * 128 * 4); cudaMalloc(&cyc, 8);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    bench<B, D><<<blocks, 128>>>(out, 12345u, cyc);
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    for (int i = 0; i < 5; ++i) bench<B, D><<<blocks, 128>>>(out, 12345u, cyc);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    std::vector<uint32_t> h((size_t)blocks * 128);
    cudaMemcpy(h.data(), out, h.size() * 4, cudaMemcpyDeviceToHost);
    uint64_t sum = 0; for (uint32_t x : h) sum += x;
    const double pairs = 5.0 * blocks * 128.0 * ITERS;
    printf("%-44s %8.1f G pixel-pairs/s  %.3f ms  checksum %llu\n", name, pairs / (ms * 1e-3) / 1e9, ms / 5, (unsigned long long)sum);
    cudaFree(out); cudaFree(cyc);
    return sum;
}

#include <vector>
int main() {
    const uint64_t a = run<IntOps, IntOps>("int / int (VIMNMX + VIMNMX3)");
    const uint64_t b = run<HalfOps, HalfOps>("half / half (HMNMX2 only)");
    const uint64_t c = run<IntOps, HalfOps>("bright int, dark half");
    const uint64_t d = run<Mix2Ops, Mix2Ops>("2-input half, 3-input int");
    const uint64_t e = run<Mix2Ops, IntOps>("bright mixed, dark int");
    const uint64_t f = run<Mix2Ops, HalfOps>("bright mixed, dark half");
    printf("identical results: %s\n", (a == b && a == c && a == d && a == e && a == f) ? "yes" : "NO");
    return 0;
}
