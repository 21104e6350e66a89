// Developer probe: issue rate of tcgen05.mma M=128 N=256 K=32 from shared memory, kind::i8 against kind::f8f6f4 (e4m3) and kind::f16,
// one CTA per SM, back-to-back MMAs on two accumulator stages.  Prints cycles per MMA instruction.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
constexpr int M = 128, N = 256, KS = 9, K = 32 * KS, LBO = 128, SBO = (K / 16) * 128;
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr) {
    return (uint64_t)((saddr >> 4) & 0x3fff) | ((uint64_t)(LBO >> 4) << 16) | ((uint64_t)(SBO >> 4) << 32) | ((uint64_t)1 << 46);
}
template <int KIND>
__global__ void __launch_bounds__(128) k_rate(long long* out, int iters) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ __align__(8) uint64_t bar;
    __shared__ uint32_t tmemBase;
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int i = tid; i < (M + N) * K / 4; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0x01010101u * (i & 1);
    if (tid == 0) { asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar))); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&tmemBase)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tm = tmemBase;
    if (tid == 0) {
        // i8: D s32, A u8, B s8.  f8f6f4: D f32 (1 << 4), A/B e4m3 (0).  f16: D f32, A/B f16 (0), K = 16 per instruction (same 32 bytes)
        const uint32_t idesc = (KIND == 0 ? ((2u << 4) | (1u << 10)) : (1u << 4)) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
        const long long t0 = clock64();
        for (int it = 0; it < iters; ++it) {
            for (int j = 0; j < KS; ++j) {
                const uint64_t da = make_desc(smem_u32(smem) + j * 2 * LBO), db = make_desc(smem_u32(smem + M * K) + j * 2 * LBO);
                const uint32_t acc = j > 0;
                const uint32_t d = tm + (uint32_t)((it & 1) * N);
                if (KIND == 0)
                    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}\n" :: "r"(d), "l"(da), "l"(db), "r"(idesc), "r"(acc) : "memory");
                else if (KIND == 1)
                    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f8f6f4 [%0], %1, %2, %3, p;\n\t}\n" :: "r"(d), "l"(da), "l"(db), "r"(idesc), "r"(acc) : "memory");
                else
                    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}\n" :: "r"(d), "l"(da), "l"(db), "r"(idesc), "r"(acc) : "memory");
            }
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
        uint32_t done = 0;
        while (!done) asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }" : "=r"(done) : "r"(smem_u32(&bar)), "r"(0) : "memory");
        if (blockIdx.x == 0) out[KIND] = clock64() - t0;
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tm) : "memory");
}
int main() {
    long long* d; cudaMalloc(&d, 64); cudaMemset(d, 0, 64);
    const size_t smem = (size_t)(M + N) * K + 1024;
    const int iters = 2000;
    cudaFuncSetAttribute(k_rate<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaFuncSetAttribute(k_rate<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaFuncSetAttribute(k_rate<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    for (int ctas : {1, 148}) {
        k_rate<0><<<ctas, 128, smem>>>(d, iters); k_rate<1><<<ctas, 128, smem>>>(d, iters); k_rate<2><<<ctas, 128, smem>>>(d, iters);
        printf("sync: %s\n", cudaGetErrorString(cudaDeviceSynchronize()));
        long long h[3]; cudaMemcpy(h, d, 24, cudaMemcpyDeviceToHost);
        const double n = (double)iters * KS;
        printf("%3d CTAs: cycles per MMA (M128 N256, 32 operand bytes of K): i8 %.1f  f8f6f4 %.1f  f16 %.1f\n", ctas, h[0] / n, h[1] / n, h[2] / n);
    }
    return 0;
}
