// Developer probe: one CTA, tcgen05.mma kind::i8 (u8 x s8 -> s32), M=128 N=256, K = 32*KS, operands in the K-major no-swizzle
// canonical layout ((8,n),2):((1,SBO),LBO) [units of 16 B], accumulator read back with tcgen05.ld.32x32b.  Checks the
// descriptor encodings the k-NN tensor-core kernel relies on against a CPU loop.
// nvcc -gencode arch=compute_100a,code=sm_100a -o tc_probe tools/tc_probe.cu && ./tc_probe
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cstdint>
#include <vector>
#include <cuda_runtime.h>

constexpr int M = 128, N = 256, KS = 9, K = 32 * KS;
constexpr int LBO = 128, SBO = (K / 16) * 128;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr >> 4) & 0x3fff);
    d |= (uint64_t)((lbo >> 4) & 0x3fff) << 16;
    d |= (uint64_t)((sbo >> 4) & 0x3fff) << 32;
    d |= (uint64_t)1 << 46;                      // version = 1 (Blackwell)
    return d;                                    // base_offset 0, lbo_mode 0, layout_type 0 (no swizzle)
}

__global__ void __launch_bounds__(128) k_probe(const uint8_t* __restrict__ A, const int8_t* __restrict__ B, int32_t* __restrict__ D) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t* sA = smem;                              // M x K
    uint8_t* sB = smem + M * K;                      // N x K
    __shared__ __align__(8) uint64_t bar;
    __shared__ uint32_t tmemBase;
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int i = tid; i < M * K; i += 128) { const int r = i / K, kb = i % K; sA[(r / 8) * SBO + (kb / 16) * LBO + (r % 8) * 16 + kb % 16] = A[i]; }
    for (int i = tid; i < N * K; i += 128) { const int r = i / K, kb = i % K; sB[(r / 8) * SBO + (kb / 16) * LBO + (r % 8) * 16 + kb % 16] = (uint8_t)B[i]; }
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 256;" ::"r"(smem_u32(&tmemBase)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");       // generic-proxy smem writes -> visible to the tensor core
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tm = tmemBase;
    // instruction descriptor: c_format S32 (2) << 4 | a_format u8 (0) << 7 | b_format s8 (1) << 10 | K-major both | N>>3 << 17 | M>>4 << 24
    const uint32_t idesc = (2u << 4) | (0u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
    if (tid == 0) {
        for (int j = 0; j < KS; ++j) {
            const uint64_t da = make_desc(smem_u32(sA) + j * 2 * LBO, LBO, SBO);
            const uint64_t db = make_desc(smem_u32(sB) + j * 2 * LBO, LBO, SBO);
            const uint32_t acc = j > 0;
            asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                         "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}\n"
                         :: "r"(tm), "l"(da), "l"(db), "r"(idesc), "r"(acc) : "memory");
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
    }
    {
        uint32_t done = 0;
        while (!done)
            asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                         : "=r"(done) : "r"(smem_u32(&bar)), "r"(0) : "memory");
    }
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    // thread = accumulator row (TMEM lane); warp w reads lanes 32w .. 32w+31
    for (int c0 = 0; c0 < N; c0 += 16) {
        uint32_t v[16];
        const uint32_t addr = tm + ((uint32_t)(warp * 32) << 16) + (uint32_t)c0;
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                     : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                       "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
                     : "r"(addr));
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        for (int i = 0; i < 16; ++i) D[tid * N + c0 + i] = (int32_t)v[i];
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 256;" ::"r"(tm) : "memory");
}

int main() {
    std::vector<uint8_t> A(M * K); std::vector<int8_t> B(N * K);
    srand(1);
    for (auto& a : A) a = (uint8_t)(rand() % 3);                 // 0, 1, 2
    for (auto& b : B) b = (int8_t)(rand() % 11 - 8);             // -8 .. 2
    uint8_t* dA; int8_t* dB; int32_t* dD;
    cudaMalloc(&dA, A.size()); cudaMalloc(&dB, B.size()); cudaMalloc(&dD, M * N * 4);
    cudaMemcpy(dA, A.data(), A.size(), cudaMemcpyHostToDevice); cudaMemcpy(dB, B.data(), B.size(), cudaMemcpyHostToDevice);
    cudaMemset(dD, 0xff, M * N * 4);
    const size_t smem = (size_t)(M + N) * K + 1024;
    cudaFuncSetAttribute(k_probe, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    k_probe<<<1, 128, smem>>>(dA, dB, dD);
    cudaError_t e = cudaDeviceSynchronize();
    printf("kernel: %s\n", cudaGetErrorString(e));
    std::vector<int32_t> D(M * N);
    cudaMemcpy(D.data(), dD, D.size() * 4, cudaMemcpyDeviceToHost);
    long bad = 0; int shown = 0;
    for (int m = 0; m < M; ++m)
        for (int n = 0; n < N; ++n) {
            int ref = 0;
            for (int k = 0; k < K; ++k) ref += (int)A[m * K + k] * (int)B[n * K + k];
            if (ref != D[m * N + n]) { ++bad; if (shown++ < 8) printf("  D[%d][%d] = %d, expected %d\n", m, n, D[m * N + n], ref); }
        }
    printf("mismatches: %ld of %d\n", bad, M * N);
    return bad != 0;
}
