// Developer probe: tcgen05.mma kind::f8f6f4 (e4m3 x e4m3) with F16 accumulators, M=128 N=256 K=288: where do the 16-bit results land in
// TMEM (packed two per 32-bit column?) and are small-integer sums exact.  Dumps raw columns and checks candidate layouts.
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <vector>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
constexpr int M = 128, N = 256, KS = 9, K = 32 * KS, LBO = 128, SBO = (K / 16) * 128;
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr) {
    return (uint64_t)((saddr >> 4) & 0x3fff) | ((uint64_t)(LBO >> 4) << 16) | ((uint64_t)(SBO >> 4) << 32) | ((uint64_t)1 << 46);
}
__global__ void __launch_bounds__(128) k_probe(const uint8_t* __restrict__ A, const uint8_t* __restrict__ B, uint32_t* __restrict__ D) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t* sA = smem; uint8_t* sB = smem + M * K;
    __shared__ __align__(8) uint64_t bar;
    __shared__ uint32_t tmemBase;
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int i = tid; i < M * K; i += 128) { const int r = i / K, kb = i % K; sA[(r / 8) * SBO + (kb / 16) * LBO + (r % 8) * 16 + kb % 16] = A[i]; }
    for (int i = tid; i < N * K; i += 128) { const int r = i / K, kb = i % K; sB[(r / 8) * SBO + (kb / 16) * LBO + (r % 8) * 16 + kb % 16] = B[i]; }
    if (tid == 0) { asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar))); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 256;" ::"r"(smem_u32(&tmemBase)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tm = tmemBase;
    const uint32_t idesc = (0u << 4) | (0u << 7) | (0u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);   // D f16, A/B e4m3
    if (tid == 0) {
        for (int j = 0; j < KS; ++j) {
            const uint64_t da = make_desc(smem_u32(sA) + j * 2 * LBO), db = make_desc(smem_u32(sB) + j * 2 * LBO);
            const uint32_t acc = j > 0;
            asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f8f6f4 [%0], %1, %2, %3, p;\n\t}\n"
                         :: "r"(tm), "l"(da), "l"(db), "r"(idesc), "r"(acc) : "memory");
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
    }
    { uint32_t done = 0; while (!done) asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }" : "=r"(done) : "r"(smem_u32(&bar)), "r"(0) : "memory"); }
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    for (int c0 = 0; c0 < 256; c0 += 16) {
        uint32_t v[16];
        const uint32_t addr = tm + ((uint32_t)(warp * 32) << 16) + (uint32_t)c0;
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                     : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                       "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]) : "r"(addr));
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        for (int i = 0; i < 16; ++i) D[tid * 256 + c0 + i] = v[i];
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 256;" ::"r"(tm) : "memory");
}
static uint8_t e4m3(int v) {        // small integers -8 .. 8
    static const uint8_t pos[9] = {0x00, 0x38, 0x40, 0x44, 0x48, 0x4a, 0x4c, 0x4e, 0x50};
    return v >= 0 ? pos[v] : (uint8_t)(pos[-v] | 0x80);
}
int main() {
    std::vector<int> Ai(M * K), Bi(N * K);
    std::vector<uint8_t> A(M * K), B(N * K);
    srand(2);
    for (int i = 0; i < M * K; ++i) { const int kb = i % K; Ai[i] = kb < 256 ? 2 * (rand() & 1) : 1; A[i] = e4m3(Ai[i]); }
    for (int i = 0; i < N * K; ++i) { const int kb = i % K; Bi[i] = kb < 256 ? (rand() & 1) : -(rand() % 9); B[i] = e4m3(Bi[i]); }
    uint8_t *dA, *dB; uint32_t* dD;
    cudaMalloc(&dA, A.size()); cudaMalloc(&dB, B.size()); cudaMalloc(&dD, M * 256 * 4);
    cudaMemcpy(dA, A.data(), A.size(), cudaMemcpyHostToDevice); cudaMemcpy(dB, B.data(), B.size(), cudaMemcpyHostToDevice);
    cudaMemset(dD, 0xee, M * 256 * 4);
    const size_t smem = (size_t)(M + N) * K + 1024;
    cudaFuncSetAttribute(k_probe, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    k_probe<<<1, 128, smem>>>(dA, dB, dD);
    printf("kernel: %s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    std::vector<uint32_t> D(M * 256);
    cudaMemcpy(D.data(), dD, D.size() * 4, cudaMemcpyDeviceToHost);
    std::vector<int> ref(M * N);
    for (int m = 0; m < M; ++m) for (int n = 0; n < N; ++n) { int s = 0; for (int k = 0; k < K; ++k) s += Ai[m * K + k] * Bi[n * K + k]; ref[m * N + n] = s; }
    auto h = [](uint16_t b) { __half_raw r; r.x = b; return (int)__half2float(__half(r)); };
    printf("row 0 raw cols 0..7: "); for (int c = 0; c < 8; ++c) printf("%08x ", D[c]); printf("\nref row 0 n 0..15: "); for (int n = 0; n < 16; ++n) printf("%d ", ref[n]); printf("\n");
    long badPacked = 0, badLow = 0, badSplit = 0;
    for (int m = 0; m < M; ++m) for (int n = 0; n < N; ++n) {
        const uint32_t w = D[m * 256 + n / 2]; const int got = h((uint16_t)(n & 1 ? w >> 16 : w & 0xffff));
        if (got != ref[m * N + n]) ++badPacked;                                            // layout 1: column n/2, half n&1
        if (h((uint16_t)(D[m * 256 + n] & 0xffff)) != ref[m * N + n]) ++badLow;           // layout 2: column n, low half
        const uint32_t w3 = D[m * 256 + (n % 128)]; if (h((uint16_t)(n >= 128 ? w3 >> 16 : w3 & 0xffff)) != ref[m * N + n]) ++badSplit;   // layout 3: halves = n / 128
    }
    printf("mismatches: packed (col n/2, half n&1) %ld | unpacked low half %ld | split halves %ld of %d\n", badPacked, badLow, badSplit, M * N);
    return 0;
}
