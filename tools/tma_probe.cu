// Developer aid: which way of handing a CUtensorMap to cp.async.bulk.tensor works on this box (run on the GPU).
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <cstring>
#include <vector>
struct alignas(64) Maps { CUtensorMap m[4]; int ok[4]; };
__device__ __forceinline__ uint32_t sa(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ void load(const CUtensorMap* tm, uint8_t* tile, uint64_t* bar, int x, int y, int f, int bytes) {
    const uint32_t b = sa(bar);
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(b));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(b), "r"(bytes) : "memory");
        asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
                     :: "r"(sa(tile)), "l"(reinterpret_cast<uint64_t>(tm)), "r"(b), "r"(x), "r"(y), "r"(f) : "memory");
    }
    __syncwarp();
    uint32_t done = 0;
    while (!done)
        asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }" : "=r"(done) : "r"(b), "r"(0) : "memory");
}
__device__ void sum_out(const uint8_t* tile, uint32_t* out) {
    uint32_t s = 0; for (int i = threadIdx.x; i < 96 * 34; i += 32) s += tile[i];
    for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(~0u, s, o);
    if (threadIdx.x == 0) *out = s;
}
__global__ void k_direct(const __grid_constant__ CUtensorMap tm, uint32_t* out, int x, int y, int f) {
    __shared__ __align__(128) uint8_t tile[96 * 34]; __shared__ __align__(8) uint64_t bar;
    load(&tm, tile, &bar, x, y, f, 96 * 34);
    sum_out(tile, out);
}
__global__ void k_array(const __grid_constant__ Maps maps, int l, uint32_t* out, int x, int y, int f) {
    __shared__ __align__(128) uint8_t tile[96 * 34]; __shared__ __align__(8) uint64_t bar;
    load(&maps.m[l], tile, &bar, x, y, f, 96 * 34);
    sum_out(tile, out);
}
__global__ void k_global(const CUtensorMap* tm, uint32_t* out, int x, int y, int f) {
    __shared__ __align__(128) uint8_t tile[96 * 34]; __shared__ __align__(8) uint64_t bar;
    load(tm, tile, &bar, x, y, f, 96 * 34);
    sum_out(tile, out);
}
typedef CUresult (*Enc)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                        CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
template <int R>
__global__ void k_rank(const __grid_constant__ CUtensorMap tm, uint32_t* out, int x, int y, int f, int bytes, int cta) {
    __shared__ __align__(1024) uint8_t tile[128 * 40]; __shared__ __align__(8) uint64_t bar;
    const uint32_t b = sa(&bar);
    for (int i = threadIdx.x; i < 128 * 40; i += 32) tile[i] = 0;
    __syncwarp();
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(b));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(b), "r"(bytes) : "memory");
        if (R == 2)
            asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                         :: "r"(sa(tile)), "l"(reinterpret_cast<uint64_t>(&tm)), "r"(b), "r"(x), "r"(y) : "memory");
        else
            asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
                         :: "r"(sa(tile)), "l"(reinterpret_cast<uint64_t>(&tm)), "r"(b), "r"(x), "r"(y), "r"(f) : "memory");
    }
    __syncwarp();
    uint32_t done = 0;
    while (!done)
        asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }" : "=r"(done) : "r"(b), "r"(0) : "memory");
    uint32_t s = 0; for (int i = threadIdx.x; i < bytes; i += 32) s += tile[i];
    for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(~0u, s, o);
    if (threadIdx.x == 0) *out = s;
}
int main(int argc, char** argv) {
    const int variant = argc > 1 ? atoi(argv[1]) : 0;
    void* p = nullptr; cudaDriverEntryPointQueryResult q;
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q);
    Enc enc = (Enc)p;
    const int W = 627, H = 400, B = 3, pitch = 640; size_t plane = (size_t)pitch * H;
    uint8_t* d; cudaMalloc(&d, plane * B);
    std::vector<uint8_t> h(plane * B); for (size_t i = 0; i < h.size(); ++i) h[i] = (uint8_t)(i * 7 + (i >> 9));
    cudaMemcpy(d, h.data(), h.size(), cudaMemcpyHostToDevice);
    int rank = 3, bw = 96, bh = 34; CUtensorMapL2promotion l2 = CU_TENSOR_MAP_L2_PROMOTION_L2_128B;
    int x = 12, y = 16, f = 1;
    if (variant == 1) rank = 2;
    if (variant == 2) { bw = 128; bh = 32; }
    if (variant == 3) { bw = 64; bh = 34; }
    if (variant == 4) l2 = CU_TENSOR_MAP_L2_PROMOTION_NONE;
    if (variant == 5) { rank = 2; bw = 128; bh = 32; l2 = CU_TENSOR_MAP_L2_PROMOTION_NONE; x = 0; y = 0; }
    if (variant == 6) { x = 16; }                 // 16-byte aligned start column
    if (variant == 7) { rank = 2; x = 16; }
    CUtensorMap tm; cuuint64_t dims[3] = {(cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)B}, str[2] = {(cuuint64_t)pitch, plane};
    cuuint32_t box[3] = {(cuuint32_t)bw, (cuuint32_t)bh, 1}, es[3] = {1, 1, 1};
    CUresult r = enc(&tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, rank, d, dims, str, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                     l2, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (rank == 2) f = 0;
    uint32_t ref = 0; for (int yy = 0; yy < bh; ++yy) for (int xx = 0; xx < bw; ++xx) if (x + xx < W && y + yy < H) ref += h[f * plane + (size_t)(y + yy) * pitch + x + xx];
    uint32_t* out; cudaMalloc(&out, 4); uint32_t got = 0;
    if (rank == 2) k_rank<2><<<1, 32>>>(tm, out, x, y, f, bw * bh, 0); else k_rank<3><<<1, 32>>>(tm, out, x, y, f, bw * bh, 0);
    cudaError_t e = cudaDeviceSynchronize(); cudaMemcpy(&got, out, 4, cudaMemcpyDeviceToHost);
    printf("variant %d rank %d box %dx%d x=%d encode=%d: %s got %u ref %u\n", variant, rank, bw, bh, x, (int)r, cudaGetErrorString(e), got, ref);
    return 0;
}
