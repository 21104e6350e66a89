// Brute-force Hamming k-NN on the 5th-generation tensor cores (A/B beside k_knn, BASELINE config 5).
//
// Hamming(q, r) = popc(q) + popc(r) - 2 <q, r> over the 256 descriptor bits, and <q, r> is a dot product of 0/1 vectors: a GEMM
// with K = 256 that the integer tensor cores compute exactly (tcgen05.mma kind::i8, u8 x s8 -> s32 in TMEM).  The bits are
// expanded to bytes ON CHIP (the database stays 32 bytes per row in HBM): a warp-specialised CTA
//   warps 0-15  epilogue   warp = (TMEM lane quarter, 64-column slice); thread = one query (TMEM lane = accumulator row) against 64
//                          rows of every tile: tcgen05.ld, then the k best (distance, row) keys in registers - the same sorted
//                          insertion as k_knn, no cross-thread step; the four slice lists of a query are merged by k_knn_merge
//   warps 16-23 expansion  thread = one database row of the tile: 32 B of bits -> 288 B of operand bytes in shared memory, in the
//                          K-major no-swizzle canonical layout ((8,n),2):((1,SBO),LBO) the tensor core reads through descriptors
//   warp  24    issue      one thread issues nine tcgen05.mma (M = 128 queries, N = 256 rows, K = 32 each) per tile and commits
//                          them to the mbarriers that free the operand stage and publish the accumulator stage
// with two operand stages and two accumulator stages (2 x 256 TMEM columns), so expansion, MMA and epilogue of consecutive tiles
// overlap.  To keep the epilogue at ~1 instruction per evaluation the operands carry the popcounts: A = 2 * bits, plus 32
// columns of ones; B = bits, plus 32 columns that sum to -popc(row).  The accumulator is then D' = 2 <q, r> - popc(r), the
// distance is popc(q) - D', and "can this row enter the list" is one signed compare of D' against a per-thread threshold.
// Result: bit-identical to k_knn (same keys, same (distance, row) tie rule; per-chunk lists merged by k_knn_merge).
#pragma once
#include "orbm_kernels.cuh"

namespace orbm {

constexpr int kTcM = 128;                        // queries per CTA (TMEM lanes)
constexpr int kTcN = 256;                        // database rows per tile (TMEM columns per accumulator stage)
constexpr int kTcKSteps = 9;                     // 8 x 32 bit-bytes + 1 x 32 popcount bytes
constexpr int kTcK = 32 * kTcKSteps;             // 288 operand bytes per row
constexpr int kTcLBO = 128;                      // bytes between core matrices adjacent in K (8 rows x 16 B each)
constexpr int kTcSBO = (kTcK / 16) * 128;        // bytes between 8-row groups
constexpr int kTcSlices = 4;                      // column slices of a tile: the TMEM -> register path moves ~7 B/clk per warp, 64 B/clk per
constexpr int kTcSliceCols = kTcN / kTcSlices;    // SM, so the accumulator is drained by 16 warps (4 lane quarters x 4 slices), not 4
constexpr int kTcEpiWarps = 4 * kTcSlices, kTcExpWarps = 8, kTcIssueWarp = kTcEpiWarps + kTcExpWarps;
constexpr int kTcThreads = (kTcIssueWarp + 1) * 32;
constexpr size_t kTcSmemA = (size_t)kTcM * kTcK, kTcSmemB = (size_t)kTcN * kTcK;
constexpr size_t kTcSmemBytes = kTcSmemA + 2 * kTcSmemB + 1024;

__device__ __forceinline__ uint64_t tc_desc(uint32_t saddr) {      // UMMA shared-memory descriptor, version 1, no swizzle
    return (uint64_t)((saddr >> 4) & 0x3fff) | ((uint64_t)(kTcLBO >> 4) << 16) | ((uint64_t)(kTcSBO >> 4) << 32) | ((uint64_t)1 << 46);
}
// kSleep: back off between polls - 24 warps spinning on their barriers cost the SM a quarter of its issue slots (ncu: 6400
// warp-instructions per tile of which ~1700 were polls)
template <bool kSleep = true>
__device__ __forceinline__ void tc_wait(uint32_t bar, uint32_t parity) {
    uint32_t done = 0;
    for (;;) {
        asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                     : "=r"(done) : "r"(bar), "r"(parity) : "memory");
        if (done) break;
        if (kSleep) __nanosleep(40);
    }
}
// four bits -> four bytes (bit i in byte i): the four shifted copies of x do not overlap, so the product has no carries
__device__ __forceinline__ uint32_t tc_nibble(uint32_t x) { return (x * 0x00204081u) & 0x01010101u; }

// one descriptor row (32 B of bits) -> operand bytes at `row` of a canonical tile: 16 chunks of 16 bit-bytes (x scale) and two
// chunks of popcount bytes (`tail` in every byte position sums to what the caller wants)
__device__ __forceinline__ void tc_expand_row(uint8_t* tile, int row, const uint4 lo, const uint4 hi, uint32_t scale, const uint4 tail0, const uint4 tail1) {
    uint8_t* base = tile + (size_t)(row >> 3) * kTcSBO + (size_t)(row & 7) * 16;
    const uint32_t w[8] = {lo.x, lo.y, lo.z, lo.w, hi.x, hi.y, hi.z, hi.w};
#pragma unroll
    for (int i = 0; i < 8; ++i) {
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const uint32_t b = w[i] >> (16 * h);
            uint4 o;
            o.x = tc_nibble(b & 15u) * scale; o.y = tc_nibble((b >> 4) & 15u) * scale;
            o.z = tc_nibble((b >> 8) & 15u) * scale; o.w = tc_nibble((b >> 12) & 15u) * scale;
            *reinterpret_cast<uint4*>(base + (size_t)(2 * i + h) * kTcLBO) = o;
        }
    }
    *reinterpret_cast<uint4*>(base + (size_t)16 * kTcLBO) = tail0;
    *reinterpret_cast<uint4*>(base + (size_t)17 * kTcLBO) = tail1;
}

// 32 consecutive accumulator columns of this thread's TMEM lane
__device__ __forceinline__ void tc_ld32(uint32_t (&v)[32], uint32_t addr) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                 "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                 "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                   "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
                   "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
                   "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
                 : "r"(addr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// grid (ceil(nq / 128), nchunks), block kTcThreads, dynamic smem kTcSmemBytes.  rowsPerChunk is a multiple of kTcN.
// part holds kTcSlices lists per (chunk, query): list index = chunk * kTcSlices + slice.
template <int kK>
__global__ void __launch_bounds__(kTcThreads, 1) k_knn_tc(const uint32_t* __restrict__ qdesc, int nq, const uint32_t* __restrict__ db,
                                                           long long ndb, long long rowOffset, int rowsPerChunk, int k,
                                                           unsigned long long* __restrict__ part, unsigned long long* __restrict__ stats) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ __align__(8) uint64_t bFull[2], bEmpty[2], dFull[2], dEmpty[2];
    __shared__ uint32_t tmemBase;
    uint8_t* sA = smem;
    uint8_t* sB = smem + kTcSmemA;
    const int tid = threadIdx.x, warp = tid >> 5;
    const long long r0 = (long long)blockIdx.y * rowsPerChunk;
    const int nrows = (int)max(0LL, min(ndb, r0 + rowsPerChunk) - r0);
    const int ntiles = (nrows + kTcN - 1) / kTcN;
    const int qBase = blockIdx.x * kTcM;
    const bool doStats = stats && blockIdx.x == 0 && blockIdx.y == 0;

    if (tid == 0) {
        for (int s = 0; s < 2; ++s) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(&bFull[s])), "r"(kTcExpWarps * 32));
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bEmpty[s])));
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&dFull[s])));
            asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(&dEmpty[s])), "r"(kTcEpiWarps * 32));
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == kTcIssueWarp) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&tmemBase)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    if (warp >= kTcEpiWarps && warp < kTcIssueWarp) {     // A: queries of this CTA, 2 * bit in the bit columns, ones in the popcount columns
        const int r = tid - kTcEpiWarps * 32;
        if (r < kTcM) {
            const int q = qBase + r;
            uint4 lo = make_uint4(0, 0, 0, 0), hi = lo;
            if (q < nq) { lo = reinterpret_cast<const uint4*>(qdesc)[2 * q]; hi = reinterpret_cast<const uint4*>(qdesc)[2 * q + 1]; }
            const uint4 ones = make_uint4(0x01010101u, 0x01010101u, 0x01010101u, 0x01010101u);
            tc_expand_row(sA, r, lo, hi, 2u, ones, ones);
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tm = tmemBase;

    if (warp == kTcIssueWarp) {
        // ---- issue: one thread drives the tensor core ----
        if ((tid & 31) == 0) {
            // instruction descriptor: D s32 (2 << 4), A u8 (0 << 7), B s8 (1 << 10), both K-major, N >> 3 at bit 17, M >> 4 at bit 24
            const uint32_t idesc = (2u << 4) | (1u << 10) | ((uint32_t)(kTcN >> 3) << 17) | ((uint32_t)(kTcM >> 4) << 24);
            const uint32_t aAddr = smem_u32(sA);
            for (int t = 0; t < ntiles; ++t) {
                const int s = t & 1;
                const uint32_t u = (uint32_t)(t >> 1);
                const long long c0s = clock64();
                tc_wait<false>(smem_u32(&dEmpty[s]), (u & 1u) ^ 1u);   // the epilogue has drained this accumulator stage
                const long long c1s = clock64();
                tc_wait<false>(smem_u32(&bFull[s]), u & 1u);           // the operand bytes of this tile are in place
                if (doStats) { atomicAdd(stats + 0, (unsigned long long)(c1s - c0s)); atomicAdd(stats + 1, (unsigned long long)(clock64() - c1s)); }
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t bAddr = smem_u32(sB + (size_t)s * kTcSmemB);
#pragma unroll
                for (int j = 0; j < kTcKSteps; ++j) {
                    const uint64_t da = tc_desc(aAddr + j * 2 * kTcLBO), dbd = tc_desc(bAddr + j * 2 * kTcLBO);
                    const uint32_t acc = j > 0;
                    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                                 "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}\n"
                                 :: "r"(tm + (uint32_t)(s * kTcN)), "l"(da), "l"(dbd), "r"(idesc), "r"(acc) : "memory");
                }
                // both arrive when the nine MMAs above have completed (commit implies tcgen05.fence::before_thread_sync)
                asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bEmpty[s])) : "memory");
                asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&dFull[s])) : "memory");
            }
        }
    } else if (warp >= kTcEpiWarps) {
        // ---- expansion: thread e owns row e of every tile ----
        const int e = tid - kTcEpiWarps * 32;
        const uint4* rows = reinterpret_cast<const uint4*>(db);
        for (int t = 0; t < ntiles; ++t) {
            const int s = t & 1;
            const uint32_t u = (uint32_t)(t >> 1);
            const int rl = t * kTcN + e;                                  // load first: it flies while the stage is still busy
            const bool ok = rl < nrows;
            uint4 lo = make_uint4(0, 0, 0, 0), hi = lo;
            if (ok) { lo = __ldg(rows + 2 * (r0 + rl)); hi = __ldg(rows + 2 * (r0 + rl) + 1); }
            const long long e0s = clock64();
            tc_wait(smem_u32(&bEmpty[s]), (u & 1u) ^ 1u);
            const long long e1s = clock64();
            // popcount columns: signed bytes that sum to -popc(row) (0 .. 256): two bytes of at most -128 each, thirty zeros.
            // Rows past the end of the chunk get -256: D' = -256, and the epilogue skips their columns anyway.
            const int pd = ok ? __popc(lo.x) + __popc(lo.y) + __popc(lo.z) + __popc(lo.w) + __popc(hi.x) + __popc(hi.y) + __popc(hi.z) + __popc(hi.w) : 256;
            const int p0 = min(pd, 128), p1 = pd - p0;
            const uint32_t tw0 = ((uint32_t)(-p0) & 0xffu) | (((uint32_t)(-p1) & 0xffu) << 8);
            tc_expand_row(sB + (size_t)s * kTcSmemB, e, lo, hi, 1u, make_uint4(tw0, 0, 0, 0), make_uint4(0, 0, 0, 0));
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // my generic-proxy stores before the tensor core's reads
            asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&bFull[s])) : "memory");
            if (doStats && e == 0) { atomicAdd(stats + 2, (unsigned long long)(e1s - e0s)); atomicAdd(stats + 3, (unsigned long long)(clock64() - e1s)); }
        }
    } else {
        // ---- epilogue: warp = (TMEM lane quarter, column slice); thread = one query against kTcSliceCols rows of every tile ----
        const int quarter = warp & 3, slice = warp >> 2;
        const int q = qBase + quarter * 32 + (tid & 31);
        int pq = 0;
        if (q < nq) {
            const uint4 a = reinterpret_cast<const uint4*>(qdesc)[2 * q], b = reinterpret_cast<const uint4*>(qdesc)[2 * q + 1];
            pq = __popc(a.x) + __popc(a.y) + __popc(a.z) + __popc(a.w) + __popc(b.x) + __popc(b.y) + __popc(b.z) + __popc(b.w);
        }
        uint32_t best[kK];
#pragma unroll
        for (int u = 0; u < kK; ++u) best[u] = 0xffffffffu;
        uint32_t worst = 0xffffffffu;
        // D' >= thr  <=>  distance <= distance of the k-th best; a padding lane (q >= nq) holds a threshold no D' reaches
        int thr = q < nq ? pq - 511 : 0x7fffffff;
        for (int t = 0; t < ntiles; ++t) {
            const int s = t & 1;
            const uint32_t u = (uint32_t)(t >> 1);
            const int valid = min(kTcN, nrows - t * kTcN);
            const long long p0s = clock64();
            tc_wait(smem_u32(&dFull[s]), u & 1u);
            const long long p1s = clock64();
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll 1
            for (int c0 = slice * kTcSliceCols; c0 < (slice + 1) * kTcSliceCols; c0 += 32) {
                uint32_t v[32];
                tc_ld32(v, tm + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(s * kTcN + c0));
                if (c0 + 32 >= (slice + 1) * kTcSliceCols) {       // my last load of this stage is in registers: the stage is free
                    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&dEmpty[s])) : "memory");
                }
                // one compare for 32 evaluations - almost never taken once the list has warmed up
                int m32 = (int)v[0];
#pragma unroll
                for (int i = 1; i < 32; ++i) m32 = max(m32, (int)v[i]);
                if (m32 >= thr) {
#pragma unroll
                    for (int i = 0; i < 32; ++i) {
                        const int col = c0 + i;
                        const int dp = (int)v[i];
                        if (dp >= thr && col < valid) {
                            const uint32_t key = ((uint32_t)(pq - dp) << 23) | (uint32_t)(t * kTcN + col);
                            if (key < worst) {
                                uint32_t carry = key;
#pragma unroll
                                for (int w2 = 0; w2 < kK; ++w2) {
                                    if (w2 < k) { const uint32_t lo2 = min(best[w2], carry), hi2 = max(best[w2], carry); best[w2] = lo2; carry = hi2; worst = lo2; }
                                }
                                thr = pq - (int)(worst >> 23);
                            }
                        }
                    }
                }
            }
            if (doStats && tid == 0) { atomicAdd(stats + 4, (unsigned long long)(p1s - p0s)); atomicAdd(stats + 5, (unsigned long long)(clock64() - p1s)); atomicAdd(stats + 6, 1ull); }
        }
        if (q < nq) {
            unsigned long long* out = part + (((size_t)blockIdx.y * kTcSlices + slice) * nq + q) * k;
#pragma unroll
            for (int u = 0; u < kK; ++u) {
                if (u < k) {
                    const uint32_t key = best[u];
                    out[u] = key == 0xffffffffu ? 0xffffffffffffffffull
                                                : (((unsigned long long)(key >> 23)) << 32) | (unsigned long long)(rowOffset + r0 + (key & 0x7fffffu));
                }
            }
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == kTcIssueWarp) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tm) : "memory");
}

}  // namespace orbm
