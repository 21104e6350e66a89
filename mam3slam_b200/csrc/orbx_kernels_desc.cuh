// Quadtree driver (E3/E4), output assembly (E8), orientation (E5) and rotated BRIEF (E7).
// Reference: src/ORBextractor.cc:874-895 (post-octree fix-up), :76-103 (IC_Angle), :106-146 (computeOrbDescriptor),
// :1086-1168 (operator() assembly, lapping order).
#pragma once
#include "orbx_kernels_image.cuh"
#include "orbx_kernels_tree.cuh"

namespace orbx {

__constant__ int c_umax[16];          // circular patch half-widths (src/ORBextractor.cc:453-468)
__constant__ char4 c_pattern[256];    // rBRIEF tests (x0,y0,x1,y1), include/orb_pattern_31.inc

__host__ __device__ inline size_t align16(size_t v) { return (v + 15) & ~(size_t)15; }

// shared-memory carve-up of the tree state; the host uses the same function to size the launch
__host__ __device__ inline size_t tree_smem_layout(int P, int scanCap, TreeSmem* S, uint8_t* base) {
    size_t o = 0;
    auto take = [&](size_t bytes) { size_t at = o; o += align16(bytes); return at; };
    size_t r0 = take((size_t)P * 8), r1 = take((size_t)P * 8);
    size_t b0 = take((size_t)P * 4), b1 = take((size_t)P * 4);
    size_t c0 = take((size_t)P * 4), c1 = take((size_t)P * 4);
    size_t pe = take((size_t)P * 16), cc = take((size_t)P * 16);
    size_t sa = take((size_t)scanCap * 4), sb = take((size_t)scanCap * 4);
    size_t pr = take((size_t)P * 2), dv = take((size_t)P);
    if (S) {
        S->rect[0] = (ushort4*)(base + r0); S->rect[1] = (ushort4*)(base + r1);
        S->beg[0] = (uint32_t*)(base + b0); S->beg[1] = (uint32_t*)(base + b1);
        S->cnt[0] = (uint32_t*)(base + c0); S->cnt[1] = (uint32_t*)(base + c1);
        S->pend = (ise::Elem*)(base + pe); S->cc = (uint4*)(base + cc);
        S->scanA = (int*)(base + sa); S->scanB = (int*)(base + sb);
        S->proc = (uint16_t*)(base + pr); S->dv = base + dv;
        S->scanCap = scanCap;
    }
    return o;
}

// One CTA per (level, frame): gather the level's FAST candidates in the reference's order (cells row-major,
// raster inside a cell), cull with the quadtree, leave the kept keys in list order.
__global__ void __launch_bounds__(kTreeMaxThreads) k_octree(const DevGeom* __restrict__ g, const FastCell* __restrict__ cells,
                                                         const uint32_t* __restrict__ cand, const int* __restrict__ cellCount,
                                                         uint64_t* qtA, uint64_t* qtB, uint64_t* lvlKp, int* lvlCnt,
                                                         int P, int scanCap, int frameBase, int pyrRoots) {
    extern __shared__ __align__(16) uint8_t smem[];
    __shared__ int sW[33];
    TreeSmem S;
    const size_t treeBytes = tree_smem_layout(P, scanCap, &S, smem);
    const int l = blockIdx.y, frame = frameBase + blockIdx.x;      // level-major launch order: the heavy fine levels start first
    const LevelGeom& L = g->lv[l];
    const int tid = threadIdx.x, wid = tid >> 5, lane = tid & 31;
    const int cb = L.cellBegin, nc = L.cellEnd - L.cellBegin;
    const int* cc = cellCount + (size_t)frame * g->nCells + cb;
    TPROF_T0();
    // count pyramid (orbx_kernels_tree.cuh): filled while the candidates are gathered, when the level's geometry fits its words
    PyrView PV{nullptr, nullptr};
    const bool pyr = pyrRoots > 0 && L.nIni <= pyrRoots && L.candCap < (1 << 20);
    if (pyr) {
        PV.best = reinterpret_cast<uint32_t*>(smem + align16(treeBytes));
        PV.cnt = PV.best + (size_t)pyrRoots * kPyrCellsPerRoot;
        for (int i = tid; i < L.nIni * kPyrCellsPerRoot; i += kTreeThreads) { PV.best[i] = 0; PV.cnt[i] = 0; }
    }
    for (int i = tid; i < nc; i += kTreeThreads) S.scanA[i] = cc[i];
    __syncthreads();
    const int n = block_exscan(S.scanA, nc, sW);
    uint64_t* bufA = qtA + (size_t)frame * g->qtCapFrame + L.qtOffset;
    uint64_t* bufB = qtB + (size_t)frame * g->qtCapFrame + L.qtOffset;
    const uint32_t* cf = cand + (size_t)frame * g->candCapFrame;
    // A thread per candidate of the level (position i of the concatenated cell lists = its place in the key array).  A warp
    // takes a contiguous stretch of positions, its lanes interleaved, so a lane's positions only grow: its cell is found by
    // bisection in the offsets just scanned ONCE and then advanced (cells hold ~30 candidates: about one step per candidate).
    // Four candidates per thread at a time with all their loads issued before anything is stored: a warp per cell walked 17
    // (512 threads) to 68 (128 threads) cells in sequence, each a chain of two L2 round trips that the shared-memory atomics
    // of the insert kept from overlapping.
    if (tid == 0) S.scanA[nc] = n;                                 // sentinel (block_exscan left the exclusive offsets in scanA[0..nc))
    __syncthreads();
    constexpr int kGatherBatch = 4;
    const int perWarp = (((n + kTreeWarps - 1) / kTreeWarps) + 31) & ~31;
    const int wBegin = wid * perWarp, wEnd = min(wBegin + perWarp, n);
    int lo = 0;
    if (wBegin + lane < wEnd) {
        const int i = wBegin + lane;
        int hi = nc - 1;
        while (lo < hi) {                                          // last cell whose offset is <= i (empty cells share the next one's)
            const int mid = (lo + hi + 1) >> 1;
            if (S.scanA[mid] <= i) lo = mid; else hi = mid - 1;
        }
    }
    for (int i0 = wBegin + lane; i0 < wEnd; i0 += kGatherBatch * 32) {
        uint32_t e[kGatherBatch];
        int rx[kGatherBatch], ry[kGatherBatch];
#pragma unroll
        for (int j = 0; j < kGatherBatch; ++j) {
            const int i = i0 + j * 32;
            e[j] = 0; rx[j] = 0; ry[j] = 0;
            if (i < wEnd) {
                while (i >= S.scanA[lo + 1]) ++lo;
                const FastCell& cell = cells[cb + lo];
                rx[j] = cell.relX; ry[j] = cell.relY;
                e[j] = cf[cell.outOffset + (i - S.scanA[lo])];
            }
        }
#pragma unroll
        for (int j = 0; j < kGatherBatch; ++j) {
            const int i = i0 + j * 32;
            if (i < wEnd) {
                const uint32_t x = (e[j] & 255u) + rx[j], y = ((e[j] >> 8) & 255u) + ry[j], sc = e[j] >> 16;
                bufA[i] = (uint64_t)x | ((uint64_t)y << 16) | ((uint64_t)sc << 32);
                if (pyr) pyr_insert(PV, x, y, sc, (uint32_t)i, L.hX, L.nIni, L.treeH);
            }
        }
    }
    __syncthreads();
    TPROF(0);
    uint64_t* const buf[2] = {bufA, bufB};
    uint64_t* out = lvlKp + (size_t)frame * g->kpCapFrame + L.kpOffset;
    int* outCount = lvlCnt + frame * g->nlevels + l;
    if (pyr) {
        pyr_reduce(PV, L.nIni);
        if (distribute_tree<true>(buf, n, L.quota, L.nIni, L.hX, L.treeH, S, out, outCount, PV)) return;
        __syncthreads();                                           // a node deeper than the pyramid: the partitioning form takes over
    }
    distribute_tree<false>(buf, n, L.quota, L.nIni, L.hX, L.treeH, S, out, outCount);
}

// standalone tree for tests: keys (x | y<<16 | score<<32) already in bufA
// usePyr: try the count-pyramid form first (outCount[1] = 1 when it completed, 0 when the partitioning form had to take over)
__global__ void __launch_bounds__(kTreeMaxThreads) k_debug_distribute(uint64_t* bufA, uint64_t* bufB, int n, int N, int nIni,
                                                                   float hX, int treeH, uint64_t* out, int* outCount,
                                                                   int P, int scanCap, int usePyr) {
    extern __shared__ __align__(16) uint8_t smem[];
    TreeSmem S;
    const size_t treeBytes = tree_smem_layout(P, scanCap, &S, smem);
    uint64_t* const buf[2] = {bufA, bufB};
    if (threadIdx.x == 0) outCount[1] = 0;
    if (usePyr && nIni <= kPyrMaxRoots && n < (1 << 20)) {
        PyrView PV;
        PV.best = reinterpret_cast<uint32_t*>(smem + align16(treeBytes));
        PV.cnt = PV.best + (size_t)kPyrMaxRoots * kPyrCellsPerRoot;
        for (int i = threadIdx.x; i < nIni * kPyrCellsPerRoot; i += kTreeThreads) { PV.best[i] = 0; PV.cnt[i] = 0; }
        __syncthreads();
        for (int i = threadIdx.x; i < n; i += kTreeThreads) {
            const uint64_t k = bufA[i];
            pyr_insert(PV, (uint32_t)(k & 0xffff), (uint32_t)((k >> 16) & 0xffff), (uint32_t)(k >> 32), (uint32_t)i, hX, nIni, treeH);
        }
        pyr_reduce(PV, nIni);
        if (distribute_tree<true>(buf, n, N, nIni, hX, treeH, S, out, outCount, PV)) { if (threadIdx.x == 0) outCount[1] = 1; return; }
        __syncthreads();
    }
    distribute_tree<false>(buf, n, N, nIni, hX, treeH, S, out, outCount);
}

// mode 0: single-thread std::sort emulation, 1: heap sort alone, 2: the warp version the quadtree uses (scratch: 3n+8 ints)
__global__ void k_debug_sort(ise::Elem* e, int n, int mode, uint32_t* scratch) {
    if (mode == 2) {
        warp_sort_nodes(e, n, scratch, scratch + n + 4, scratch + 2 * n + 8);
    } else if (threadIdx.x == 0 && blockIdx.x == 0) {
        if (mode == 1) ise::heap_sort(e, 0, n);
        else ise::sort(e, n);
    }
}

struct WorkItem { short level; short pad; int x, y; int dst; };   // keypoint in level coordinates + output slot

// E4 + E8: one CTA per frame.  Walk the kept lists level by level (the reference's order), add the border offset,
// scale to level-0 coordinates, and place every keypoint where operator() would (:1143-1164): points inside
// [lap0, lap1] fill the output from the BACK in reverse, the others from the front; monoIndex = number at the front.
constexpr int kFinalThreads = 256;
__global__ void __launch_bounds__(kFinalThreads) k_finalize(const DevGeom* __restrict__ g, const uint64_t* __restrict__ lvlKp,
                                                            const int* __restrict__ lvlCnt, int lap0, int lap1,
                                                            orbx_keypoint* outKp, int* outCnt, WorkItem* work, int frameBase) {
    __shared__ int sOff[kMaxLevels + 1];
    __shared__ int sWarp[kFinalThreads / 32 + 1];
    const int frame = frameBase + blockIdx.x, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int L = g->nlevels;
    if (tid == 0) {
        int o = 0;
        for (int l = 0; l < L; ++l) { sOff[l] = o; o += lvlCnt[frame * L + l]; }
        sOff[L] = o;
    }
    __syncthreads();
    const int n = min(sOff[L], g->outCap);
    const int per = (n + kFinalThreads - 1) / kFinalThreads;
    const int b = min(tid * per, n), e = min(b + per, n);
    const float fl0 = (float)lap0, fl1 = (float)lap1;
    const uint64_t* kf = lvlKp + (size_t)frame * g->kpCapFrame;
    // pass 1: count lapping keypoints in my chunk
    int myLap = 0;
    int l = 0;
    for (int o = b; o < e; ++o) {
        while (o >= sOff[l + 1]) ++l;
        const uint64_t key = kf[g->lv[l].kpOffset + (o - sOff[l])];
        float x = (float)((int)(key & 0xffff) + kMinBorder);
        if (l != 0) x = __fmul_rn(x, g->lv[l].scale);
        myLap += (x >= fl0 && x <= fl1) ? 1 : 0;
    }
    int incl = myLap;
#pragma unroll
    for (int s = 1; s < 32; s <<= 1) { int t = __shfl_up_sync(0xffffffffu, incl, s); if (lane >= s) incl += t; }
    if (lane == 31) sWarp[wid] = incl;
    __syncthreads();
    if (wid == 0) {
        int v = lane < kFinalThreads / 32 ? sWarp[lane] : 0, iv = v;
#pragma unroll
        for (int s = 1; s < 32; s <<= 1) { int t = __shfl_up_sync(0xffffffffu, iv, s); if (lane >= s) iv += t; }
        if (lane < kFinalThreads / 32) sWarp[lane] = iv - v;
        if (lane == 31) sWarp[kFinalThreads / 32] = iv;
    }
    __syncthreads();
    int lapBefore = sWarp[wid] + incl - myLap;
    const int lapTotal = sWarp[kFinalThreads / 32];
    orbx_keypoint* ok = outKp + (size_t)frame * g->outCap;
    WorkItem* wk = work + (size_t)frame * g->outCap;
    l = 0;
    for (int o = b; o < e; ++o) {
        while (o >= sOff[l + 1]) ++l;
        const LevelGeom& G = g->lv[l];
        const uint64_t key = kf[G.kpOffset + (o - sOff[l])];
        const int xi = (int)(key & 0xffff) + kMinBorder, yi = (int)((key >> 16) & 0xffff) + kMinBorder;
        float x = (float)xi, y = (float)yi;
        if (l != 0) { x = __fmul_rn(x, G.scale); y = __fmul_rn(y, G.scale); }
        const bool lap = (x >= fl0 && x <= fl1);
        const int dst = lap ? (n - 1 - lapBefore) : (o - lapBefore);
        lapBefore += lap ? 1 : 0;
        orbx_keypoint k;
        k.x = x; k.y = y; k.size = G.sizeF; k.angle = -1.f; k.response = (float)(int)((key >> 32) & 0xff);
        k.octave = l; k.class_id = -1;
        ok[dst] = k;
        WorkItem w; w.level = (short)l; w.pad = 0; w.x = xi; w.y = yi; w.dst = dst;
        wk[o] = w;
    }
    if (tid == 0) { outCnt[2 * frame] = n; outCnt[2 * frame + 1] = n - lapTotal; }
}

// cv::fastAtan2 scalar path, evaluated without FMA contraction so that it matches a plain float CPU build.
__device__ __forceinline__ float fast_atan2_deg(float y, float x) {
    constexpr float p1 = 0.9997878412794807f * (float)(180 / 3.14159265358979323846);
    constexpr float p3 = -0.3258083974640975f * (float)(180 / 3.14159265358979323846);
    constexpr float p5 = 0.1555786518463281f * (float)(180 / 3.14159265358979323846);
    constexpr float p7 = -0.04432655554792128f * (float)(180 / 3.14159265358979323846);
    constexpr float eps = (float)2.2204460492503131e-16;
    const float ax = fabsf(x), ay = fabsf(y);
    float a;
    if (ax >= ay) {
        const float c = __fdiv_rn(ay, __fadd_rn(ax, eps));
        const float c2 = __fmul_rn(c, c);
        a = __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c);
    } else {
        const float c = __fdiv_rn(ax, __fadd_rn(ay, eps));
        const float c2 = __fmul_rn(c, c);
        a = __fsub_rn(90.f, __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c));
    }
    if (x < 0) a = __fsub_rn(180.f, a);
    if (y < 0) a = __fsub_rn(360.f, a);
    return a;
}

// E5 + E7: one warp per GROUP of kDescGroup keypoints.
//   1. orientation of each keypoint in turn: lanes = columns u = -15..15 of the radius-15 disc on the UNBLURRED level, int32
//      moments, shuffle reduction; lane j keeps the moments of keypoint j.
//   2. the first kDescGroup lanes finish their own keypoint: fastAtan2, then cos / sin in double narrowed to float - one
//      instruction stream for the whole group.
//   3. descriptor of each keypoint in turn: lane i produces byte i (8 of the 256 tests) on the BLURRED level; the lane's 8
//      tests live in registers as floats for the whole group (no per-keypoint table reads or int->float conversions);
//      pattern points rotated in float with separately rounded products (no FMA), cvRound = round-half-even.
#ifndef ORBX_DESC_GROUP
#define ORBX_DESC_GROUP 2      /* 4 -> 2 (with 2 warps per CTA in the TMA kernel): +1.4 % per 64-frame step, a third of the shared memory per CTA; 1 and 8 are slower */
#endif
constexpr int kDescWarps = 8, kDescGroup = ORBX_DESC_GROUP;
__global__ void __launch_bounds__(kDescWarps * 32) k_orient_brief(const DevGeom* __restrict__ g, Level0 l0,
                                                                  const WorkItem* __restrict__ work,
                                                                  const int* __restrict__ outCnt, orbx_keypoint* outKp,
                                                                  uint8_t* outDesc, int frameBase) {
    const int frame = frameBase + blockIdx.y;
    const int o0 = (blockIdx.x * kDescWarps + (threadIdx.x >> 5)) * kDescGroup;
    const int lane = threadIdx.x & 31;
    const int total = outCnt[2 * frame];
    if (o0 >= total) return;
    const int ng = min(kDescGroup, total - o0);
    // lane j < ng holds keypoint j of the group
    WorkItem mine; mine.level = 0; mine.pad = 0; mine.x = 0; mine.y = 0; mine.dst = 0;
    if (lane < ng) mine = work[(size_t)frame * g->outCap + o0 + lane];
    // this lane's 8 tests
    float px0[8], py0[8], px1[8], py1[8];
#pragma unroll
    for (int bit = 0; bit < 8; ++bit) {
        const char4 t = c_pattern[lane * 8 + bit];
        px0[bit] = (float)t.x; py0[bit] = (float)t.y; px1[bit] = (float)t.z; py1[bit] = (float)t.w;
    }
    const int u = lane - kHalfPatch;
    const int vm = lane < 31 ? c_umax[abs(u)] : -1;          // lane 31 holds no column
    int myM10 = 0, myM01 = 0;
    for (int j = 0; j < ng; ++j) {
        const int lev = __shfl_sync(0xffffffffu, (int)mine.level, j);
        const int kx = __shfl_sync(0xffffffffu, mine.x, j), ky = __shfl_sync(0xffffffffu, mine.y, j);
        int pitch;
        const uint8_t* img = level_ptr(g, l0, lev, frame, &pitch);
        const uint8_t* ctr = img + (size_t)ky * pitch + kx;
        {
            // the descriptor phase will sample the blurred level in a 37x37 window whose position is known now but whose
            // sample offsets depend on the angle: pull the window towards L1 / L2 while the orientation is being computed
            const LevelGeom& G = g->lv[lev];
            const uint8_t* bw = g->blur[lev] + (size_t)frame * G.planeBytes + (size_t)(ky - 18) * G.pitch + (kx - 18);
            const uint8_t* r0p = bw + (size_t)lane * G.pitch;
            asm volatile("prefetch.global.L1 [%0];" ::"l"(r0p));
            asm volatile("prefetch.global.L1 [%0];" ::"l"(r0p + 36));
            if (lane < 5) {
                const uint8_t* r1p = bw + (size_t)(32 + lane) * G.pitch;
                asm volatile("prefetch.global.L1 [%0];" ::"l"(r1p));
                asm volatile("prefetch.global.L1 [%0];" ::"l"(r1p + 36));
            }
        }
        int m10 = 0, m01 = 0;
        // the disc mask is symmetric in (u, v) (:453-468 makes it so): column u spans the rows |v| <= umax[|u|].  Fully
        // unrolled so that the 31 byte loads of a lane are in flight together.
#pragma unroll
        for (int v = -kHalfPatch; v <= kHalfPatch; ++v) {
            if ((v < 0 ? -v : v) <= vm) {
                const int val = ctr[v * pitch + u];
                m10 += u * val;
                m01 += v * val;
            }
        }
#pragma unroll
        for (int s = 16; s > 0; s >>= 1) {
            m10 += __shfl_xor_sync(0xffffffffu, m10, s);
            m01 += __shfl_xor_sync(0xffffffffu, m01, s);
        }
        if (lane == j) { myM10 = m10; myM01 = m01; }
    }
    float ca = 0.f, sb = 0.f;
    if (lane < ng) {
        const float angle = fast_atan2_deg((float)myM01, (float)myM10);
        outKp[(size_t)frame * g->outCap + mine.dst].angle = angle;
        constexpr float factorPI = (float)(3.14159265358979323846 / 180.f);
        const double rad = (double)__fmul_rn(angle, factorPI);
        double sd, cd;
        sincos(rad, &sd, &cd);
        ca = (float)cd; sb = (float)sd;
    }
    for (int j = 0; j < ng; ++j) {
        const float a = __shfl_sync(0xffffffffu, ca, j), b = __shfl_sync(0xffffffffu, sb, j);
        const int lev = __shfl_sync(0xffffffffu, (int)mine.level, j);
        const int kx = __shfl_sync(0xffffffffu, mine.x, j), ky = __shfl_sync(0xffffffffu, mine.y, j);
        const int dst = __shfl_sync(0xffffffffu, mine.dst, j);
        const LevelGeom& G = g->lv[lev];
        const int bp = G.pitch;
        const uint8_t* bc = g->blur[lev] + (size_t)frame * G.planeBytes + (size_t)ky * bp + kx;
        int byte = 0;
#pragma unroll
        for (int bit = 0; bit < 8; ++bit) {
            const int r0 = __float2int_rn(__fadd_rn(__fmul_rn(px0[bit], b), __fmul_rn(py0[bit], a)));
            const int c0 = __float2int_rn(__fsub_rn(__fmul_rn(px0[bit], a), __fmul_rn(py0[bit], b)));
            const int r1 = __float2int_rn(__fadd_rn(__fmul_rn(px1[bit], b), __fmul_rn(py1[bit], a)));
            const int c1 = __float2int_rn(__fsub_rn(__fmul_rn(px1[bit], a), __fmul_rn(py1[bit], b)));
            const int v0 = bc[r0 * bp + c0], v1 = bc[r1 * bp + c1];
            byte |= (v0 < v1 ? 1 : 0) << bit;
        }
        outDesc[((size_t)frame * g->outCap + dst) * 32 + lane] = (uint8_t)byte;
    }
}

}  // namespace orbx
