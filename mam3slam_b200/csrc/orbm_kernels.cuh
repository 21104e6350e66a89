// Hamming-matching kernels of the B200-native ORBmatcher (include/orbm.h).
//
// k_grid_*        Frame::AssignFeaturesToGrid (src/Frame.cc:385-416) as a stable counting sort -> CSR grid.
// k_search        every guided search of src/ORBmatcher.cc (M1-M7, M9 of SURVEY.md 8a) behind one engine:
//                   candidate source   : 64x48 frame grid window (Frame/KeyFrame::GetFeaturesInArea order) | BoW node run
//                   per-candidate test : level range, square window, chi2 gate, eligibility, epipolar predicate
//                   distance           : 256-bit Hamming, 8 x (xor, popc) on the integer pipe
//                   reduction          : first minimum (+ second best and their levels) | last minimum
//                   acceptance         : TH_HIGH / TH_LOW / ORBdist, NN-ratio variants of the reference
//                   state              : once-only slot occupancy | "steal if strictly better" (SearchForInitialization)
//                 The reference loops are SEQUENTIAL over queries (an accepted match occupies its keypoint for all later
//                 queries).  Here every warp gathers candidates and computes distances for its own query in parallel
//                 (the expensive part), parks them in a shared-memory slab, and then the queries are RESOLVED in query
//                 order by passing a token through shared memory: the resolve step only reads the slab, tests the
//                 occupancy bytes and does two warp reductions.  Searches without inter-query state (triangulation,
//                 fuse) skip the token and spread over many CTAs.
// k_rot_filter    rotation histogram + ComputeThreeMaxima (src/ORBmatcher.cc:2012-2053) epilogue.
// k_knn / k_knn_merge   brute-force k-NN (BASELINE config 5): thread = query, database rows broadcast from shared memory
//                 tiles filled with TMA bulk copies; 256-bit popcount through a carry-save adder tree (4 POPC instead of
//                 8: POPC issues at 16/clk/SM on B200, LOP3 at 128/clk/SM — tools/microbench.cu).
#pragma once
#include <climits>
#ifdef ORBM_STATS
#include <cstdio>
#endif
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/orbm.h"

namespace orbm {

constexpr int kGridCols = 64, kGridRows = 48, kGridCells = kGridCols * kGridRows;

struct FrameDev {
    int n;
    const float4* kp4;        // x, y, angle, octave (int bits)
    const uint32_t* desc;     // n x 8
    const int* cellOff;       // kGridCells + 1
    const int* cellIdx;       // n
    float minX, minY, invW, invH;
};

// ---------------------------------------------------------------------------------------------------------------------
// grid build: cell of every keypoint (PosInGrid uses round(), half away from zero), then a stable counting sort
__global__ void k_grid_prepare(int n, const orbx_keypoint* __restrict__ kps, float4* __restrict__ kp4, int* __restrict__ cellOf,
                               float minX, float minY, float invW, float invH) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const orbx_keypoint k = kps[i];
    kp4[i] = make_float4(k.x, k.y, k.angle, __int_as_float(k.octave));
    const int px = (int)roundf(__fmul_rn(__fsub_rn(k.x, minX), invW));
    const int py = (int)roundf(__fmul_rn(__fsub_rn(k.y, minY), invH));
    cellOf[i] = (px < 0 || px >= kGridCols || py < 0 || py >= kGridRows) ? -1 : px * kGridRows + py;
}

// one CTA: histogram, exclusive scan over the 3072 cells, then a STABLE scatter (ascending keypoint index inside a cell,
// the reference's push_back order).  n is a few thousand; the per-cell order is produced by letting each thread own a
// contiguous range of cells and scan the keypoints of those cells in index order via a per-cell linked pass.
constexpr int kGridThreads = 1024, kGridTmp = 4096;
__device__ __forceinline__ void grid_build_cta(int n, const int* __restrict__ cellOf, int* __restrict__ cellOff,
                                               int* __restrict__ cellIdx) {
    __shared__ int cnt[kGridCells];
    __shared__ int warpSum[32];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    for (int c = tid; c < kGridCells; c += kGridThreads) cnt[c] = 0;
    __syncthreads();
    for (int i = tid; i < n; i += kGridThreads) { const int c = cellOf[i]; if (c >= 0) atomicAdd(&cnt[c], 1); }
    __syncthreads();
    // exclusive scan of cnt[0..3072): 3 cells per thread
    constexpr int per = kGridCells / kGridThreads;
    int v[per], s = 0;
#pragma unroll
    for (int k = 0; k < per; ++k) { v[k] = cnt[tid * per + k]; s += v[k]; }
    int incl = s;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { int t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
    if (lane == 31) warpSum[wid] = incl;
    __syncthreads();
    if (wid == 0) {
        int w = warpSum[lane], iw = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { int t = __shfl_up_sync(0xffffffffu, iw, o); if (lane >= o) iw += t; }
        warpSum[lane] = iw - w;
    }
    __syncthreads();
    int run = warpSum[wid] + incl - s;
#pragma unroll
    for (int k = 0; k < per; ++k) { cnt[tid * per + k] = run; cellOff[tid * per + k] = run; run += v[k]; }
    if (tid == kGridThreads - 1) cellOff[kGridCells] = run;
    __syncthreads();
    // stable scatter.  Usual case (n <= kGridTmp): every keypoint takes any free slot of its cell's segment (shared-memory
    // atomic), then finds its place as the number of smaller indices in that segment - cells hold a handful of keypoints, so
    // this is two short parallel passes instead of one warp walking all keypoints in order.
    if (n <= kGridTmp) {
        __shared__ int tmp[kGridTmp];
        for (int i = tid; i < n; i += kGridThreads) { const int c = cellOf[i]; if (c >= 0) tmp[atomicAdd(&cnt[c], 1)] = i; }
        __syncthreads();                              // cnt[c] is now the END of cell c's segment = the begin of cell c + 1
        for (int i = tid; i < n; i += kGridThreads) {
            const int c = cellOf[i];
            if (c < 0) continue;
            const int b = c ? cnt[c - 1] : 0, e = cnt[c];
            int rank = 0;
            for (int p = b; p < e; ++p) rank += tmp[p] < i;
            cellIdx[b + rank] = i;
        }
        return;
    }
    // large frames: warp 0 walks keypoints in index order in chunks of 32; within a chunk, a keypoint's rank inside its
    // cell = number of lower lanes with the same cell (match_any)
    if (wid == 0) {
        for (int base = 0; base < n; base += 32) {
            const int i = base + lane;
            const int c = i < n ? cellOf[i] : -1;
            const unsigned peers = __match_any_sync(0xffffffffu, c);
            if (c >= 0) {
                const int rank = __popc(peers & ((1u << lane) - 1));
                const int leader = __ffs(peers) - 1;
                int basePos = 0;
                if (lane == leader) { basePos = cnt[c]; cnt[c] = basePos + __popc(peers); }
                basePos = __shfl_sync(peers, basePos, leader);
                cellIdx[basePos + rank] = i;
            }
            __syncwarp();
        }
    }
}

__global__ void __launch_bounds__(kGridThreads) k_grid_build(int n, const int* __restrict__ cellOf, int* __restrict__ cellOff,
                                                             int* __restrict__ cellIdx) {
    grid_build_cta(n, cellOf, cellOff, cellIdx);
}

// k_grid_prepare + k_grid_build of one frame in one launch (one CTA: the frames of single searches hold a few thousand keypoints)
__global__ void __launch_bounds__(kGridThreads) k_frame_grid(int n, const orbx_keypoint* __restrict__ kps, float4* __restrict__ kp4,
                                                             int* __restrict__ cellOf, int* __restrict__ cellOff, int* __restrict__ cellIdx,
                                                             float minX, float minY, float invW, float invH) {
    for (int i = threadIdx.x; i < n; i += kGridThreads) {
        const orbx_keypoint k = kps[i];
        kp4[i] = make_float4(k.x, k.y, k.angle, __int_as_float(k.octave));
        const int px = (int)roundf(__fmul_rn(__fsub_rn(k.x, minX), invW));
        const int py = (int)roundf(__fmul_rn(__fsub_rn(k.y, minY), invH));
        cellOf[i] = (px < 0 || px >= kGridCols || py < 0 || py >= kGridRows) ? -1 : px * kGridRows + py;
    }
    __syncthreads();
    grid_build_cta(n, cellOf, cellOff, cellIdx);
}

// ---------------------------------------------------------------------------------------------------------------------
enum Rule : int {
    RULE_LOCAL_MAP = 0,   // M1: best <= TH_HIGH, reject iff bestLevel == bestLevel2 && best > nnratio * best2
    RULE_BEST_LE = 1,     // M2/M3/M7: (float)best <= thF
    RULE_RATIO_LE = 2,    // M4a: best <= thI && (float)best < nnratio * (float)best2
    RULE_RATIO_LT = 3,    // M4b: best <  thI && (float)best < nnratio * (float)best2
    RULE_INIT = 4,        // M5: best <= thI && best < (float)best2 * nnratio
    RULE_LAST_MIN = 5,    // M6: any surviving candidate; the LAST minimum wins
    RULE_NONE = 6         // M9: report best index / distance, no acceptance
};

struct SearchParams {
    FrameDev tgt;
    int nq;
    int source;                  // 0 = grid window, 1 = BoW run
    const float4* qwin;          // window: u, v, r, -          (M6: kp1.x, kp1.y, -, -)
    const int2* qlev;            // window: lo, hi octave bounds (inclusive)
    const int2* qrun;            // BoW: [begin, end) into bowIdx
    const int* bowIdx;
    const uint32_t* qdesc;       // nq x 8
    const float* qangle;
    const uint8_t* elig;         // target-side eligibility (may be NULL)
    uint8_t* occ;                // target-side occupancy bytes, in/out (NULL when state == 0)
    const uint8_t* qocc;         // state 1: query q's accepted match blocks its keypoint for later queries iff qocc[q] (the
                                 // assigned MapPoint has Observations() > 0, src/ORBmatcher.cc:92-94, 1745-1747); NULL = always
    int state;                   // 0 = none, 1 = occupancy (claim on accept), 2 = steal-if-better (M5)
    int stageTgt;                // 1 = the launch has room to stage the target frame (grid, kp4, descriptors) in shared memory
    int rule;
    float thF; int thI; float nnratio; int none2;
    int checkOri;
    int chi2; const float* invSigma2;
    int tri; int coarse; float ep0, ep1; float F[9]; const float* scale2; const float* sigma2;
    int* qResult;                // accepted target index or -1
    int* qDist;                  // best distance
    int* qBin;                   // rotation bin of the accept event or -1
    unsigned long long* evals;
    // node-parallel BoW searches (k_search_nodes): queries [nodeQ[j].x, nodeQ[j].y) belong to shared node j and all of them
    // scan the same run; the query descriptor is row qsrc[q] of qdescBase (the source frame's descriptors in HBM)
    const int2* nodeQ; int nNodes;
    const int* qsrc; const uint32_t* qdescBase;
    unsigned int* ticket;        // CTAs-done counter for the fused rotation epilogue (zeroed by the host)
};

#ifndef ORBM_SEARCH_THREADS
#define ORBM_SEARCH_THREADS 256      /* 512 measured 2.8 % slower per 64-frame step: a 512 x 122-register CTA holds an SM's whole register file for the 107 us of the in-order commit; 128: 0.22 ms tracking, no gain */
#endif
constexpr int kSearchThreads = ORBM_SEARCH_THREADS, kSearchWarps = kSearchThreads / 32, kSlab = 192, kSlots = 64;

// dynamic shared memory of a search CTA: slabs + match state (+ the staged target frame); host and device agree on this
// window: grid-window searches need only the 32-batch ring of Top4 lists (24 KB), run searches the 64 candidate slabs (48 KB)
constexpr size_t kSearchRingBytes = 32 * 32 * (16 + 4 + 4);
__host__ __device__ inline size_t search_slab_bytes(bool window) { return window ? kSearchRingBytes : (size_t)kSlots * kSlab * 4; }
__host__ __device__ inline size_t search_smem_bytes(int n, int state, bool stage, bool window) {
    size_t b = search_slab_bytes(window);
    if (state == 1) b += ((size_t)n + 15) & ~(size_t)15;
    if (state == 2) b += (((size_t)n * 2 + 15) & ~(size_t)15) + (((size_t)n * 4 + 15) & ~(size_t)15);
    if (stage) b = ((b + 15) & ~(size_t)15) + (size_t)n * 32 + (size_t)n * 16 + (size_t)(kGridCells + 1) * 4 + (size_t)n * 4 + 16;
    return b;
}
constexpr size_t kSearchSmemMax = 200 * 1024;

__device__ __forceinline__ int hamming8(const uint32_t (&q)[8], const uint4 a, const uint4 b) {
    return __popc(q[0] ^ a.x) + __popc(q[1] ^ a.y) + __popc(q[2] ^ a.z) + __popc(q[3] ^ a.w) +
           __popc(q[4] ^ b.x) + __popc(q[5] ^ b.y) + __popc(q[6] ^ b.z) + __popc(q[7] ^ b.w);
}

// entry packing: dist (9 bits) << 23 | octave (4 bits) << 19 | idx (19 bits)
__device__ __forceinline__ uint32_t pack_entry(int dist, int oct, int idx) { return ((uint32_t)dist << 23) | ((uint32_t)(oct & 15) << 19) | (uint32_t)idx; }

// Walks the candidates of query q in the reference's order; calls consume(active, entry) warp-synchronously for every
// chunk of 32 candidate slots (active = this lane holds a candidate that passed all per-candidate tests).
template <class Consume>
__device__ __forceinline__ void walk_candidates(const SearchParams& P, const FrameDev& T, int q, const uint32_t (&qd)[8], unsigned long long& nEval,
                                                Consume&& consume) {
    const int lane = threadIdx.x & 31;
    const uint4* drows = reinterpret_cast<const uint4*>(T.desc);
    if (P.source == 0) {
        const float4 w = P.qwin[q];
        const int2 lv = P.qlev[q];
        const float x = w.x, y = w.y, r = w.z;
        // Frame::GetFeaturesInArea (src/Frame.cc:665-687): cell range, early outs
        const int minCX = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(x, T.minX), r), T.invW)));
        if (minCX >= kGridCols) return;
        const int maxCX = min(kGridCols - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(x, T.minX), r), T.invW)));
        if (maxCX < 0) return;
        const int minCY = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(y, T.minY), r), T.invH)));
        if (minCY >= kGridRows) return;
        const int maxCY = min(kGridRows - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(y, T.minY), r), T.invH)));
        if (maxCY < 0) return;
        if (minCY > maxCY) return;
        // The reference visits the cells column by column (ix outer, iy inner) and, inside a cell, in insertion order; with
        // the column-major CSR grid the candidates of one column are ONE contiguous run.  Lane c fetches the run of column
        // minCX + c, a warp prefix sum flattens the runs, and the lanes then enumerate the flattened sequence (same order),
        // so the dependent-load chain (offsets -> index -> keypoint -> descriptor) is walked once per 32 candidates
        // instead of once per column.
        const int ncol = maxCX - minCX + 1;
        for (int c0 = 0; c0 < ncol; c0 += 32) {
            int rb = 0, len = 0;
            if (c0 + lane < ncol) {
                const int ix = minCX + c0 + lane;
                rb = T.cellOff[ix * kGridRows + minCY];
                len = T.cellOff[ix * kGridRows + maxCY + 1] - rb;
            }
            int incl = len;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
            const int total = __shfl_sync(0xffffffffu, incl, 31);
            for (int base = 0; base < total; base += 32) {
                const int p = base + lane;
                // column of position p = first lane whose inclusive sum exceeds p
                int j = 0;
#pragma unroll
                for (int step = 16; step > 0; step >>= 1) {
                    const int t = __shfl_sync(0xffffffffu, incl, j + step - 1);
                    if (t <= p) j += step;
                }
                j = min(j, 31);
                const int rbj = __shfl_sync(0xffffffffu, rb, j), inj = __shfl_sync(0xffffffffu, incl, j), lnj = __shfl_sync(0xffffffffu, len, j);
                bool ok = false; uint32_t entry = 0;
                if (p < total) {
                    const int e = rbj + (p - (inj - lnj));
                    const int idx = T.cellIdx[e];
                    const float4 k = T.kp4[idx];
                    const int oct = __float_as_int(k.w);
                    ok = oct >= lv.x && oct <= lv.y && fabsf(__fsub_rn(k.x, x)) < r && fabsf(__fsub_rn(k.y, y)) < r;
                    if (ok && P.chi2) {
                        const float ex = __fsub_rn(x, k.x), ey = __fsub_rn(y, k.y);
                        const float e2 = __fadd_rn(__fmul_rn(ex, ex), __fmul_rn(ey, ey));
                        ok = !((double)__fmul_rn(e2, P.invSigma2[oct]) > 5.99);
                    }
                    if (ok && P.elig) ok = P.elig[idx] != 0;
                    if (ok) {
                        const int d = hamming8(qd, drows[2 * idx], drows[2 * idx + 1]);
                        ++nEval;
                        entry = pack_entry(d, oct, idx);
                    }
                }
                consume(ok, entry);
            }
        }
    } else {
        const int2 run = P.qrun[q];
        const float4 w = P.tri ? P.qwin[q] : make_float4(0, 0, 0, 0);
        float la = 0, lb = 0, lc = 0;
        if (P.tri && !P.coarse) {   // epipolar line l = x1' F12 (src/CameraModels/Pinhole.cpp:115-117), no FMA contraction
            la = __fadd_rn(__fadd_rn(__fmul_rn(w.x, P.F[0]), __fmul_rn(w.y, P.F[3])), P.F[6]);
            lb = __fadd_rn(__fadd_rn(__fmul_rn(w.x, P.F[1]), __fmul_rn(w.y, P.F[4])), P.F[7]);
            lc = __fadd_rn(__fadd_rn(__fmul_rn(w.x, P.F[2]), __fmul_rn(w.y, P.F[5])), P.F[8]);
        }
        for (int base = run.x; base < run.y; base += 32) {
            const int e = base + lane;
            bool ok = false; uint32_t entry = 0;
            if (e < run.y) {
                const int idx = P.bowIdx[e];
                ok = !P.elig || P.elig[idx] != 0;
                if (ok) {
                    const int d = hamming8(qd, drows[2 * idx], drows[2 * idx + 1]);
                    ++nEval;
                    const float4 k = T.kp4[idx];
                    const int oct = __float_as_int(k.w);
                    if (P.tri) {
                        // SearchForTriangulation (:1017-1034, :1072): dist <= TH_LOW, epipole distance, epipolar constraint.
                        // (the running "dist > bestDist" test is applied by the LAST_MIN reduction)
                        ok = d <= ORBM_TH_LOW;
                        if (ok) {
                            const float dx = __fsub_rn(P.ep0, k.x), dy = __fsub_rn(P.ep1, k.y);
                            ok = !(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)) < __fmul_rn(100.f, P.scale2[oct]));
                        }
                        if (ok && !P.coarse) {
                            const float num = __fadd_rn(__fadd_rn(__fmul_rn(la, k.x), __fmul_rn(lb, k.y)), lc);
                            const float den = __fadd_rn(__fmul_rn(la, la), __fmul_rn(lb, lb));
                            if (den == 0.f) ok = false;
                            else ok = (double)__fdiv_rn(__fmul_rn(num, num), den) < 3.84 * (double)P.sigma2[oct];
                        }
                    }
                    entry = pack_entry(d, oct, idx);
                }
            }
            consume(ok, entry);
        }
    }
}

struct Best { uint32_t k1, k2; uint32_t e1, e2; };   // lane-private best / second-best keys (dist<<20|pos) and entries

// state tests of one candidate at resolve time
__device__ __forceinline__ bool blocked(const SearchParams& P, const uint8_t* occS, const uint16_t* mdS, uint32_t entry) {
    const int idx = entry & 0x7ffff;
    if (P.state == 1) return occS[idx] != 0;
    if (P.state == 2) return mdS[idx] <= (entry >> 23);           // vMatchedDistance[i2] <= dist (:687)
    return false;
}

__device__ __forceinline__ void best_update(Best& b, uint32_t key, uint32_t entry) {
    if (key < b.k1) { b.k2 = b.k1; b.e2 = b.e1; b.k1 = key; b.e1 = entry; }
    else if (key < b.k2) { b.k2 = key; b.e2 = entry; }
}

constexpr uint32_t kNoEntry = 0xffffffffu, kOverflowEntry = 0xfffffffeu;     // real entries are < 0x80800000 (dist <= 256)

__device__ __forceinline__ bool rule_needs_second(int rule) {
    return rule == RULE_LOCAL_MAP || rule == RULE_RATIO_LE || rule == RULE_RATIO_LT || rule == RULE_INIT;
}

// Best and second-best candidate of query q among those not blocked by the match state.  Warp-collective; every lane gets
// the result (kNoEntry = none).  `slab` holds the candidates in scan order; more than kSlab of them (rare) are re-walked.
__device__ __forceinline__ void scan_candidates(const SearchParams& P, const FrameDev& T, int q, const uint32_t* slab, int count,
                                                const uint8_t* occS, const uint16_t* mdS, uint32_t& e1out, uint32_t& e2out) {
    const int lane = threadIdx.x & 31;
    const bool lastMin = P.rule == RULE_LAST_MIN;
    Best b; b.k1 = b.k2 = 0xffffffffu; b.e1 = b.e2 = kNoEntry;
    if (count <= kSlab) {
        for (int p = lane; p < count; p += 32) {
            const uint32_t e = slab[p];
            if (blocked(P, occS, mdS, e)) continue;
            const uint32_t key = ((e >> 23) << 20) | (uint32_t)(lastMin ? (0xfffff - p) : p);
            best_update(b, key, e);
        }
    } else {
        uint32_t qd[8];
        const uint4 a = reinterpret_cast<const uint4*>(P.qdesc)[2 * q], c = reinterpret_cast<const uint4*>(P.qdesc)[2 * q + 1];
        qd[0] = a.x; qd[1] = a.y; qd[2] = a.z; qd[3] = a.w; qd[4] = c.x; qd[5] = c.y; qd[6] = c.z; qd[7] = c.w;
        int pos0 = 0;
        unsigned long long dummy = 0;
        walk_candidates(P, T, q, qd, dummy, [&](bool ok, uint32_t entry) {
            const unsigned m = __ballot_sync(0xffffffffu, ok);
            if (ok && !blocked(P, occS, mdS, entry)) {
                const int p = pos0 + __popc(m & ((1u << lane) - 1));
                const uint32_t key = ((entry >> 23) << 20) | (uint32_t)(lastMin ? (0xfffff - p) : p);
                best_update(b, key, entry);
            }
            pos0 += __popc(m);
        });
    }
    // warp merge: best = min k1; second = min over {k2 of the winner lane, k1 of the others}
    const uint32_t m1 = __reduce_min_sync(0xffffffffu, b.k1);
    const bool win = (b.k1 == m1) && (m1 != 0xffffffffu);
    const unsigned wmask = __ballot_sync(0xffffffffu, win);
    e1out = wmask ? __shfl_sync(0xffffffffu, b.e1, __ffs(wmask) - 1) : kNoEntry;
    e2out = kNoEntry;
    if (rule_needs_second(P.rule)) {
        const uint32_t cand2 = win ? b.k2 : b.k1;
        const uint32_t m2 = __reduce_min_sync(0xffffffffu, cand2);
        const unsigned smask = __ballot_sync(0xffffffffu, cand2 == m2 && m2 != 0xffffffffu);
        if (smask) e2out = __shfl_sync(0xffffffffu, win ? b.e2 : b.e1, __ffs(smask) - 1);
    }
}

// Acceptance test of the reference's rule on the best / second-best surviving candidates (kNoEntry = none).
__device__ __forceinline__ bool decide_query(const SearchParams& P, uint32_t e1, uint32_t e2) {
    const bool has1 = e1 != kNoEntry, has2 = e2 != kNoEntry;
    const int bestDist = has1 ? (int)(e1 >> 23) : 256;
    const int bestLevel = has1 ? (int)((e1 >> 19) & 15) : -1;
    const int bestDist2 = has2 ? (int)(e2 >> 23) : P.none2;
    const int bestLevel2 = has2 ? (int)((e2 >> 19) & 15) : -1;
    // an if-chain on purpose: a jump table (BRX) in this very large kernel faulted with "illegal instruction" on sm_100a
    const int rule = P.rule;
    if (rule == RULE_BEST_LE) return has1 && (float)bestDist <= P.thF;
    if (rule == RULE_LOCAL_MAP)
        return has1 && bestDist <= ORBM_TH_HIGH &&
               !(bestLevel == bestLevel2 && (float)bestDist > __fmul_rn(P.nnratio, (float)(has2 ? bestDist2 : 256)));
    if (rule == RULE_RATIO_LE) return has1 && bestDist <= P.thI && (float)bestDist < __fmul_rn(P.nnratio, (float)bestDist2);
    if (rule == RULE_RATIO_LT) return has1 && bestDist < P.thI && (float)bestDist < __fmul_rn(P.nnratio, (float)bestDist2);
    if (rule == RULE_INIT) return has1 && bestDist <= P.thI && (float)bestDist < __fmul_rn((float)bestDist2, P.nnratio);
    if (rule == RULE_LAST_MIN) return has1;
    return false;
}

// Match-state update of an accepted query (state 2, M5: the new owner is recorded in m21S; the queries that lost their
// keypoint are cleared after the loop).
__device__ __forceinline__ void update_state(const SearchParams& P, int q, uint32_t e1, uint8_t* occS, uint16_t* mdS, int* m21S) {
    const int bestIdx = (int)(e1 & 0x7ffff);
    if (P.state == 1 && (!P.qocc || P.qocc[q])) occS[bestIdx] = 1;
    if (P.state == 2) { m21S[bestIdx] = q; mdS[bestIdx] = (uint16_t)(e1 >> 23); }      // (:706-713)
}

// Result write of one query: accepted keypoint (or -1), best distance, rotation-histogram bin of the accept event.
__device__ __forceinline__ void write_result(const SearchParams& P, const FrameDev& T, int q, uint32_t e1, bool accept, float qa) {
    const bool has1 = e1 != kNoEntry;
    const int bestDist = has1 ? (int)(e1 >> 23) : 256;
    const int bestIdx = has1 ? (int)(e1 & 0x7ffff) : -1;
    int bin = -1;
    if (accept && P.checkOri) {
        float rot = __fsub_rn(qa, T.kp4[bestIdx].z);
        if (rot < 0.0f) rot = __fadd_rn(rot, 360.0f);
        bin = (int)roundf(__fmul_rn(rot, 1.0f / ORBM_HISTO_LENGTH));
        if (bin == ORBM_HISTO_LENGTH) bin = 0;
    }
    P.qResult[q] = (accept || P.rule == RULE_NONE) ? bestIdx : -1;
    P.qDist[q] = bestDist;
    P.qBin[q] = bin;
}

// Executed by ONE thread.
__device__ __forceinline__ void apply_query(const SearchParams& P, const FrameDev& T, int q, uint32_t e1, bool accept, float qa,
                                            uint8_t* occS, uint16_t* mdS, int* m21S) {
    if (accept) update_state(P, q, e1, occS, mdS, m21S);
    write_result(P, T, q, e1, accept, qa);
}

__device__ __forceinline__ void commit_query(const SearchParams& P, const FrameDev& T, int q, uint32_t e1, uint32_t e2, float qa,
                                             uint8_t* occS, uint16_t* mdS, int* m21S) {
    apply_query(P, T, q, e1, decide_query(P, e1, e2), qa, occS, mdS, m21S);
}

// ---- lane-per-query walk of a grid window ---------------------------------------------------------------------------
// A tracking window holds a handful of candidates, so a warp per query leaves most lanes idle.  Here every LANE walks the
// window of its own query in the reference's order (Frame::GetFeaturesInArea, src/Frame.cc:657-723: columns outer, rows
// inner, insertion order inside a cell) and keeps the first four candidates in (distance, scan position) order that the
// match state does not block.  `count` = how many unblocked candidates were seen; count <= 4 means the list is complete.
struct Top4 { uint32_t t0, t1, t2, t3; int count; };

__device__ __forceinline__ void top4_insert(Top4& s, uint32_t e) {
    const uint32_t d = e >> 23;                       // strict '<' keeps earlier candidates first among equal distances
    if (d < (s.t3 >> 23)) {
        if (d < (s.t2 >> 23)) {
            s.t3 = s.t2;
            if (d < (s.t1 >> 23)) {
                s.t2 = s.t1;
                if (d < (s.t0 >> 23)) { s.t1 = s.t0; s.t0 = e; } else s.t1 = e;
            } else s.t2 = e;
        } else s.t3 = e;
    }
}

__device__ __forceinline__ void lane_walk(const SearchParams& P, const FrameDev& T, int q, const uint8_t* occS, const uint16_t* mdS,
                                          Top4& s, unsigned long long& nEval) {
    s.t0 = s.t1 = s.t2 = s.t3 = kNoEntry; s.count = 0;
    const float4 w = P.qwin[q];
    const int2 lv = P.qlev[q];
    const float x = w.x, y = w.y, r = w.z;
    const int minCX = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(x, T.minX), r), T.invW)));
    if (minCX >= kGridCols) return;
    const int maxCX = min(kGridCols - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(x, T.minX), r), T.invW)));
    if (maxCX < 0) return;
    const int minCY = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(y, T.minY), r), T.invH)));
    if (minCY >= kGridRows) return;
    const int maxCY = min(kGridRows - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(y, T.minY), r), T.invH)));
    if (maxCY < 0 || minCY > maxCY) return;
    uint32_t qd[8];
    {
        const uint4 a = reinterpret_cast<const uint4*>(P.qdesc)[2 * q], b = reinterpret_cast<const uint4*>(P.qdesc)[2 * q + 1];
        qd[0] = a.x; qd[1] = a.y; qd[2] = a.z; qd[3] = a.w; qd[4] = b.x; qd[5] = b.y; qd[6] = b.z; qd[7] = b.w;
    }
    const uint4* drows = reinterpret_cast<const uint4*>(T.desc);
    int ix = minCX, e = 0, re = 0;
    while (true) {
        if (e >= re) {                                  // next column: one contiguous CSR run (column-major grid)
            if (ix > maxCX) break;
            e = T.cellOff[ix * kGridRows + minCY];
            re = T.cellOff[ix * kGridRows + maxCY + 1];
            ++ix;
            continue;
        }
        const int idx = T.cellIdx[e++];
        const float4 k = T.kp4[idx];
        const int oct = __float_as_int(k.w);
        bool ok = oct >= lv.x && oct <= lv.y && fabsf(__fsub_rn(k.x, x)) < r && fabsf(__fsub_rn(k.y, y)) < r;
        if (ok && P.chi2) {
            const float ex = __fsub_rn(x, k.x), ey = __fsub_rn(y, k.y);
            const float e2 = __fadd_rn(__fmul_rn(ex, ex), __fmul_rn(ey, ey));
            ok = !((double)__fmul_rn(e2, P.invSigma2[oct]) > 5.99);
        }
        if (ok && P.elig) ok = P.elig[idx] != 0;
        if (!ok) continue;
        const int d = hamming8(qd, drows[2 * idx], drows[2 * idx + 1]);
        ++nEval;
        const uint32_t entry = pack_entry(d, oct, idx);
        if (blocked(P, occS, mdS, entry)) continue;
        ++s.count;
        top4_insert(s, entry);
    }
}

// first and second entry of a Top4 list that the CURRENT state does not block; returns how many were found (0..2)
__device__ __forceinline__ int top4_pick(const SearchParams& P, const Top4& s, const uint8_t* occS, const uint16_t* mdS,
                                         uint32_t& u1, uint32_t& u2) {
    // the four state reads are independent; an absent entry counts as blocked
    const bool k0 = s.t0 != kNoEntry && !(P.state != 0 && blocked(P, occS, mdS, s.t0));
    const bool k1 = s.t1 != kNoEntry && !(P.state != 0 && blocked(P, occS, mdS, s.t1));
    const bool k2 = s.t2 != kNoEntry && !(P.state != 0 && blocked(P, occS, mdS, s.t2));
    const bool k3 = s.t3 != kNoEntry && !(P.state != 0 && blocked(P, occS, mdS, s.t3));
    const unsigned m = (unsigned)k0 | ((unsigned)k1 << 1) | ((unsigned)k2 << 2) | ((unsigned)k3 << 3);
    const int a = __ffs(m) - 1;                       // first survivor (or -1)
    const unsigned m2 = m & (m - 1);
    const int b = __ffs(m2) - 1;                      // second survivor (or -1)
    u1 = a == 0 ? s.t0 : a == 1 ? s.t1 : a == 2 ? s.t2 : a == 3 ? s.t3 : kNoEntry;
    u2 = b == 1 ? s.t1 : b == 2 ? s.t2 : b == 3 ? s.t3 : kNoEntry;
    return (m != 0) + (m2 != 0);
}

// gather + distances of query q into a slab (scan order preserved by ballot compaction); returns the candidate count
__device__ __forceinline__ int gather_query(const SearchParams& P, const FrameDev& T, int q, uint32_t* slab, unsigned long long& nEval) {
    const int lane = threadIdx.x & 31;
    uint32_t qd[8];
    const uint4 a = reinterpret_cast<const uint4*>(P.qdesc)[2 * q], b = reinterpret_cast<const uint4*>(P.qdesc)[2 * q + 1];
    qd[0] = a.x; qd[1] = a.y; qd[2] = a.z; qd[3] = a.w; qd[4] = b.x; qd[5] = b.y; qd[6] = b.z; qd[7] = b.w;
    int count = 0;
    walk_candidates(P, T, q, qd, nEval, [&](bool ok, uint32_t entry) {
        const unsigned m = __ballot_sync(0xffffffffu, ok);
        if (ok) {
            const int pos = count + __popc(m & ((1u << lane) - 1));
            if (pos < kSlab) slab[pos] = entry;
        }
        count += __popc(m);
    });
    __syncwarp();
    return count;
}

// One CTA per search (state != 0) or per slice of queries (state == 0).
// dynamic smem: kSlots slabs | match state (occupancy bytes [n], or matched distance u16 [n] + match21 int [n]) | staged frame.
//
// Stateful searches reproduce the reference's sequential loop by SPECULATION.  The match state only ever blocks more
// candidates (an occupied keypoint stays occupied; vMatchedDistance only decreases), so the best / second-best candidate a
// query saw against an OLDER state is still its best / second-best as long as neither has been blocked since.
//   warps 1..15 (producers): gather candidates + distances of query q into slab q % kSlots, reduce them against the state
//       as it is at that moment, publish (best, second, angle) -> ready[slot] = q + 1.
//   warp 0 (resolver): takes 32 consecutive queries at a time, one per lane.  A query is "clean" when its speculative
//       best / second are still unblocked and no earlier query of the same batch wins one of them; clean prefixes commit
//       in parallel (distinct keypoints), the rest of the batch is replayed one query at a time (slab re-scan only when
//       the speculation failed).  `done` frees the slabs.
// kWindow selects the code that is compiled in: true = grid-window searches (lane per query), false = BoW / triangulation
// runs (warp per query).  Two kernels instead of one: with both bodies in a single ~8000-instruction kernel the BoW path
// faulted with "illegal instruction" on sm_100a (nvcc 12.9) although only dead code had been added.
template <bool kWindow>
__device__ __forceinline__ void search_cta(const SearchParams& P, uint8_t* dyn, int ctaIdx, int nCtas) {
    __shared__ volatile int ready[kSlots];
    __shared__ volatile int done;
    __shared__ int slotCount[kSlots];
    __shared__ float slotAngle[kSlots];
    __shared__ uint32_t slotE1[kSlots], slotE2[kSlots];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
#ifdef ORBM_STATS
    const long long kStart = clock64();
#endif
    const int n = P.tgt.n;
    uint32_t* slabs = reinterpret_cast<uint32_t*>(dyn);
    uint8_t* stateBase = dyn + search_slab_bytes(kWindow);
    uint8_t* occS = stateBase;                                        // state 1
    uint16_t* mdS = reinterpret_cast<uint16_t*>(stateBase);           // state 2
    int* m21S = reinterpret_cast<int*>(stateBase + (((size_t)n * 2 + 15) & ~(size_t)15));
    // The candidate walk is a chain of dependent loads (cell offsets -> keypoint index -> keypoint -> descriptor row).  When the
    // target frame fits (a few thousand keypoints), stage all four arrays in shared memory once per CTA so that the chain
    // runs at shared-memory latency instead of L2 latency.
    FrameDev T = P.tgt;
    if (P.stageTgt) {
        const size_t so = search_smem_bytes(n, P.state, false, kWindow);
        uint4* sDesc = reinterpret_cast<uint4*>(dyn + ((so + 15) & ~(size_t)15));
        float4* sKp4 = reinterpret_cast<float4*>(sDesc + 2 * (size_t)n);
        int* sOff = reinterpret_cast<int*>(sKp4 + n);
        int* sIdx = sOff + kGridCells + 1;
        const uint4* gDesc = reinterpret_cast<const uint4*>(P.tgt.desc);
        for (int i = tid; i < 2 * n; i += kSearchThreads) sDesc[i] = gDesc[i];
        for (int i = tid; i < n; i += kSearchThreads) { sKp4[i] = P.tgt.kp4[i]; sIdx[i] = P.tgt.cellIdx[i]; }
        for (int i = tid; i <= kGridCells; i += kSearchThreads) sOff[i] = P.tgt.cellOff[i];
        T.desc = reinterpret_cast<const uint32_t*>(sDesc); T.kp4 = sKp4; T.cellOff = sOff; T.cellIdx = sIdx;
    }
    if (P.state == 1) for (int i = tid; i < n; i += kSearchThreads) occS[i] = P.occ[i];
    if (P.state == 2) for (int i = tid; i < n; i += kSearchThreads) { mdS[i] = 0xffff; m21S[i] = -1; }
    if (tid < kSlots) ready[tid] = 0;
    if (tid == 0) done = 0;
    __syncthreads();

    const bool needSecond = rule_needs_second(P.rule);
    unsigned long long nEval = 0;
#ifdef ORBM_STATS
    const long long kT0 = clock64();
    if (tid == 0) atomicAdd(P.evals + 10, (unsigned long long)(kT0 - kStart));
#endif
    if (kWindow && P.state == 0) {
        // grid windows, independent queries: one query per lane
        for (int b = ctaIdx + wid * nCtas; b * 32 < P.nq; b += kSearchWarps * nCtas) {      // spread batches over CTAs first
            const int q = b * 32 + lane;
            if (q < P.nq) {
                Top4 t;
                lane_walk(P, T, q, occS, mdS, t, nEval);
                commit_query(P, T, q, t.t0, t.t1, P.checkOri ? P.qangle[q] : 0.f, occS, mdS, m21S);
            }
        }
    } else if (kWindow) {
        // grid windows against a match state: producers walk 32 queries per warp (one per lane) against the state as it is
        // and publish each query's Top4; the resolver validates and commits a batch of 32 at a time (see above; with four
        // candidates on file a blocked winner usually has its successor at hand and no second walk is needed).
        constexpr int kRingB = 32;
        uint4* hdrT = reinterpret_cast<uint4*>(dyn);                    // reuses the slab region: kRingB * 32 * 24 B = 24 KB
        int* hdrCount = reinterpret_cast<int*>(hdrT + kRingB * 32);
        float* hdrAngle = reinterpret_cast<float*>(hdrCount + kRingB * 32);
        const int nb = (P.nq + 31) >> 5;
        if (wid == 0) {
            const int needed = needSecond ? 2 : 1;
#ifdef ORBM_STATS
            long long tWait = 0, tRes = 0; int nIter = 0, nWalk = 0;
#endif
            for (int b = 0; b < nb; ++b) {
#ifdef ORBM_STATS
                const long long c0 = clock64();
#endif
                const int q = b * 32 + lane;
                const int at = (b % kRingB) * 32 + lane;
                if (lane == 0) while (ready[b % kRingB] != b + 1) { }
                __syncwarp();
#ifdef ORBM_STATS
                const long long c1 = clock64(); tWait += c1 - c0;
#endif
                bool pending = q < P.nq;
                Top4 t; float qa = 0.f;
                { const uint4 v = hdrT[at]; t.t0 = v.x; t.t1 = v.y; t.t2 = v.z; t.t3 = v.w; t.count = hdrCount[at]; qa = hdrAngle[at]; }
                uint32_t r1 = kNoEntry; bool racc = false;           // this lane's outcome, written once after the loop
                while (true) {
                    const unsigned R = __ballot_sync(0xffffffffu, pending);
                    if (!R) break;
#ifdef ORBM_STATS
                    ++nIter; const long long i0 = clock64();
#endif
                    uint32_t u1 = kNoEntry, u2 = kNoEntry;
                    bool insufficient = false;
                    if (pending) insufficient = top4_pick(P, t, occS, mdS, u1, u2) < needed && t.count > 4;
                    const int lowest = __ffs(R) - 1;
                    if (__shfl_sync(0xffffffffu, (int)insufficient, lowest)) {
                        // every earlier query is committed and the four candidates on file do not settle the lowest pending
                        // one: the whole warp walks its window again against the state as it is now (exact)
                        uint32_t e1, e2;
                        scan_candidates(P, T, b * 32 + lowest, nullptr, INT_MAX, occS, mdS, e1, e2);
                        if (lane == lowest) { u1 = e1; u2 = e2; insufficient = false; }
#ifdef ORBM_STATS
                        if (lane == 0) atomicAdd(P.evals + 5, 1ull);
#endif
                    }
                    const bool acc = pending && !insufficient && decide_query(P, u1, u2);
#ifdef ORBM_STATS
                    const long long i1c = clock64();
#endif
                    // an earlier pending query that is about to claim my best or second best invalidates my view
                    const int claim = acc ? (int)(u1 & 0x7ffff) : -1;
                    const int w1 = (pending && u1 != kNoEntry) ? (int)(u1 & 0x7ffff) : -2 - lane;     // unique when absent
                    const int w2 = (pending && u2 != kNoEntry) ? (int)(u2 & 0x7ffff) : -2;
                    const unsigned lower = (1u << lane) - 1;
                    const unsigned accMask = __ballot_sync(0xffffffffu, acc);
                    bool conflict = (__match_any_sync(0xffffffffu, w1) & accMask & lower) != 0;      // claim == that lane's w1
                    if (needSecond) {
#pragma unroll
                        for (int j = 0; j < 31; ++j) {
                            const int cj = __shfl_sync(0xffffffffu, claim, j);
                            if (j < lane && cj >= 0 && cj == w2) conflict = true;
                        }
                    }
                    const unsigned badMask = __ballot_sync(0xffffffffu, pending && (insufficient || conflict));
                    const int first = badMask ? __ffs(badMask) - 1 : 32;      // never the lowest pending lane: progress
#ifdef ORBM_STATS
                    const long long i2c = clock64();
#endif
                    if (pending && lane < first) {
                        if (acc) update_state(P, q, u1, occS, mdS, m21S);
                        r1 = u1; racc = acc; pending = false;
                    }
                    __syncwarp();
#ifdef ORBM_STATS
                    if (lane == 0) { atomicAdd(P.evals + 11, (unsigned long long)(i1c - i0)); atomicAdd(P.evals + 12, (unsigned long long)(i2c - i1c)); atomicAdd(P.evals + 13, (unsigned long long)(clock64() - i2c)); }
#endif
                }
                if (q < P.nq) write_result(P, T, q, r1, racc, qa);
                if (lane == 0) done = b + 1;
#ifdef ORBM_STATS
                tRes += clock64() - c1;
#endif
            }
#ifdef ORBM_STATS
            if (lane == 0) { atomicAdd(P.evals + 1, (unsigned long long)tWait); atomicAdd(P.evals + 3, (unsigned long long)tRes); atomicAdd(P.evals + 4, (unsigned long long)nIter); atomicAdd(P.evals + 6, 1ull); }
#endif
        } else {
#ifdef ORBM_STATS
            long long pWait = 0, pWork = 0;
#endif
            for (int b = wid - 1; b < nb; b += kSearchWarps - 1) {
#ifdef ORBM_STATS
                const long long p0 = clock64();
#endif
                const int q = b * 32 + lane;
                const int at = (b % kRingB) * 32 + lane;
                if (lane == 0) while (b - done >= kRingB) { __nanosleep(32); }
                __syncwarp();
#ifdef ORBM_STATS
                const long long p1 = clock64(); pWait += p1 - p0;
#endif
                Top4 t; t.t0 = t.t1 = t.t2 = t.t3 = kNoEntry; t.count = 0;
                float qa = 0.f;
                if (q < P.nq) { lane_walk(P, T, q, occS, mdS, t, nEval); if (P.checkOri) qa = P.qangle[q]; }
                hdrT[at] = make_uint4(t.t0, t.t1, t.t2, t.t3); hdrCount[at] = t.count; hdrAngle[at] = qa;
                __syncwarp();
                if (lane == 0) { __threadfence_block(); ready[b % kRingB] = b + 1; }
#ifdef ORBM_STATS
                pWork += clock64() - p1;
#endif
            }
#ifdef ORBM_STATS
            if (lane == 0) { atomicAdd(P.evals + 7, (unsigned long long)pWait); atomicAdd(P.evals + 8, (unsigned long long)pWork); }
#endif
        }
    } else if (!kWindow && P.state == 0) {
        // independent queries: every warp gathers and resolves its own
        uint32_t* slab = slabs + (size_t)wid * kSlab;
        for (int q = ctaIdx + wid * nCtas; q < P.nq; q += kSearchWarps * nCtas) {
            const int count = gather_query(P, T, q, slab, nEval);
            uint32_t e1, e2;
            scan_candidates(P, T, q, slab, count, occS, mdS, e1, e2);
            if (lane == 0) commit_query(P, T, q, e1, e2, P.checkOri ? P.qangle[q] : 0.f, occS, mdS, m21S);
            __syncwarp();
        }
    } else if (!kWindow && wid == 0) {
#ifdef ORBM_STATS
        long long tWait = 0, tPar = 0, tRep = 0; int nRep = 0, nRedo = 0;
#endif
        for (int q0 = 0; q0 < P.nq; q0 += 32) {
#ifdef ORBM_STATS
            long long c0 = clock64();
#endif
            const int q = q0 + lane;
            const bool live = q < P.nq;
            const int slot = q % kSlots;
            if (live) while (ready[slot] != q + 1) { }
            __syncwarp();
#ifdef ORBM_STATS
            long long c1 = clock64(); tWait += c1 - c0;
#endif
            uint32_t e1 = kNoEntry, e2 = kNoEntry;
            float qa = 0.f;
            bool bad = false;
            if (live) {
                e1 = slotE1[slot]; e2 = slotE2[slot]; qa = slotAngle[slot];
                if (e1 == kOverflowEntry) bad = true;
                else {
                    if (e1 != kNoEntry && blocked(P, occS, mdS, e1)) bad = true;
                    if (e2 != kNoEntry && blocked(P, occS, mdS, e2)) bad = true;
                }
            }
            // in-batch conflicts: an earlier query's winner is my winner or my runner-up
            const int i1 = (e1 < kOverflowEntry) ? (int)(e1 & 0x7ffff) : -1;
            const int i2 = (e1 < kOverflowEntry && e2 != kNoEntry) ? (int)(e2 & 0x7ffff) : -1;
            const unsigned same = __match_any_sync(0xffffffffu, i1);
            if (i1 >= 0 && (same & ((1u << lane) - 1))) bad = true;
            if (needSecond) {
                for (int j = 0; j < 31; ++j) {
                    const int o = __shfl_sync(0xffffffffu, i1, j);
                    if (j < lane && o >= 0 && o == i2) bad = true;
                }
            }
            const unsigned badMask = __ballot_sync(0xffffffffu, bad);
            const int first = badMask ? __ffs(badMask) - 1 : 32;
            if (live && lane < first) commit_query(P, T, q, e1, e2, qa, occS, mdS, m21S);
            __syncwarp();
#ifdef ORBM_STATS
            long long c2 = clock64(); tPar += c2 - c1;
#endif
            for (int i = first; i < 32 && q0 + i < P.nq; ++i) {          // replay in order, whole warp per query
                const int qq = q0 + i, s = qq % kSlots;
                uint32_t a1 = slotE1[s], a2 = slotE2[s];
                const bool redo = a1 == kOverflowEntry || (a1 != kNoEntry && blocked(P, occS, mdS, a1)) ||
                                  (a2 != kNoEntry && blocked(P, occS, mdS, a2));
#ifdef ORBM_STATS
                ++nRep; nRedo += redo;
#endif
                if (redo) scan_candidates(P, T, qq, slabs + (size_t)s * kSlab, slotCount[s], occS, mdS, a1, a2);
                if (lane == 0) commit_query(P, T, qq, a1, a2, slotAngle[s], occS, mdS, m21S);
                __syncwarp();
            }
            if (lane == 0) done = min(q0 + 32, P.nq);
#ifdef ORBM_STATS
            tRep += clock64() - c2;
#endif
        }
#ifdef ORBM_STATS
        if (lane == 0) { atomicAdd(P.evals + 1, (unsigned long long)tWait); atomicAdd(P.evals + 2, (unsigned long long)tPar); atomicAdd(P.evals + 3, (unsigned long long)tRep); atomicAdd(P.evals + 4, (unsigned long long)nRep); atomicAdd(P.evals + 5, (unsigned long long)nRedo); atomicAdd(P.evals + 6, 1ull); }
#endif
    } else {
#ifdef ORBM_STATS
        long long pWait = 0, pWork = 0;
#endif
        for (int q = wid - 1; q < P.nq; q += kSearchWarps - 1) {
            const int slot = q % kSlots;
#ifdef ORBM_STATS
            long long p0 = clock64();
#endif
            if (lane == 0) while (q - done >= kSlots) { __nanosleep(32); }
            __syncwarp();
#ifdef ORBM_STATS
            long long p1 = clock64(); pWait += p1 - p0;
#endif
            uint32_t* slab = slabs + (size_t)slot * kSlab;
            const int count = gather_query(P, T, q, slab, nEval);
            uint32_t e1 = kOverflowEntry, e2 = kNoEntry;
            if (count <= kSlab) scan_candidates(P, T, q, slab, count, occS, mdS, e1, e2);
            if (lane == 0) {
                slotCount[slot] = count;
                slotAngle[slot] = P.checkOri ? P.qangle[q] : 0.f;
                slotE1[slot] = e1; slotE2[slot] = e2;
                __threadfence_block();
                ready[slot] = q + 1;
            }
            __syncwarp();
#ifdef ORBM_STATS
            pWork += clock64() - p1;
#endif
        }
#ifdef ORBM_STATS
        if (lane == 0) { atomicAdd(P.evals + 7, (unsigned long long)pWait); atomicAdd(P.evals + 8, (unsigned long long)pWork); }
#endif
    }
#ifdef ORBM_STATS
    if (tid == 0) atomicAdd(P.evals + 9, (unsigned long long)(clock64() - kT0));
#endif
    // evaluation count (reporting only)
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) nEval += __shfl_xor_sync(0xffffffffu, nEval, o);
    if (lane == 0 && nEval) atomicAdd(P.evals, nEval);
    if (P.state != 0) __syncthreads();
    if (P.state == 1) for (int i = tid; i < n; i += kSearchThreads) P.occ[i] = occS[i];
    if (P.state == 2) {
        // a query keeps its match only if it is still the owner of the keypoint (vnMatches12[vnMatches21[i2]] = -1, :708)
        for (int q = tid; q < P.nq; q += kSearchThreads) {
            const int r = P.qResult[q];
            if (r >= 0 && m21S[r] != q) P.qResult[q] = -1;
        }
    }
}


template <bool kWindow>
__global__ void __launch_bounds__(kSearchThreads, 1) k_search(const SearchParams P) {
    extern __shared__ __align__(16) uint8_t dyn[];
    search_cta<kWindow>(P, dyn, blockIdx.x, gridDim.x);
}

// rotation-consistency epilogue (e.g. src/ORBmatcher.cc:1862-1884): histogram of accept events, three maxima, drop the rest
// encode: a dropped query leaves -(k + 2) instead of -1, k = the keypoint it had been assigned: with non-occupying matches
// (qocc) two queries can have been assigned the same keypoint, and the reference clears the KEYPOINT of every dropped
// histogram entry (CurrentFrame.mvpMapPoints[rotHist[i][j]] = NULL, :1876), so the host needs to know which one that was
__device__ __forceinline__ void rot_filter_cta(int nq, const int* __restrict__ qBin, int* __restrict__ qResult, bool encode = false) {
    __shared__ int hist[ORBM_HISTO_LENGTH];
    __shared__ int keep[3];
    const int tid = threadIdx.x;
    if (tid < ORBM_HISTO_LENGTH) hist[tid] = 0;
    __syncthreads();
    for (int q = tid; q < nq; q += blockDim.x) { const int b = qBin[q]; if (b >= 0) atomicAdd(&hist[b], 1); }
    __syncthreads();
    if (tid == 0) {
        int max1 = 0, max2 = 0, max3 = 0, ind1 = -1, ind2 = -1, ind3 = -1;
        for (int i = 0; i < ORBM_HISTO_LENGTH; ++i) {
            const int s = hist[i];
            if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
            else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
            else if (s > max3) { max3 = s; ind3 = i; }
        }
        if ((float)max2 < __fmul_rn(0.1f, (float)max1)) { ind2 = -1; ind3 = -1; }
        else if ((float)max3 < __fmul_rn(0.1f, (float)max1)) { ind3 = -1; }
        keep[0] = ind1; keep[1] = ind2; keep[2] = ind3;
    }
    __syncthreads();
    for (int q = tid; q < nq; q += blockDim.x) {
        const int b = qBin[q];
        if (b >= 0 && b != keep[0] && b != keep[1] && b != keep[2]) qResult[q] = (encode && qResult[q] >= 0) ? -2 - qResult[q] : -1;
    }
}
__global__ void __launch_bounds__(256) k_rot_filter(int nq, const int* __restrict__ qBin, int* __restrict__ qResult) {
    rot_filter_cta(nq, qBin, qResult, true);
}

// ---------------------------------------------------------------------------------------------------------------------
// SearchByBoW with one WARP PER SHARED NODE (M4a / M4b).  The match state of these two functions is the occupancy of side 2
// (vpMapPointMatches[realIdxF] != NULL, src/ORBmatcher.cc:276; vbMatched2[idx2], :826), and a feature of side 2 belongs to
// exactly ONE node of its FeatureVector, so queries of different nodes can never block each other: only the queries INSIDE a
// node are order-dependent.  A warp walks the queries of its node in the reference's order; when the node's run of side 2 has
// at most 32 features (the usual case: ~16 per node at levelsup 4) every lane keeps one candidate descriptor in registers and
// a query costs one uniform descriptor load, 8 POPC and two warp min-reductions.  The host checks that no side-2 feature is
// listed under two nodes (a hand-built orbm_featvec could do that) and otherwise uses the single-CTA path (k_search<false>).
// Fused epilogue: the last CTA to finish runs the rotation-consistency filter (no second launch on the latency path).
constexpr int kNodeWarps = 2;

// rotation bin of an accept event (e.g. src/ORBmatcher.cc:310-318): qa = query angle, ta = angle of the accepted keypoint
__device__ __forceinline__ int rot_bin(float qa, float ta) {
    float rot = __fsub_rn(qa, ta);
    if (rot < 0.0f) rot = __fadd_rn(rot, 360.0f);
    int bin = (int)roundf(__fmul_rn(rot, 1.0f / ORBM_HISTO_LENGTH));
    if (bin == ORBM_HISTO_LENGTH) bin = 0;
    return bin;
}

// Register path of a node: lane l keeps candidates l, l + 32, ... (kC of them: descriptor, index, octave, angle) for the whole
// node and the queries arrive 32 at a time, one per lane, then broadcast by shuffles: the sequential loop touches no memory
// except its own result stores.
template <int kC>
__device__ __forceinline__ void bow_node_regs(const SearchParams& P, int2 qr, int2 run, unsigned long long& nEval) {
    const int lane = threadIdx.x & 31;
    const FrameDev& T = P.tgt;
    const uint4* drows = reinterpret_cast<const uint4*>(T.desc);
    const uint4* qrows = reinterpret_cast<const uint4*>(P.qdescBase);
    const bool needSecond = rule_needs_second(P.rule);
    const int len = run.y - run.x;
    uint4 ca[kC], cb[kC];
    int cidx[kC], coct[kC];
    float cang[kC];
    unsigned haveM = 0, blk = 0;
#pragma unroll
    for (int j = 0; j < kC; ++j) {
        const int p = lane + 32 * j;
        ca[j] = cb[j] = make_uint4(0, 0, 0, 0); cidx[j] = 0; coct[j] = 0; cang[j] = 0.f;
        if (p < len) {
            const int idx = P.bowIdx[run.x + p];
            cidx[j] = idx;
            if (!P.elig || P.elig[idx] != 0) haveM |= 1u << j;
            ca[j] = drows[2 * idx]; cb[j] = drows[2 * idx + 1];
            const float4 k = T.kp4[idx];
            coct[j] = __float_as_int(k.w); cang[j] = k.z;
            if (P.occ[idx] != 0) blk |= 1u << j;
        }
    }
    const int nHave = __popc(haveM);
    for (int q0 = qr.x; q0 < qr.y; q0 += 32) {
        const int nb = min(32, qr.y - q0);
        uint4 ma = make_uint4(0, 0, 0, 0), mb = ma;
        float mang = 0.f;
        int mocc = 1;
        if (lane < nb) {
            const int src = P.qsrc[q0 + lane];
            ma = qrows[2 * src]; mb = qrows[2 * src + 1];
            if (P.checkOri) mang = P.qangle[q0 + lane];
            if (P.qocc) mocc = P.qocc[q0 + lane] != 0;
        }
        for (int i = 0; i < nb; ++i) {
            uint32_t qd[8];
            qd[0] = __shfl_sync(0xffffffffu, ma.x, i); qd[1] = __shfl_sync(0xffffffffu, ma.y, i);
            qd[2] = __shfl_sync(0xffffffffu, ma.z, i); qd[3] = __shfl_sync(0xffffffffu, ma.w, i);
            qd[4] = __shfl_sync(0xffffffffu, mb.x, i); qd[5] = __shfl_sync(0xffffffffu, mb.y, i);
            qd[6] = __shfl_sync(0xffffffffu, mb.z, i); qd[7] = __shfl_sync(0xffffffffu, mb.w, i);
            const float qa = __shfl_sync(0xffffffffu, mang, i);
            const int occupies = __shfl_sync(0xffffffffu, mocc, i);
            Best b; b.k1 = b.k2 = 0xffffffffu; b.e1 = b.e2 = kNoEntry;
#pragma unroll
            for (int j = 0; j < kC; ++j) {
                const int d = hamming8(qd, ca[j], cb[j]);
                if (((haveM & ~blk) >> j) & 1u) best_update(b, ((uint32_t)d << 20) | (uint32_t)(lane + 32 * j), pack_entry(d, coct[j], cidx[j]));
            }
            nEval += nHave;
            const uint32_t m1 = __reduce_min_sync(0xffffffffu, b.k1);
            const bool win = (b.k1 == m1) && (m1 != 0xffffffffu);          // keys are distinct: at most one lane
            const unsigned wmask = __ballot_sync(0xffffffffu, win);
            const int wl = wmask ? __ffs(wmask) - 1 : 0;
            const int wj = (int)((m1 & 0xfffffu) >> 5);
            float ta = 0.f;
#pragma unroll
            for (int j = 0; j < kC; ++j) if (j == wj) ta = cang[j];
            ta = __shfl_sync(0xffffffffu, ta, wl);
            uint32_t e1 = __shfl_sync(0xffffffffu, b.e1, wl), e2 = kNoEntry;
            if (!wmask) e1 = kNoEntry;
            if (needSecond) {
                const uint32_t cand2 = win ? b.k2 : b.k1;
                const uint32_t m2 = __reduce_min_sync(0xffffffffu, cand2);
                const unsigned smask = __ballot_sync(0xffffffffu, cand2 == m2 && m2 != 0xffffffffu);
                const uint32_t t2 = __shfl_sync(0xffffffffu, win ? b.e2 : b.e1, smask ? __ffs(smask) - 1 : 0);
                if (smask) e2 = t2;
            }
            const bool accept = decide_query(P, e1, e2);
            if (lane == 0) {
                const int q = q0 + i;
                const bool has1 = e1 != kNoEntry;
                P.qResult[q] = accept ? (int)(e1 & 0x7ffff) : -1;
                P.qDist[q] = has1 ? (int)(e1 >> 23) : 256;
                P.qBin[q] = (accept && P.checkOri) ? rot_bin(qa, ta) : -1;
            }
            if (accept && win && occupies) { blk |= 1u << wj; P.occ[e1 & 0x7ffff] = 1; }
        }
    }
}

__global__ void __launch_bounds__(kNodeWarps * 32) k_search_nodes(const SearchParams P) {
    const int lane = threadIdx.x & 31;
    const int node = blockIdx.x * kNodeWarps + (threadIdx.x >> 5);
    unsigned long long nEval = 0;
    if (node < P.nNodes) {
        const int2 qr = P.nodeQ[node];
        const int2 run = P.qrun[qr.x];
        const FrameDev& T = P.tgt;
        const uint4* drows = reinterpret_cast<const uint4*>(T.desc);
        const uint4* qrows = reinterpret_cast<const uint4*>(P.qdescBase);
        volatile uint8_t* occV = P.occ;
        const bool needSecond = rule_needs_second(P.rule);
        const int len = run.y - run.x;
        if (len <= 32) bow_node_regs<1>(P, qr, run, nEval);
        else if (len <= 64) bow_node_regs<2>(P, qr, run, nEval);
        else if (len <= 128) bow_node_regs<4>(P, qr, run, nEval);
        else {
            // very long run: candidates re-read per query (L1 / L2), occupancy through memory
            for (int q = qr.x; q < qr.y; ++q) {
                const int src = P.qsrc[q];
                const uint4 qa4 = qrows[2 * src], qb4 = qrows[2 * src + 1];
                const uint32_t qd[8] = {qa4.x, qa4.y, qa4.z, qa4.w, qb4.x, qb4.y, qb4.z, qb4.w};
                Best b; b.k1 = b.k2 = 0xffffffffu; b.e1 = b.e2 = kNoEntry;
                for (int p = lane; p < len; p += 32) {
                    const int idx = P.bowIdx[run.x + p];
                    if (P.elig && P.elig[idx] == 0) continue;
                    const int d = hamming8(qd, drows[2 * idx], drows[2 * idx + 1]);
                    ++nEval;
                    if (occV[idx] != 0) continue;
                    best_update(b, ((uint32_t)d << 20) | (uint32_t)p, pack_entry(d, __float_as_int(T.kp4[idx].w), idx));
                }
                const uint32_t m1 = __reduce_min_sync(0xffffffffu, b.k1);
                const bool win = (b.k1 == m1) && (m1 != 0xffffffffu);
                const unsigned wmask = __ballot_sync(0xffffffffu, win);
                uint32_t e1 = kNoEntry, e2 = kNoEntry;
                if (wmask) e1 = __shfl_sync(0xffffffffu, b.e1, __ffs(wmask) - 1);
                if (needSecond) {
                    const uint32_t cand2 = win ? b.k2 : b.k1;
                    const uint32_t m2 = __reduce_min_sync(0xffffffffu, cand2);
                    const unsigned smask = __ballot_sync(0xffffffffu, cand2 == m2 && m2 != 0xffffffffu);
                    if (smask) e2 = __shfl_sync(0xffffffffu, win ? b.e2 : b.e1, __ffs(smask) - 1);
                }
                if (lane == 0) { commit_query(P, T, q, e1, e2, P.checkOri ? P.qangle[q] : 0.f, P.occ, nullptr, nullptr); __threadfence_block(); }
                __syncwarp();
            }
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) nEval += __shfl_xor_sync(0xffffffffu, nEval, o);
    if (lane == 0 && nEval) atomicAdd(P.evals, nEval);
    if (!P.checkOri) return;
    // last CTA done: rotation-consistency filter over all queries (results of the other CTAs are visible after the fence)
    __shared__ int isLast;
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) isLast = atomicAdd(P.ticket, 1u) == gridDim.x - 1;
    __syncthreads();
    if (!isLast) return;
    __threadfence();
    rot_filter_cta(P.nq, P.qBin, P.qResult, true);
}

// ---------------------------------------------------------------------------------------------------------------------
// Device-resident tracking step (BASELINE config[1]): SearchByProjection(Cur = frame i, Last = frame i-1, th, bMono)
// (src/ORBmatcher.cc:1676-1887) for every consecutive pair of a batch of extractor results that never left HBM.
// Every keypoint of the last frame stands for a map point projected at its own position (identity motion prior).
struct TrackBatch {
    int nframes, stride, countStride;
    const orbx_keypoint* kps; const uint32_t* desc; const int* counts;     // extractor outputs (outKp / outDesc / outCnt)
    float minX, minY, invW, invH;
    float th, thF; int checkOri; int stageTgt;
    float scale[16];
    // per-frame work buffers, [nframes x stride] unless noted
    float4* kp4; int* cellOf; int* cellIdx; int* cellOff /* nframes x (kGridCells+1) */;
    float4* qwin; int2* qlev; float* qangle; uint8_t* occ;
    int* qResult; int* qDist; int* qBin;
    int* kpMatch; int* nMatches /* nframes */;
    unsigned long long* evals;
    uint4* lists; int* listCounts;          // claim rounds (orbm_kernels_rounds.cuh): the short list of every query, [nframes x stride]
};

__global__ void k_track_prepare(const TrackBatch T) {
    const int f = blockIdx.y, i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= T.stride) return;
    const size_t at = (size_t)f * T.stride + i;
    T.kpMatch[at] = -1; T.occ[at] = 0; T.qBin[at] = -1; T.qResult[at] = -1;
    const int n = min(T.counts[f * T.countStride], T.stride);
    if (i >= n) return;
    const orbx_keypoint k = T.kps[at];
    T.kp4[at] = make_float4(k.x, k.y, k.angle, __int_as_float(k.octave));
    const int px = (int)roundf(__fmul_rn(__fsub_rn(k.x, T.minX), T.invW));
    const int py = (int)roundf(__fmul_rn(__fsub_rn(k.y, T.minY), T.invH));
    T.cellOf[at] = (px < 0 || px >= kGridCols || py < 0 || py >= kGridRows) ? -1 : px * kGridRows + py;
    // the keypoint as a query into the NEXT frame: window th * scale[octave], levels [octave-1, octave+1] (:1731-1741)
    const int o = k.octave;
    T.qwin[at] = make_float4(k.x, k.y, __fmul_rn(T.th, T.scale[o & 15]), 0.f);
    T.qlev[at] = make_int2(o - 1, o + 1);            // bCheckLevels is true: maxLevel = o+1 >= 0
    T.qangle[at] = k.angle;
}

__global__ void __launch_bounds__(kGridThreads) k_track_grid(const TrackBatch T) {
    const int f = blockIdx.x;
    const int n = min(T.counts[f * T.countStride], T.stride);
    grid_build_cta(n, T.cellOf + (size_t)f * T.stride, T.cellOff + (size_t)f * (kGridCells + 1), T.cellIdx + (size_t)f * T.stride);
}

// the search of frame pair (last = pair, cur = pair + 1) in the terms of the generic engine
__device__ __forceinline__ void track_pair_params(const TrackBatch& T, int pair, SearchParams& P) {
    const int cur = pair + 1, last = pair;
    P.tgt.n = min(T.counts[cur * T.countStride], T.stride);
    P.tgt.kp4 = T.kp4 + (size_t)cur * T.stride;
    P.tgt.desc = T.desc + (size_t)cur * T.stride * 8;
    P.tgt.cellOff = T.cellOff + (size_t)cur * (kGridCells + 1);
    P.tgt.cellIdx = T.cellIdx + (size_t)cur * T.stride;
    P.tgt.minX = T.minX; P.tgt.minY = T.minY; P.tgt.invW = T.invW; P.tgt.invH = T.invH;
    P.nq = min(T.counts[last * T.countStride], T.stride);
    P.source = 0;
    P.qwin = T.qwin + (size_t)last * T.stride; P.qlev = T.qlev + (size_t)last * T.stride;
    P.qrun = nullptr; P.bowIdx = nullptr;
    P.qdesc = T.desc + (size_t)last * T.stride * 8;
    P.qangle = T.qangle + (size_t)last * T.stride;
    P.elig = nullptr;
    P.occ = T.occ + (size_t)cur * T.stride;
    P.qocc = nullptr;                  // every keypoint of the last frame stands for an observed map point: a match occupies
    P.stageTgt = T.stageTgt;
    P.state = 1; P.rule = RULE_BEST_LE; P.thF = T.thF; P.thI = 0; P.nnratio = 0.f; P.none2 = 256;
    P.checkOri = T.checkOri; P.chi2 = 0; P.invSigma2 = nullptr;
    P.tri = 0; P.coarse = 0; P.ep0 = P.ep1 = 0.f; P.scale2 = nullptr; P.sigma2 = nullptr;
    P.qResult = T.qResult + (size_t)cur * T.stride; P.qDist = T.qDist + (size_t)cur * T.stride; P.qBin = T.qBin + (size_t)cur * T.stride;
    P.evals = T.evals;
    P.nodeQ = nullptr; P.nNodes = 0; P.qsrc = nullptr; P.qdescBase = nullptr; P.ticket = nullptr;
}

__global__ void __launch_bounds__(kSearchThreads, 1) k_track_search(const TrackBatch T) {
    extern __shared__ __align__(16) uint8_t dyn[];
    SearchParams P;
    track_pair_params(T, blockIdx.x, P);
    search_cta<true>(P, dyn, 0, 1);
}

// rotation filter + scatter into the reference's output (CurrentFrame.mvpMapPoints[i2] = pMP  ->  kpMatch[i2] = query)
__global__ void __launch_bounds__(256) k_track_finish(const TrackBatch T) {
    __shared__ int cnt;
    const int cur = blockIdx.x + 1, last = blockIdx.x;
    const int nq = min(T.counts[last * T.countStride], T.stride);
    int* qRes = T.qResult + (size_t)cur * T.stride;
    if (threadIdx.x == 0) cnt = 0;
    if (T.checkOri) rot_filter_cta(nq, T.qBin + (size_t)cur * T.stride, qRes);
    __syncthreads();
    int mine = 0;
    for (int q = threadIdx.x; q < nq; q += blockDim.x) {
        const int r = qRes[q];
        if (r >= 0) { T.kpMatch[(size_t)cur * T.stride + r] = q; ++mine; }
    }
    if (mine) atomicAdd(&cnt, mine);
    __syncthreads();
    if (threadIdx.x == 0) { T.nMatches[cur] = cnt; if (cur == 1) T.nMatches[0] = 0; }
}

// ---------------------------------------------------------------------------------------------------------------------
// SearchForTriangulation with a caller-side epipolar predicate (any GeometricCamera, e.g. KannalaBrandt8): the candidates of
// every query keypoint of KF1 that survive everything the reference tests BEFORE pCamera1->epipolarConstrain
// (src/ORBmatcher.cc:1000-1033: no map point in KF2, DescriptorDistance <= TH_LOW, the epipole-distance test), in the
// reference's scan order (the KF2 feature list of the common vocabulary node).  The host then runs the reference's inner loop
// over these few candidates with the caller's own camera object.  One warp per query; a lane takes one candidate per round;
// the survivors of a round are written in scan order by a ballot rank.
__global__ void __launch_bounds__(256) k_tri_candidates(int nq, const int* __restrict__ qsrc, const int2* __restrict__ qrun,
                                                        const int* __restrict__ bowIdx, const uint32_t* __restrict__ desc1,
                                                        const uint32_t* __restrict__ desc2, const float4* __restrict__ kp4_2,
                                                        const uint8_t* __restrict__ free2, const float* __restrict__ scale2,
                                                        float ep0, float ep1, int th, int cap, int* __restrict__ outIdx,
                                                        int* __restrict__ outDist, int* __restrict__ outCount,
                                                        unsigned long long* __restrict__ evals) {
    const int q = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (q >= nq) return;
    const int i1 = qsrc[q];
    const int2 run = qrun[q];
    uint32_t d1[8];
    {
        const uint4 a = reinterpret_cast<const uint4*>(desc1 + (size_t)8 * i1)[0], b = reinterpret_cast<const uint4*>(desc1 + (size_t)8 * i1)[1];
        d1[0] = a.x; d1[1] = a.y; d1[2] = a.z; d1[3] = a.w; d1[4] = b.x; d1[5] = b.y; d1[6] = b.z; d1[7] = b.w;
    }
    int count = 0, ev = 0;
    for (int base = run.x; base < run.y; base += 32) {
        const int p = base + lane;
        bool ok = false;
        int idx2 = -1, dist = 0;
        if (p < run.y) {
            idx2 = bowIdx[p];
            if (free2[idx2]) {                                                  // :1001-1005 (vbMatched2 is never set in the reference)
                const uint4* r = reinterpret_cast<const uint4*>(desc2 + (size_t)8 * idx2);
                dist = hamming8(d1, r[0], r[1]);
                ++ev;
                if (dist <= th) {                                               // :1017 (the running best is applied by the host)
                    const float4 k = kp4_2[idx2];
                    const float dx = __fsub_rn(ep0, k.x), dy = __fsub_rn(ep1, k.y);      // :1026-1032, no FMA contraction
                    ok = !(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)) < __fmul_rn(100.f, scale2[__float_as_int(k.w)]));
                }
            }
        }
        const unsigned m = __ballot_sync(0xffffffffu, ok);
        const int pos = count + __popc(m & ((1u << lane) - 1u));
        if (ok && pos < cap) { outIdx[(size_t)q * cap + pos] = idx2; outDist[(size_t)q * cap + pos] = dist; }
        count += __popc(m);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) ev += __shfl_xor_sync(0xffffffffu, ev, o);
    if (lane == 0) { outCount[q] = count; if (ev) atomicAdd(evals, (unsigned long long)ev); }
}

// brute-force k-NN.  grid (query blocks, db chunks); block kKnnThreads; thread = query.
constexpr int kKnnThreads = 256, kKnnTileRows = 128, kKnnMaxK = 16;

// 256-bit Hamming distance with a carry-save adder tree: 8 xor words -> 4 words of weight 1, 1, 2, 4 -> 4 popc
__device__ __forceinline__ void csa(uint32_t& h, uint32_t& l, uint32_t a, uint32_t b, uint32_t c) {
    l = a ^ b ^ c;                      // LOP3 0x96
    h = (a & b) | (c & (a | b));        // LOP3 0xe8 (majority)
}
__device__ __forceinline__ int hamming_csa(const uint32_t (&q)[8], const uint4 a, const uint4 b) {
    const uint32_t x0 = q[0] ^ a.x, x1 = q[1] ^ a.y, x2 = q[2] ^ a.z, x3 = q[3] ^ a.w;
    const uint32_t x4 = q[4] ^ b.x, x5 = q[5] ^ b.y, x6 = q[6] ^ b.z, x7 = q[7] ^ b.w;
    uint32_t t0, o0, t1, o1, t2, o2, f0, tw;
    csa(t0, o0, x0, x1, x2);
    csa(t1, o1, x3, x4, x5);
    csa(t2, o2, o0, o1, x6);            // weight 1 left: o2 and x7
    csa(f0, tw, t0, t1, t2);            // weight 2: tw, weight 4: f0
    // four words remain (weights 1, 1, 2, 4): stopping here costs the same 4 POPC as a full ones/twos/fours/eights
    // reduction but 6 LOP3 less (LOP3 issues at 64 /clk/SM on sm_100a, POPC at 16: 16 LOP3 + 4 POPC balance the two pipes)
    return __popc(o2) + __popc(x7) + 2 * __popc(tw) + 4 * __popc(f0);
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// Optional fused epilogue of k_knn (kFused).  The CTA that finishes the LAST chunk of a query block (a done-counter per query
// block) merges the per-chunk lists of its queries right away - no k_knn_merge launch - and writes either the final distances /
// row ids (world == 0) or, for the sharded search, this rank's merged (distance, global row) keys straight into EVERY rank's
// exchange buffer over NVLink peer stores, followed by one release-add on each rank's arrival flag (orbm_knn_sharded; the
// receiving side is k_knn_xchg_merge).  That is the all-gather of the sharded search done by the kernel that produced the
// data: no NCCL call, no extra launch, the transfer of one query block overlaps the scan of the others.
constexpr int kXchgMaxRanks = 16;
struct KnnFused {
    unsigned int* bound;                              // per query: (smallest k-th best distance any chunk has published + 1) << 23, or all ones; NULL = off
    unsigned int* counter;
    uint16_t* dist; int32_t* idx;
    int world;
    unsigned long long* peerSlot[kXchgMaxRanks];      // rank r's buffer at [parity][this rank]
    unsigned int* peerFlag[kXchgMaxRanks];            // rank r's arrival counter [parity][this rank]
};

// dist/idx partial lists: part[(chunk * nq + q) * k + t] = key (dist << 32 | global row)
// kQ queries per thread (q, q + kKnnThreads, ...): one pair of 16-byte row loads serves kQ evaluations, and the kQ independent
// carry-save trees give the scheduler instruction-level parallelism; kK = compile-time bound of k (list kept in registers).
template <int kQ, int kK, bool kFused>
__global__ void __launch_bounds__(kKnnThreads, kFused ? 4 : 5) k_knn(      // fused: the merge epilogue must not cost the scan its 4 CTAs per SM
    const uint32_t* __restrict__ qdesc, int nq, const uint32_t* __restrict__ db,
                                                     long long ndb, long long rowOffset, int rowsPerChunk, int k,
                                                     unsigned long long* __restrict__ part, const KnnFused F) {
    __shared__ __align__(128) uint4 tile[2][kKnnTileRows * 2];
    __shared__ __align__(8) unsigned long long mbar[2];
    const int tid = threadIdx.x;
    const int qBase = blockIdx.x * kKnnThreads * kQ + tid;
    const long long r0 = (long long)blockIdx.y * rowsPerChunk;
    const long long r1 = min(ndb, r0 + rowsPerChunk);
    const int nrows = (int)max(0LL, r1 - r0);
    uint32_t qd[kQ][8];
    uint32_t best[kQ][kK];       // keys dist << 23 | chunk-local row, ascending
    uint32_t worst[kQ];
    // Chunks of one query block run side by side and each would warm its list up from scratch (~64 ln(rows / 64) warp-level
    // insertions); they share a bound instead: a row farther than the k-th best distance ANY chunk has seen cannot be in the final
    // top-k (equal distances stay: the lower row wins), so `worst` = min(own k-th best key, published bound).
    uint32_t gb[kQ];
#pragma unroll
    for (int j = 0; j < kQ; ++j) {
        const int qq = min(qBase + j * kKnnThreads, nq - 1);
        const uint4 a = reinterpret_cast<const uint4*>(qdesc)[2 * qq], b = reinterpret_cast<const uint4*>(qdesc)[2 * qq + 1];
        qd[j][0] = a.x; qd[j][1] = a.y; qd[j][2] = a.z; qd[j][3] = a.w; qd[j][4] = b.x; qd[j][5] = b.y; qd[j][6] = b.z; qd[j][7] = b.w;
#pragma unroll
        for (int t = 0; t < kK; ++t) best[j][t] = 0xffffffffu;
        worst[j] = 0xffffffffu;
        gb[j] = 0xffffffffu;
    }

    const int ntiles = (nrows + kKnnTileRows - 1) / kKnnTileRows;
    if (tid == 0) {
        for (int s = 0; s < 2; ++s)
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&mbar[s])));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    auto issue = [&](int t) {     // one elected thread: TMA bulk copy of a tile of rows into stage t & 1
        const int s = t & 1;
        const int rows = min(kKnnTileRows, nrows - t * kKnnTileRows);
        const uint32_t bytes = (uint32_t)rows * 32u;
        const uint32_t bar = smem_u32(&mbar[s]);
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(smem_u32(&tile[s][0])), "l"(db + (size_t)(r0 + (long long)t * kKnnTileRows) * 8), "r"(bytes), "r"(bar)
                     : "memory");
    };
    if (tid == 0 && ntiles > 0) issue(0);
    for (int t = 0; t < ntiles; ++t) {
        const int s = t & 1;
        if (tid == 0 && t + 1 < ntiles) issue(t + 1);
        {   // wait for stage s (phase parity = (t >> 1) & 1)
            const uint32_t bar = smem_u32(&mbar[s]);
            const uint32_t parity = (t >> 1) & 1;
            uint32_t done = 0;
            while (!done)
                asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                             : "=r"(done) : "r"(bar), "r"(parity) : "memory");
        }
        const int rows = min(kKnnTileRows, nrows - t * kKnnTileRows);
        const uint32_t rowBase = (uint32_t)(t * kKnnTileRows);
        if (F.bound) {      // once per tile: take the others' bound, publish mine if it is tighter
#pragma unroll
            for (int j = 0; j < kQ; ++j) {
                const int qq = qBase + j * kKnnThreads;
                if (qq < nq) {
                    const uint32_t seen = __ldcg(F.bound + qq);
                    uint32_t kth = 0xffffffffu;
#pragma unroll
                    for (int u = 0; u < kK; ++u) if (u == k - 1) kth = best[j][u];
                    const uint32_t mine = kth == 0xffffffffu ? 0xffffffffu : ((kth >> 23) + 1u) << 23;
                    if (mine < seen) atomicMin(F.bound + qq, mine);
                    gb[j] = min(seen, mine);
                    worst[j] = min(worst[j], gb[j]);
                }
            }
        }
        auto offer = [&](int j, uint32_t key) {
            // sorted insertion into the k best: bubble the new key through, drop the largest
            uint32_t carry = key;
#pragma unroll
            for (int u = 0; u < kK; ++u) {
                if (u < k) { const uint32_t lo = min(best[j][u], carry), hi = max(best[j][u], carry); best[j][u] = lo; carry = hi; worst[j] = min(lo, gb[j]); }
            }
        };
        // four rows at a time: one comparison of the smallest of their keys against the current k-th best decides whether
        // any of them can enter the list (almost never, once the list has warmed up)
        int r = 0;
        for (; r + 4 <= rows; r += 4) {
#pragma unroll
            for (int j = 0; j < kQ; ++j) {
                uint32_t key[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int d = hamming_csa(qd[j], tile[s][2 * (r + i)], tile[s][2 * (r + i) + 1]);
                    key[i] = ((uint32_t)d << 23) | (rowBase + (uint32_t)(r + i));
                }
                const uint32_t m = min(min(key[0], key[1]), min(key[2], key[3]));
                if (m < worst[j]) {
#pragma unroll
                    for (int i = 0; i < 4; ++i) if (key[i] < worst[j]) offer(j, key[i]);
                }
            }
        }
        for (; r < rows; ++r) {
#pragma unroll
            for (int j = 0; j < kQ; ++j) {
                const int d = hamming_csa(qd[j], tile[s][2 * r], tile[s][2 * r + 1]);
                const uint32_t key = ((uint32_t)d << 23) | (rowBase + (uint32_t)r);
                if (key < worst[j]) offer(j, key);
            }
        }
        __syncthreads();      // everyone is done with stage s before it is refilled (issue(t+2) happens next iteration)
    }
#pragma unroll
    for (int j = 0; j < kQ; ++j) {
        const int q = qBase + j * kKnnThreads;
        if (q < nq) {
            unsigned long long* out = part + ((size_t)blockIdx.y * nq + q) * k;
#pragma unroll
            for (int u = 0; u < kK; ++u) {
                if (u < k) {
                    const uint32_t key = best[j][u];
                    out[u] = key == 0xffffffffu ? 0xffffffffffffffffull
                                                : (((unsigned long long)(key >> 23)) << 32) | (unsigned long long)(rowOffset + r0 + (key & 0x7fffffu));
                }
            }
        }
    }
    if (!kFused) return;
    // fused merge: last chunk of this query block done -> this CTA merges the block's lists (thread = query, keys ascending =
    // (distance, global row): the strict-< rule of a single scan)
    __shared__ int sLast;
    __threadfence();
    __syncthreads();
    if (tid == 0) sLast = atomicAdd(&F.counter[blockIdx.x], 1u) == gridDim.y - 1;
    __syncthreads();
    if (!sLast) return;
    if (tid == 0) F.counter[blockIdx.x] = 0;                    // re-armed for the next call on this stream
    __threadfence();
#pragma unroll 1
    for (int j = 0; j < kQ; ++j) {
        const int q = qBase + j * kKnnThreads;
        if (q >= nq) continue;
        if (F.bound) F.bound[q] = 0xffffffffu;
        unsigned long long top[kK];
#pragma unroll
        for (int u = 0; u < kK; ++u) top[u] = 0xffffffffffffffffull;
        unsigned long long worst = 0xffffffffffffffffull;        // top[k - 1]
        // eight independent loads in flight per round (a dependent load per key would cost ~150 L2 round trips per query)
        const int nkeys = (int)gridDim.y * k;
#pragma unroll 1
        for (int e0 = 0; e0 < nkeys; e0 += 8) {
            unsigned long long key[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int e = e0 + i, c = e / k, u = e - c * k;
                key[i] = e < nkeys ? __ldcg(part + ((size_t)c * nq + q) * k + u) : 0xffffffffffffffffull;
            }
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                unsigned long long carry = key[i];
                if (carry < worst) {
#pragma unroll
                    for (int v = 0; v < kK; ++v) {
                        if (v < k) { const unsigned long long lo = top[v] < carry ? top[v] : carry, hi = top[v] < carry ? carry : top[v]; top[v] = lo; carry = hi; worst = lo; }
                    }
                }
            }
        }
        if (F.world == 0) {
#pragma unroll
            for (int u = 0; u < kK; ++u) {
                if (u < k) {
                    const bool none = top[u] == 0xffffffffffffffffull;
                    F.dist[(size_t)q * k + u] = none ? (uint16_t)65535 : (uint16_t)(top[u] >> 32);
                    F.idx[(size_t)q * k + u] = none ? -1 : (int32_t)(top[u] & 0xffffffffull);
                }
            }
        } else {
            for (int r = 0; r < F.world; ++r) {
                unsigned long long* dst = F.peerSlot[r] + (size_t)q * k;
#pragma unroll
                for (int u = 0; u < kK; ++u) if (u < k) dst[u] = top[u];
            }
        }
    }
    if (F.world > 0) {
        // every thread's peer stores are ordered before the arrival signal: fence (system scope), CTA barrier, one release-add per rank
        __threadfence_system();
        __syncthreads();
        if (tid < F.world) asm volatile("red.release.sys.global.add.u32 [%0], %1;" ::"l"(F.peerFlag[tid]), "r"(1u) : "memory");
    }
}

// Sending side of the sharded search, unfused form (the default): the per-chunk lists of one query are merged by one warp
// (k_knn_merge's rule) and the merged keys go straight to EVERY rank's exchange buffer - lane r stores to rank r - followed by
// one release-add per rank and CTA on the arrival counters.  250 CTAs work in parallel here, where the fused epilogue of
// k_knn leaves the merge of a query block to ONE CTA at the very end of the scan (measured +60 us on a 125 k-row shard).
__global__ void __launch_bounds__(256) k_knn_merge_push(const unsigned long long* __restrict__ part, int nchunks, int nq, int k, const KnnFused F) {
    const int q = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (q < nq) {
        if (F.bound && lane == 0) F.bound[q] = 0xffffffffu;
        const int total = nchunks * k;
        unsigned long long prev = 0; bool first = true;
        for (int t = 0; t < k; ++t) {
            unsigned long long m = 0xffffffffffffffffull;
            for (int e = lane; e < total; e += 32) {
                const int c = e / k, u = e - c * k;
                const unsigned long long key = part[((size_t)c * nq + q) * k + u];
                if ((first || key > prev) && key < m) m = key;
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) { const unsigned long long x = __shfl_xor_sync(0xffffffffu, m, o); m = x < m ? x : m; }
            if (lane < F.world) F.peerSlot[lane][(size_t)q * k + t] = m;          // all-ones = no neighbour (the lists end there)
            if (m == 0xffffffffffffffffull) {
                for (int u = t + 1; u < k; ++u) if (lane < F.world) F.peerSlot[lane][(size_t)q * k + u] = m;
                break;
            }
            prev = m; first = false;
        }
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x < F.world) asm volatile("red.release.sys.global.add.u32 [%0], %1;" ::"l"(F.peerFlag[threadIdx.x]), "r"(1u) : "memory");
}

// Receiving side of the sharded search: waits until every rank's arrival counter of this parity has reached `expected` (all
// query blocks of all ranks have landed in THIS rank's exchange buffer), then merges the world lists of each query: one warp
// per query, keys ascending.  A rank that never shows up must not hang the GPU: after timeoutNs the kernel gives up, raises
// *status and writes "no neighbour".
__global__ void __launch_bounds__(256) k_knn_xchg_merge(const unsigned long long* __restrict__ slots, size_t slotKeys,
                                                        const unsigned int* __restrict__ flags, unsigned int expected, int world, int nq,
                                                        int k, uint16_t* __restrict__ outDist, int32_t* __restrict__ outIdx,
                                                        unsigned int* __restrict__ status, unsigned long long timeoutNs) {
    __shared__ int sOk;
    if (threadIdx.x == 0) {
        unsigned long long t0, t1;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
        int ok = 1;
        for (int r = 0; r < world && ok; ++r) {
            unsigned int v;
            do {
                asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(flags + r) : "memory");
                if ((int)(v - expected) >= 0) break;
                asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
                if (t1 - t0 > timeoutNs) { ok = 0; atomicExch(status, 1u); }
                else __nanosleep(200);
            } while (ok);
        }
        sOk = ok;
    }
    __syncthreads();
    const int q = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (q >= nq) return;
    const int total = world * k;
    unsigned long long prev = 0; bool first = true;
    for (int t = 0; t < k; ++t) {
        unsigned long long m = 0xffffffffffffffffull;
        if (sOk) {
            for (int e = lane; e < total; e += 32) {
                const int c = e / k, u = e - c * k;
                const unsigned long long key = __ldcg(slots + (size_t)c * slotKeys + (size_t)q * k + u);     // L2: where peer stores land
                if ((first || key > prev) && key < m) m = key;
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) { const unsigned long long x = __shfl_xor_sync(0xffffffffu, m, o); m = x < m ? x : m; }
        if (lane == 0) {
            if (m == 0xffffffffffffffffull) { outDist[(size_t)q * k + t] = 65535; outIdx[(size_t)q * k + t] = -1; }
            else { outDist[(size_t)q * k + t] = (uint16_t)(m >> 32); outIdx[(size_t)q * k + t] = (int32_t)(m & 0xffffffffull); }
        }
        if (m == 0xffffffffffffffffull) { for (int u = t + 1; u < k && lane == 0; ++u) { outDist[(size_t)q * k + u] = 65535; outIdx[(size_t)q * k + u] = -1; } break; }
        prev = m; first = false;
    }
}

// merge the per-chunk lists of one query: k smallest (dist, row) keys.  One warp per query.
__global__ void __launch_bounds__(256) k_knn_merge(const unsigned long long* __restrict__ part, int nchunks, int nq, int k,
                                                   uint16_t* __restrict__ dist, int32_t* __restrict__ idx, unsigned int* __restrict__ bound) {
    const int q = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (q >= nq) return;
    if (bound && lane == 0) bound[q] = 0xffffffffu;       // the scan's shared bound is back at rest for the next call
    const int total = nchunks * k;
    unsigned long long prev = 0; bool first = true;
    for (int t = 0; t < k; ++t) {
        // smallest key strictly greater than the previous pick (keys are unique: distinct rows)
        unsigned long long m = 0xffffffffffffffffull;
        for (int e = lane; e < total; e += 32) {
            const int c = e / k, u = e - c * k;
            const unsigned long long key = part[((size_t)c * nq + q) * k + u];
            if ((first || key > prev) && key < m) m = key;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) { const unsigned long long x = __shfl_xor_sync(0xffffffffu, m, o); m = x < m ? x : m; }
        if (lane == 0) {
            if (m == 0xffffffffffffffffull) { dist[(size_t)q * k + t] = 65535; idx[(size_t)q * k + t] = -1; }
            else { dist[(size_t)q * k + t] = (uint16_t)(m >> 32); idx[(size_t)q * k + t] = (int32_t)(m & 0xffffffffull); }
        }
        prev = m; first = false;
        if (m == 0xffffffffffffffffull) { for (int u = t + 1; u < k && lane == 0; ++u) { dist[(size_t)q * k + u] = 65535; idx[(size_t)q * k + u] = -1; } break; }
    }
}

// merge of the per-SHARD top-k lists after the all-gather (config 5): shard s keeps its nq x k distances at
// dist + s * strideBytes and its global row ids at idx + s * strideBytes.  Same rule as the single scan: (distance, global
// row) ascending, so the result equals the unsharded search bit for bit.  One warp per query.
__global__ void __launch_bounds__(256) k_knn_merge_shards(const uint8_t* __restrict__ dist, const uint8_t* __restrict__ idx,
                                                          size_t strideBytes, int nshards, int nq, int k,
                                                          uint16_t* __restrict__ outDist, int32_t* __restrict__ outIdx) {
    const int q = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (q >= nq) return;
    const int total = nshards * k;
    unsigned long long prev = 0; bool first = true;
    for (int t = 0; t < k; ++t) {
        unsigned long long m = 0xffffffffffffffffull;
        for (int e = lane; e < total; e += 32) {
            const int c = e / k, u = e - c * k;
            const int32_t row = reinterpret_cast<const int32_t*>(idx + (size_t)c * strideBytes)[(size_t)q * k + u];
            const uint16_t d = reinterpret_cast<const uint16_t*>(dist + (size_t)c * strideBytes)[(size_t)q * k + u];
            if (row < 0) continue;
            const unsigned long long key = ((unsigned long long)d << 32) | (uint32_t)row;
            if ((first || key > prev) && key < m) m = key;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) { const unsigned long long x = __shfl_xor_sync(0xffffffffu, m, o); m = x < m ? x : m; }
        if (lane == 0) {
            if (m == 0xffffffffffffffffull) { outDist[(size_t)q * k + t] = 65535; outIdx[(size_t)q * k + t] = -1; }
            else { outDist[(size_t)q * k + t] = (uint16_t)(m >> 32); outIdx[(size_t)q * k + t] = (int32_t)(m & 0xffffffffull); }
        }
        if (m == 0xffffffffffffffffull) { for (int u = t + 1; u < k && lane == 0; ++u) { outDist[(size_t)q * k + u] = 65535; outIdx[(size_t)q * k + u] = -1; } break; }
        prev = m; first = false;
    }
}

// MapPoint::ComputeDistinctiveDescriptors selection (src/MapPoint.cc:366-396): one warp per map point (N <= 32 rows
// handled lane = row; larger N loops).  median = sorted row distances [0.5*(N-1)] (self distance 0 included).
__global__ void __launch_bounds__(256) k_distinctive(const uint32_t* __restrict__ desc, const int* __restrict__ off, int npoints,
                                                     int* __restrict__ best) {
    const int p = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (p >= npoints) return;
    const int b = off[p], N = off[p + 1] - b;
    if (N <= 0) { if (lane == 0) best[p] = -1; return; }
    const int mpos = (int)(0.5 * (N - 1));
    const uint4* rows = reinterpret_cast<const uint4*>(desc);
    unsigned bestKey = 0xffffffffu;       // median << 16 | row
    for (int i = lane; i < N; i += 32) {
        uint32_t qd[8];
        const uint4 a = rows[2 * (b + i)], c = rows[2 * (b + i) + 1];
        qd[0] = a.x; qd[1] = a.y; qd[2] = a.z; qd[3] = a.w; qd[4] = c.x; qd[5] = c.y; qd[6] = c.z; qd[7] = c.w;
        // the median = value v with  #(d < v) <= mpos < #(d <= v): count-based selection over the 257 possible values
        int lo = 0, hi = 256;
        while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            int le = 0;
            for (int j = 0; j < N; ++j) le += hamming8(qd, rows[2 * (b + j)], rows[2 * (b + j) + 1]) <= mid;
            if (le > mpos) hi = mid; else lo = mid + 1;
        }
        const unsigned key = ((unsigned)lo << 16) | (unsigned)i;
        bestKey = min(bestKey, key);
    }
    bestKey = __reduce_min_sync(0xffffffffu, bestKey);
    if (lane == 0) best[p] = (int)(bestKey & 0xffff);
}

}  // namespace orbm
