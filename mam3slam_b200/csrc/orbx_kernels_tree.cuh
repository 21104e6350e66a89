// E3: DistributeOctTree on device (src/ORBextractor.cc:555-779; DivideNode :480-536; compareNodes :538-553).
//
// The reference walks a std::list of nodes that own std::vector<KeyPoint> copies.  Here one CTA owns one
// (frame, level) and keeps:
//   * the keys as 64-bit words (x | y<<16 | score<<32) in two ping-pong HBM/L2 buffers; a node owns a contiguous
//     segment [beg, beg+cnt) of one of them, and splitting a node is a STABLE 4-way partition of that segment into
//     the other buffer (stable = the reference's push_back order inside every child);
//   * the list as an ARRAY in shared memory that is rebuilt after every pass: a pass of the reference only ever
//     pushes children to the list front and erases the split parents, so the new list is
//         [children of the nodes split in this pass, last split first, each as n4,n3,n2,n1] ++ [unsplit nodes in old order],
//     which is computed with prefix sums instead of pointer surgery;
//   * the "split the largest first" phase (:689-753) with an exact emulation of libstdc++ std::sort
//     (introsort_emul.h), because compareNodes ties are resolved by the sort's element movement.
// Every partition is speculative-safe: children are written to the other buffer, so a node that ends up not being
// split (the >= N early exit) still owns intact data.
#pragma once
#include "orbx_internal.h"
#include "introsort_emul.h"

namespace orbx {

// The CTA size is a launch parameter (128 ... 512): the tree is latency-bound, so a small batch wants many threads per level while
// a 64-frame batch runs faster overall with 128 - a 512-thread CTA pins half an SM's registers for ~100 us at 40 % issue
// utilisation and keeps the other kernels of the step off that SM (measured: 512 -> 256 -> 128 threads = 112.8 k -> 118.8 k ->
// 122.0 k frames/s per 64-frame step, while the tree of ONE frame takes 70 / 82 / ~100 us with 512 / 256 / 128 threads).
constexpr int kTreeMaxThreads = 512, kTreeMaxWarps = kTreeMaxThreads / 32;
#define kTreeThreads ((int)blockDim.x)
#define kTreeWarps ((int)(blockDim.x >> 5))
#ifndef ORBX_CTA_PART_MIN
#define ORBX_CTA_PART_MIN 1024
#endif
constexpr int kCtaPartitionMin = ORBX_CTA_PART_MIN;   // nodes with at least this many keys are partitioned by the whole CTA

struct TreeSmem {                   // carved out of dynamic shared memory, sized by poolCap P and scanCap
    ushort4* rect[2];               // x0,y0,x1,y1
    uint32_t* beg[2];               // segment start | buffer flag << 31
    uint32_t* cnt[2];
    ise::Elem* pend;                // pending (splittable) nodes of the current list: key=(cnt<<32)|x0, id=list index
    uint4* cc;                      // child key counts of a node being split, indexed by list position
    int* scanA;                     // generic scan array (>= max(P, cells per level)+1)
    int* scanB;
    uint16_t* proc;                 // processing order (list positions)
    uint8_t* dv;                    // node was split in this pass
    int scanCap;                    // ints in scanA / scanB
};

__device__ __forceinline__ int warp_incl_scan(int v, int lane) {
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { int t = __shfl_up_sync(0xffffffffu, v, o); if (lane >= o) v += t; }
    return v;
}

// In-place exclusive scan of a[0..n) by the whole CTA; returns the total.  Ends with a barrier.
__device__ int block_exscan(int* a, int n, int* sWarp /* >= 33 ints */) {
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int per = (n + kTreeThreads - 1) / kTreeThreads;
    const int b = min(tid * per, n), e = min(b + per, n);
    int sum = 0;
    for (int i = b; i < e; ++i) sum += a[i];
    const int incl = warp_incl_scan(sum, lane);
    if (lane == 31) sWarp[wid] = incl;
    __syncthreads();
    if (wid == 0) {
        int v = lane < kTreeWarps ? sWarp[lane] : 0;
        int iv = warp_incl_scan(v, lane);
        sWarp[lane] = iv - v;
        if (lane == 31) sWarp[32] = iv;
    }
    __syncthreads();
    int run = sWarp[wid] + incl - sum;
    for (int i = b; i < e; ++i) { int t = a[i]; a[i] = run; run += t; }
    const int total = sWarp[32];
    __syncthreads();
    return total;
}

// class of a key for DivideNode: n1=0 (x<mx,y<my), n2=1 (x>=mx,y<my), n3=2 (x<mx,y>=my), n4=3
struct QuadClass {
    int mx, my;
    __device__ __forceinline__ int operator()(uint64_t k) const {
        const int x = (int)(k & 0xffff), y = (int)((k >> 16) & 0xffff);
        return (x < mx ? 0 : 1) + (y < my ? 0 : 2);
    }
};
// class of a key for the root assignment vpIniNodes[kp.pt.x/hX] (:585), restricted to roots [base, base+4)
struct RootClass {
    float hX; int base;
    __device__ __forceinline__ int operator()(uint64_t k) const {
        const float x = (float)(int)(k & 0xffff);
        return (int)__fdiv_rn(x, hX) - base;        // outside 0..3 = not in this group
    }
};

// ---- count pyramid (second form of the tree, used when it fits) -------------------------------------------------------
// DivideNode's split points depend on the node's rectangle only, never on its keys: the cells of depth d under a root are a
// fixed (irregular, ceil-halved) 2^d x 2^d grid.  So instead of physically partitioning the keys pass after pass, every key
// descends kPyrDepth levels ONCE (2 * kPyrDepth comparisons), bumps the count of its finest cell and offers itself as that
// cell's best key; four-to-one sums / maxima give every coarser depth.  The list logic below (which nodes are split in which
// order - the part of DistributeOctTree that decides the result) then runs on counts alone: child counts are four loads, the
// kept key of a final node is one load.  A node the logic wants to split at depth kPyrDepth (clustered keys, or fewer keys
// than the quota) aborts this form and the partitioning form runs instead: the keys themselves were never moved.
constexpr int kPyrDepth = 5;
constexpr int kPyrCellsPerRoot = (4 * 4 * 4 * 4 * 4 * 4 - 1) / 3;      // depths 0..5: 1365 cells
constexpr int kPyrMaxRoots = 4;
struct PyrView { uint32_t* cnt; uint32_t* best; };
__host__ __device__ inline int pyr_offset(int d) { return ((1 << (2 * d)) - 1) / 3; }
// node address in pyramid form (kept in TreeSmem::beg): root << 28 | depth << 24 | cy << 12 | cx
__device__ __forceinline__ uint32_t pyr_addr(int root, int d, int cx, int cy) { return ((uint32_t)root << 28) | ((uint32_t)d << 24) | ((uint32_t)cy << 12) | (uint32_t)cx; }
__device__ __forceinline__ int pyr_cell(uint32_t a) {
    const int root = (int)(a >> 28), d = (int)((a >> 24) & 15), cy = (int)((a >> 12) & 0xfff), cx = (int)(a & 0xfff);
    return root * kPyrCellsPerRoot + pyr_offset(d) + (cy << d) + cx;
}
// best-key word (32 bits, so that the shared-memory atomicMax is one native instruction): score << 20 | (0xfffff - position in
// vToDistributeKeys).  The maximum is the highest response and, among equals, the FIRST key in list order (:757-776); the key
// itself is read back from the gathered list by its position.
__device__ __forceinline__ uint32_t pyr_best_word(uint32_t score, uint32_t pos) { return (score << 20) | (0xfffffu - pos); }
__device__ __forceinline__ uint32_t pyr_best_pos(uint32_t w) { return 0xfffffu - (w & 0xfffffu); }
// one key into the pyramid's finest depth (any thread)
__device__ __forceinline__ void pyr_insert(const PyrView& PV, uint32_t x, uint32_t y, uint32_t score, uint32_t pos, float hX, int nIni, int treeH) {
    score = min(score, 0xfffu);
    const int root = (int)__fdiv_rn((float)(int)x, hX);
    if (root < 0 || root >= nIni) return;                              // RootClass: a key outside every root is in no node
    int ax = (int)__fmul_rn(hX, (float)root), bx = (int)__fmul_rn(hX, (float)(root + 1)), ay = 0, by = treeH, cx = 0, cy = 0;
#pragma unroll
    for (int d = 0; d < kPyrDepth; ++d) {
        const int mx = ax + ((bx - ax + 1) >> 1), my = ay + ((by - ay + 1) >> 1);      // split_of
        if ((int)x < mx) { bx = mx; cx = 2 * cx; } else { ax = mx; cx = 2 * cx + 1; }
        if ((int)y < my) { by = my; cy = 2 * cy; } else { ay = my; cy = 2 * cy + 1; }
    }
    const int cell = root * kPyrCellsPerRoot + pyr_offset(kPyrDepth) + (cy << kPyrDepth) + cx;
    atomicAdd(PV.cnt + cell, 1u);
    atomicMax(PV.best + cell, pyr_best_word(score, pos));
}
// coarser depths from the finest one (whole CTA; ends with a barrier)
__device__ void pyr_reduce(const PyrView& PV, int nIni) {
    for (int d = kPyrDepth - 1; d >= 0; --d) {
        __syncthreads();
        const int w = 1 << d, cells = w * w;
        for (int i = threadIdx.x; i < nIni * cells; i += blockDim.x) {
            const int root = i / cells, c = i - root * cells, cy = c >> d, cx = c & (w - 1);
            const int child = root * kPyrCellsPerRoot + pyr_offset(d + 1) + ((2 * cy) << (d + 1)) + 2 * cx;
            const int w2 = 2 * w;
            PV.cnt[root * kPyrCellsPerRoot + pyr_offset(d) + c] = PV.cnt[child] + PV.cnt[child + 1] + PV.cnt[child + w2] + PV.cnt[child + w2 + 1];
            PV.best[root * kPyrCellsPerRoot + pyr_offset(d) + c] = max(max(PV.best[child], PV.best[child + 1]), max(PV.best[child + w2], PV.best[child + w2 + 1]));
        }
    }
    __syncthreads();
}
// child counts of the nodes proc[0..T) from the pyramid (threads firstThread.. of the CTA); a node at the finest depth raises *abort
__device__ void pyr_children(TreeSmem& S, int cur, int T, const PyrView& PV, int firstThread, int* abort) {
    for (int t = (int)threadIdx.x - firstThread; t >= 0 && t < T; t += kTreeThreads - firstThread) {
        const int node = S.proc[t];
        const uint32_t a = S.beg[cur][node];
        const int root = (int)(a >> 28), d = (int)((a >> 24) & 15), cy = (int)((a >> 12) & 0xfff), cx = (int)(a & 0xfff);
        if (d >= kPyrDepth) { *abort = 1; S.cc[node] = make_uint4(0, 0, 0, 0); continue; }
        const int w2 = 2 << d;
        const uint32_t* c = PV.cnt + root * kPyrCellsPerRoot + pyr_offset(d + 1) + (2 * cy) * w2 + 2 * cx;
        S.cc[node] = make_uint4(c[0], c[1], c[w2], c[w2 + 1]);          // n1 (x<,y<), n2 (x>=,y<), n3 (x<,y>=), n4
    }
}

// Stable 4-way partition of src[beg, beg+cnt) into dst starting at outBase, by one warp.  Returns class counts.
// The loops are bound by the L2 round trip of their key loads, not by the ballots: keys are therefore fetched kPartU
// blocks of 32 at a time (all loads of a group are issued before the first one is used), and a node of at most
// 32 * kPartU keys — every node from the third pass on — is read exactly once, its keys staying in registers between the
// counting and the scattering step.
constexpr int kPartU = 8;
template <class F>
__device__ __forceinline__ void partition_count_block(int q, int& n0, int& n1, int& n2, int& n3) {
    n0 += __popc(__ballot_sync(0xffffffffu, q == 0));
    n1 += __popc(__ballot_sync(0xffffffffu, q == 1));
    n2 += __popc(__ballot_sync(0xffffffffu, q == 2));
    n3 += __popc(__ballot_sync(0xffffffffu, q == 3));
}
__device__ __forceinline__ void partition_scatter_block(uint64_t* dst, uint64_t k, int q, uint32_t lt, int& o0, int& o1, int& o2, int& o3) {
    const uint32_t b0 = __ballot_sync(0xffffffffu, q == 0), b1 = __ballot_sync(0xffffffffu, q == 1);
    const uint32_t b2 = __ballot_sync(0xffffffffu, q == 2), b3 = __ballot_sync(0xffffffffu, q == 3);
    if (q == 0) dst[o0 + __popc(b0 & lt)] = k;
    else if (q == 1) dst[o1 + __popc(b1 & lt)] = k;
    else if (q == 2) dst[o2 + __popc(b2 & lt)] = k;
    else if (q == 3) dst[o3 + __popc(b3 & lt)] = k;
    o0 += __popc(b0); o1 += __popc(b1); o2 += __popc(b2); o3 += __popc(b3);
}
// src[lo, hi) in groups of kPartU blocks: fn(blockStart, key, valid) for every block of 32, in order
template <class Fn>
__device__ __forceinline__ void for_key_groups(const uint64_t* src, int lo, int hi, Fn fn) {
    const int lane = threadIdx.x & 31;
    for (int base = lo; base < hi; base += 32 * kPartU) {
        uint64_t k[kPartU];
#pragma unroll
        for (int u = 0; u < kPartU; ++u) {
            const int i = base + 32 * u + lane;
            k[u] = i < hi ? src[i] : 0;
        }
#pragma unroll
        for (int u = 0; u < kPartU; ++u) {
            if (base + 32 * u >= hi) break;                     // warp-uniform
            fn(k[u], base + 32 * u + lane < hi);
        }
    }
}
template <class F>
__device__ __forceinline__ uint4 warp_partition4(const uint64_t* src, uint64_t* dst, int beg, int cnt, int outBase, F cls) {
    const int lane = threadIdx.x & 31;
    const uint32_t lt = (1u << lane) - 1;
    int n0 = 0, n1 = 0, n2 = 0, n3 = 0;
    if (cnt <= 32 * kPartU) {
        uint64_t k[kPartU]; int q[kPartU];
#pragma unroll
        for (int u = 0; u < kPartU; ++u) {
            const int i = 32 * u + lane;
            k[u] = i < cnt ? src[beg + i] : 0;
        }
#pragma unroll
        for (int u = 0; u < kPartU; ++u) {
            q[u] = -1;
            if (32 * u >= cnt) continue;                        // warp-uniform
            if (32 * u + lane < cnt) q[u] = cls(k[u]);
            partition_count_block<F>(q[u], n0, n1, n2, n3);
        }
        int o0 = outBase, o1 = o0 + n0, o2 = o1 + n1, o3 = o2 + n2;
#pragma unroll
        for (int u = 0; u < kPartU; ++u) {
            if (32 * u >= cnt) continue;
            partition_scatter_block(dst, k[u], q[u], lt, o0, o1, o2, o3);
        }
        return make_uint4(n0, n1, n2, n3);
    }
    for_key_groups(src, beg, beg + cnt, [&](uint64_t k, bool valid) { partition_count_block<F>(valid ? cls(k) : -1, n0, n1, n2, n3); });
    int o0 = outBase, o1 = o0 + n0, o2 = o1 + n1, o3 = o2 + n2;
    for_key_groups(src, beg, beg + cnt, [&](uint64_t k, bool valid) { partition_scatter_block(dst, k, valid ? cls(k) : -1, lt, o0, o1, o2, o3); });
    return make_uint4(n0, n1, n2, n3);
}

// Same, by the whole CTA (warps own contiguous chunks).  All threads must call it; ends with a barrier.
template <class F>
__device__ uint4 cta_partition4(const uint64_t* src, uint64_t* dst, int beg, int cnt, int outBase, F cls,
                                int (*sCnt)[4] /* [kTreeWarps+1][4] */) {
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const uint32_t lt = (1u << lane) - 1;
    int chunk = (cnt + kTreeWarps - 1) / kTreeWarps;
    chunk = (chunk + 31) & ~31;
    const int cb = min(wid * chunk, cnt), ce = min(cb + chunk, cnt);
    int n0 = 0, n1 = 0, n2 = 0, n3 = 0;
    for_key_groups(src, beg + cb, beg + ce, [&](uint64_t k, bool valid) { partition_count_block<F>(valid ? cls(k) : -1, n0, n1, n2, n3); });
    if (lane == 0) { sCnt[wid][0] = n0; sCnt[wid][1] = n1; sCnt[wid][2] = n2; sCnt[wid][3] = n3; }
    __syncthreads();
    int t0 = 0, t1 = 0, t2 = 0, t3 = 0, p0 = 0, p1 = 0, p2 = 0, p3 = 0;
    for (int w = 0; w < kTreeWarps; ++w) {
        if (w == wid) { p0 = t0; p1 = t1; p2 = t2; p3 = t3; }
        t0 += sCnt[w][0]; t1 += sCnt[w][1]; t2 += sCnt[w][2]; t3 += sCnt[w][3];
    }
    int o0 = outBase + p0, o1 = outBase + t0 + p1, o2 = outBase + t0 + t1 + p2, o3 = outBase + t0 + t1 + t2 + p3;
    for_key_groups(src, beg + cb, beg + ce, [&](uint64_t k, bool valid) { partition_scatter_block(dst, k, valid ? cls(k) : -1, lt, o0, o1, o2, o3); });
    __syncthreads();
    return make_uint4(t0, t1, t2, t3);
}

__device__ __forceinline__ int nonempty4(uint4 c) { return (c.x > 0) + (c.y > 0) + (c.z > 0) + (c.w > 0); }
__device__ __forceinline__ int splittable4(uint4 c) { return (c.x > 1) + (c.y > 1) + (c.z > 1) + (c.w > 1); }

// Split points of DivideNode: halfX = ceil(float(UR.x-UL.x)/2) (:482-483)
__device__ __forceinline__ QuadClass split_of(ushort4 r) {
    QuadClass q;
    q.mx = r.x + (((int)r.z - (int)r.x + 1) >> 1);
    q.my = r.y + (((int)r.w - (int)r.y + 1) >> 1);
    return q;
}

// Partition every node listed in proc[0..T) (list positions of pool `cur`) into its four children.
// Big nodes by the whole CTA, the rest one warp each.  warpBegin lets warp 0 be busy elsewhere (sorting).
__device__ void split_nodes(TreeSmem& S, int cur, int T, uint64_t* const (&buf)[2], int (*sCnt)[4], int firstWarp) {
    const int wid = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (wid >= firstWarp) {
        for (int t = wid - firstWarp; t < T; t += kTreeWarps - firstWarp) {
            const int node = S.proc[t];
            const int cnt = (int)S.cnt[cur][node];
            if (cnt >= kCtaPartitionMin) continue;
            const uint32_t bf = S.beg[cur][node];
            const int beg = (int)(bf & 0x7fffffffu), flag = (int)(bf >> 31);
            const uint4 c = warp_partition4(buf[flag], buf[flag ^ 1], beg, cnt, beg, split_of(S.rect[cur][node]));
            if (lane == 0) S.cc[node] = c;
        }
    }
}
__device__ void split_big_nodes(TreeSmem& S, int cur, int T, uint64_t* const (&buf)[2], int (*sCnt)[4]) {
    // Big nodes exist only in the first two or three passes; the list of them is gathered first (order is irrelevant: the
    // nodes are disjoint) so that the CTA does not walk all T entries in lock step just to skip them.
    __shared__ int sBig[32];
    __shared__ int sNBig;
    if (threadIdx.x == 0) sNBig = 0;
    __syncthreads();
    for (int t = threadIdx.x; t < T; t += kTreeThreads)
        if ((int)S.cnt[cur][S.proc[t]] >= kCtaPartitionMin) { const int at = atomicAdd(&sNBig, 1); if (at < 32) sBig[at] = t; }
    __syncthreads();
    const int nBig = sNBig;
    if (nBig == 0) return;
    if (nBig > 32) {                 // more than the list holds (n >= 33 * 1024 keys in one level): the plain walk
        for (int t = 0; t < T; ++t) {
            const int node = S.proc[t];
            const int cnt = (int)S.cnt[cur][node];
            if (cnt < kCtaPartitionMin) continue;        // uniform across the CTA
            const uint32_t bf = S.beg[cur][node];
            const int beg = (int)(bf & 0x7fffffffu), flag = (int)(bf >> 31);
            const uint4 c = cta_partition4(buf[flag], buf[flag ^ 1], beg, cnt, beg, split_of(S.rect[cur][node]), sCnt);
            if (threadIdx.x == 0) S.cc[node] = c;
        }
        return;
    }
    for (int i = 0; i < nBig; ++i) {
        const int node = S.proc[sBig[i]];
        const int cnt = (int)S.cnt[cur][node];
        const uint32_t bf = S.beg[cur][node];
        const int beg = (int)(bf & 0x7fffffffu), flag = (int)(bf >> 31);
        const uint4 c = cta_partition4(buf[flag], buf[flag ^ 1], beg, cnt, beg, split_of(S.rect[cur][node]), sCnt);
        if (threadIdx.x == 0) S.cc[node] = c;
    }
}

// Rebuild the list after the nodes proc[0..T) (in processing order) were split.  Returns new size; *mOut = number
// of splittable children (the new pending list, in the reference's push_back order).
template <bool kPyr>
__device__ int apply_splits(TreeSmem& S, int cur, int listSize, int T, int* sWarp, int* mOut) {
    const int tid = threadIdx.x;
    const int nxt = cur ^ 1;
    for (int i = tid; i < listSize; i += kTreeThreads) S.dv[i] = 0;
    __syncthreads();
    for (int t = tid; t < T; t += kTreeThreads) {
        const int node = S.proc[t];
        S.dv[node] = 1;
        const uint4 c = S.cc[node];
        S.scanA[t] = nonempty4(c);
        S.scanB[t] = splittable4(c);
    }
    __syncthreads();
    const int neTot = block_exscan(S.scanA, T, sWarp);
    const int exTot = block_exscan(S.scanB, T, sWarp);
    for (int t = tid; t < T; t += kTreeThreads) {
        const int node = S.proc[t];
        const ushort4 r = S.rect[cur][node];
        const QuadClass sp = split_of(r);
        const uint32_t bf = S.beg[cur][node];
        uint32_t b = bf & 0x7fffffffu;
        const uint32_t childFlag = ((bf >> 31) ^ 1u) << 31;
        const uint4 c = S.cc[node];
        const uint32_t cn[4] = {c.x, c.y, c.z, c.w};
        const ushort4 cr[4] = {make_ushort4(r.x, r.y, sp.mx, sp.my), make_ushort4(sp.mx, r.y, r.z, sp.my),
                               make_ushort4(r.x, sp.my, sp.mx, r.w), make_ushort4(sp.mx, sp.my, r.z, r.w)};
        int rl = 0, el = 0;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            if (cn[q] > 0) {
                const int idx = neTot - 1 - (S.scanA[t] + rl);
                S.rect[nxt][idx] = cr[q];
                if (kPyr) S.beg[nxt][idx] = pyr_addr((int)(bf >> 28), (int)((bf >> 24) & 15) + 1, 2 * (int)(bf & 0xfff) + (q & 1), 2 * (int)((bf >> 12) & 0xfff) + (q >> 1));
                else S.beg[nxt][idx] = b | childFlag;
                S.cnt[nxt][idx] = cn[q];
                if (cn[q] > 1) {
                    ise::Elem e;
                    e.key = ((uint64_t)cn[q] << 32) | (uint32_t)cr[q].x;
                    e.id = idx; e.pad = 0;
                    S.pend[S.scanB[t] + el] = e;
                    ++el;
                }
                ++rl;
            }
            b += cn[q];
        }
    }
    __syncthreads();
    for (int i = tid; i < listSize; i += kTreeThreads) S.scanA[i] = S.dv[i] ? 0 : 1;
    __syncthreads();
    block_exscan(S.scanA, listSize, sWarp);
    for (int i = tid; i < listSize; i += kTreeThreads) {
        if (S.dv[i]) continue;
        const int idx = neTot + S.scanA[i];
        S.rect[nxt][idx] = S.rect[cur][i];
        S.beg[nxt][idx] = S.beg[cur][i];
        S.cnt[nxt][idx] = S.cnt[cur][i];
    }
    __syncthreads();
    *mOut = exTot;
    return neTot + listSize - T;
}

// std::sort emulation by ONE WARP (see introsort_emul.h: the ranges of the introsort loop are independent work items and the
// final insertion pass is a stable insertion sort of each leaf range).  Rounds: lane i partitions the i-th active range
// (> 16 elements) once; the children go to the next round's list or, when small, to the leaf list; slots are handed out
// with ballots, so no atomics and no ordering assumptions.  listA / listB: 2 ints per active range (first | last << 16,
// depth), at most n / 17 ranges each; leaves: 1 int per leaf, at most n.  Critical path ~2n element steps instead of
// ~n log2(n / 16) + the insertion pass of a single thread.
__device__ void warp_sort_nodes(ise::Elem* e, int n, uint32_t* listA, uint32_t* listB, uint32_t* leaves) {
    const int lane = threadIdx.x & 31;
    const unsigned lower = (1u << lane) - 1;
    if (n <= 16) { if (lane == 0) ise::insertion_sort(e, 0, n); __syncwarp(); return; }
    int lg = 0;
    for (int t = n; t > 1; t >>= 1) ++lg;
    if (lane == 0) { listA[0] = (uint32_t)n << 16; listA[1] = 2 * lg; }
    __syncwarp();
    uint32_t* cur = listA; uint32_t* nxt = listB;
    int nA = 1, nLeaf = 0;
    while (nA > 0) {
        int nN = 0;
        for (int base = 0; base < nA; base += 32) {
            const int i = base + lane;
            uint32_t c0 = 0, c1 = 0, d = 0;
            bool has = false;
            if (i < nA) {
                const uint32_t r = cur[2 * i];
                const int first = (int)(r & 0xffffu), last = (int)(r >> 16);
                d = cur[2 * i + 1];
                if (d == 0) {
                    ise::heap_sort(e, first, last);                 // depth limit reached: std::__partial_sort of the range
                } else {
                    --d;
                    const int cut = ise::split_step(e, first, last);
                    c0 = (uint32_t)first | ((uint32_t)cut << 16);
                    c1 = (uint32_t)cut | ((uint32_t)last << 16);
                    has = true;
                }
            }
            const int s0 = (int)(c0 >> 16) - (int)(c0 & 0xffffu), s1 = (int)(c1 >> 16) - (int)(c1 & 0xffffu);
            const bool a0 = has && s0 > 16, a1 = has && s1 > 16;
            const bool l0 = has && !a0 && s0 > 1, l1 = has && !a1 && s1 > 1;
            const unsigned mA0 = __ballot_sync(0xffffffffu, a0), mA1 = __ballot_sync(0xffffffffu, a1);
            const unsigned mL0 = __ballot_sync(0xffffffffu, l0), mL1 = __ballot_sync(0xffffffffu, l1);
            if (a0) { const int s = nN + __popc(mA0 & lower); nxt[2 * s] = c0; nxt[2 * s + 1] = d; }
            if (a1) { const int s = nN + __popc(mA0) + __popc(mA1 & lower); nxt[2 * s] = c1; nxt[2 * s + 1] = d; }
            nN += __popc(mA0) + __popc(mA1);
            if (l0) leaves[nLeaf + __popc(mL0 & lower)] = c0;
            if (l1) leaves[nLeaf + __popc(mL0) + __popc(mL1 & lower)] = c1;
            nLeaf += __popc(mL0) + __popc(mL1);
        }
        __syncwarp();
        uint32_t* t = cur; cur = nxt; nxt = t;
        nA = nN;
    }
    for (int i = lane; i < nLeaf; i += 32) {
        const uint32_t r = leaves[i];
        ise::insertion_sort(e, (int)(r & 0xffffu), (int)(r >> 16));
    }
    __syncwarp();
}

#ifdef ORBX_TREE_PROFILE
__device__ long long g_treeProf[16];
#define TPROF_T0() long long tp_ = clock64()
#define TPROF(i) do { __syncthreads(); if (threadIdx.x == 0 && blockIdx.x == 0 && blockIdx.y == 0) { long long n_ = clock64(); g_treeProf[i] += n_ - tp_; tp_ = n_; } } while (0)
#else
#define TPROF_T0()
#define TPROF(i)
#endif

// The whole DistributeOctTree for one level.  Keys are in buf[0][0..n) in the reference's vToDistributeKeys order.
// Writes the kept keys (one per final node, list order) to out[0..*outCount).
// kPyr: the count-pyramid form (PV filled by pyr_insert / pyr_reduce; buf is not touched); returns false when a node at the
// pyramid's finest depth would have to be split - nothing has been written then and the caller runs the partitioning form.
template <bool kPyr>
__device__ bool distribute_tree(uint64_t* const (&buf)[2], int n, int N, int nIni, float hX, int treeH,
                                TreeSmem& S, uint64_t* out, int* outCount, const PyrView PV = PyrView{nullptr, nullptr}) {
    __shared__ int sWarp[33];
    __shared__ int sCnt[kTreeMaxWarps + 1][4];
    __shared__ int sT;
    __shared__ int sAbort;
    const int tid = threadIdx.x, wid = tid >> 5, lane = tid & 31;
    if (n == 0) { if (tid == 0) *outCount = 0; return true; }
    TPROF_T0();

    if (kPyr) {
        if (tid == 0) sAbort = 0;
        if (tid < nIni) S.scanB[tid] = (int)PV.cnt[tid * kPyrCellsPerRoot];
    } else {
        // roots: stable grouping of the keys by (int)(x / hX) into buf[1]
        int outBase = 0;
        for (int gb = 0; gb < nIni; gb += 4) {
            RootClass rc; rc.hX = hX; rc.base = gb;
            const uint4 c = cta_partition4(buf[0], buf[1], 0, n, outBase, rc, sCnt);
            if (tid == 0) { S.scanB[gb] = c.x; S.scanB[gb + 1] = c.y; S.scanB[gb + 2] = c.z; S.scanB[gb + 3] = c.w; }
            outBase += c.x + c.y + c.z + c.w;
        }
    }
    __syncthreads();
    int cur = 0;
    if (tid == 0) {
        int sz = 0, b = 0;
        for (int i = 0; i < nIni; ++i) {
            const int c = S.scanB[i];
            if (c > 0) {
                S.rect[0][sz] = make_ushort4((unsigned short)(int)__fmul_rn(hX, (float)i), 0,
                                             (unsigned short)(int)__fmul_rn(hX, (float)(i + 1)), (unsigned short)treeH);
                S.beg[0][sz] = kPyr ? pyr_addr(i, 0, 0, 0) : ((uint32_t)b | 0x80000000u);
                S.cnt[0][sz] = c;
                ++sz;
            }
            b += c;
        }
        sT = sz;
    }
    __syncthreads();
    int listSize = sT;
    TPROF(1);

    bool finish = false;
    while (!finish) {
        const int prevSize = listSize;
        // all splittable nodes in list order
        for (int i = tid; i < listSize; i += kTreeThreads) S.scanA[i] = S.cnt[cur][i] > 1 ? 1 : 0;
        __syncthreads();
        const int T = block_exscan(S.scanA, listSize, sWarp);
        for (int i = tid; i < listSize; i += kTreeThreads)
            if (S.cnt[cur][i] > 1) S.proc[S.scanA[i]] = (uint16_t)i;
        __syncthreads();
        TPROF(2);
        if (kPyr) {
            pyr_children(S, cur, T, PV, 0, &sAbort);
            __syncthreads();
            if (sAbort) return false;
        } else {
            split_big_nodes(S, cur, T, buf, sCnt);
            TPROF(3);
            split_nodes(S, cur, T, buf, sCnt, 0);
            __syncthreads();
        }
        TPROF(4);
        int m;
        listSize = apply_splits<kPyr>(S, cur, listSize, T, sWarp, &m);
        cur ^= 1;
        TPROF(5);
        if (listSize >= N || listSize == prevSize) {
            finish = true;
        } else if (listSize + 3 * m > N) {
            while (!finish) {
                const int prev2 = listSize;
                for (int e = tid; e < m; e += kTreeThreads) S.proc[e] = (uint16_t)S.pend[e].id;
                __syncthreads();
                // std::sort emulation by warp 0 while the other warps pre-split every pending node
                if (wid == 0) warp_sort_nodes(S.pend, m, reinterpret_cast<uint32_t*>(S.scanB), reinterpret_cast<uint32_t*>(S.scanB) + (S.scanCap >> 1),
                                              reinterpret_cast<uint32_t*>(S.scanA));
                if (kPyr) {
                    pyr_children(S, cur, m, PV, 32, &sAbort);
                    __syncthreads();
                    if (sAbort) return false;
                } else {
                    split_nodes(S, cur, m, buf, sCnt, 1);
                    __syncthreads();
                    split_big_nodes(S, cur, m, buf, sCnt);
                }
                // processing order = sorted order from the back; stop as soon as the list reaches N
                for (int t = tid; t < m; t += kTreeThreads) {
                    const int node = S.pend[m - 1 - t].id;
                    S.proc[t] = (uint16_t)node;
                    S.scanA[t] = nonempty4(S.cc[node]) - 1;
                }
                if (tid == 0) sT = m;
                __syncthreads();
                block_exscan(S.scanA, m, sWarp);
                for (int t = tid; t < m; t += kTreeThreads) {
                    const int v = nonempty4(S.cc[S.proc[t]]) - 1;
                    if (listSize + S.scanA[t] + v >= N) atomicMin(&sT, t + 1);
                }
                __syncthreads();
                const int T2 = sT;
                __syncthreads();
                listSize = apply_splits<kPyr>(S, cur, listSize, T2, sWarp, &m);
                cur ^= 1;
                TPROF(7);
                if (listSize >= N || listSize == prev2) finish = true;
            }
        }
    }

    // keep the best response of every node; first maximum wins (:757-776)
    if (kPyr) {
        for (int i = tid; i < listSize; i += kTreeThreads) out[i] = buf[0][pyr_best_pos(PV.best[pyr_cell(S.beg[cur][i])])];
        if (tid == 0) *outCount = listSize;
        TPROF(8);
        return true;
    }
    for (int i = wid; i < listSize; i += kTreeWarps) {
        const uint32_t bf = S.beg[cur][i];
        const uint64_t* seg = buf[bf >> 31] + (bf & 0x7fffffffu);
        const int cnt = (int)S.cnt[cur][i];
        // order by (score desc, position asc): maximise score<<32 | ~pos
        uint64_t best = 0; uint64_t bestKey = 0;      // every real candidate has ord > 0
        for (int k = lane; k < cnt; k += 32) {
            const uint64_t key = seg[k];
            const uint64_t ord = ((key >> 32) << 32) | (uint32_t)(0x7fffffff - k);
            if (ord > best) { best = ord; bestKey = key; }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const uint64_t ob = __shfl_xor_sync(0xffffffffu, best, o);
            const uint64_t ok = __shfl_xor_sync(0xffffffffu, bestKey, o);
            if (ob > best) { best = ob; bestKey = ok; }
        }
        if (lane == 0) out[i] = bestKey;
    }
    if (tid == 0) *outCount = listSize;
    TPROF(8);
    return true;
}

}  // namespace orbx
