// Internal layout of the B200-native ORB extractor (not part of the C ABI).
//
// HBM layout per handle (B = max_batch frames in flight; everything is allocated once per image shape):
//   lvl[l]      B planes of level l, unpadded, row pitch = align128(w_l).  Level 0 is the input image itself
//               (copied in on the host path, aliased on the device-resident path).
//   blur[l]     B planes, same pitch: 7x7 fixed-point Gaussian of level l (the reference's per-level workingMat).
//   score[l]    B planes, same pitch: FAST strength of every tested pixel (k_fast_score); cbits: one 'strength > minThFAST'
//               bit per pixel, B x bitsPerFrame uint2 (layout in orbx_kernels_fast.cuh).
//   cand        B x candCap packed FAST candidates (u32: x_local | y_local<<8 | score<<16 inside the owning cell),
//               one fixed worst-case slot range per cell; cellCount B x nCells.
//   qtA/qtB     B x sum_l candCapLevel[l] u64 quadtree keys (x | y<<16 | score<<32), ping-pong buffers.
//   lvlKp       B x L x kpCapLevel kept keypoints of a level (u64 keys, list order); lvlCnt B x L.
//   outKp/outDesc/outCnt  B x outCap orbx_keypoint (28 B), B x outCap x 32 B, B x 2 ints (n, monoIndex).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <map>
#include <mutex>
#include <string>
#include <utility>
#include <vector>
#include "../../include/orbx.h"

namespace orbx {

constexpr int kEdge = 19;          // EDGE_THRESHOLD (src/ORBextractor.cc:73)
constexpr int kMinBorder = 16;     // EDGE_THRESHOLD-3 (:789)
constexpr int kHalfPatch = 15;
constexpr int kMaxLevels = 16;

struct LevelGeom {
    int w, h, pitch;               // unpadded size, row pitch in bytes
    size_t planeBytes;             // pitch*h rounded to 256
    int cellBegin, cellEnd;        // range in the cell table
    int nCols, nRows, wCell, hCell;
    int quota;                     // mnFeaturesPerLevel[l]
    int candCap;                   // worst-case number of FAST candidates of this level
    size_t qtOffset;               // offset (elements) of this level inside a frame's quadtree buffer
    int kpCap;                     // capacity of the kept list
    int kpOffset;                  // offset (elements) inside a frame's lvlKp
    int nIni;                      // quadtree roots
    float hX;
    int treeW, treeH;              // maxBorder-minBorder
    int poolCap;                   // node pool capacity
    float scale;                   // mvScaleFactor[l]
    float sizeF;                   // (float)(int)(31*scale)
    int tabX, tabY;                // offsets into the resize tables
    int areaFast;                  // exact 2x decimation -> INTER_AREA path
};

struct FastCell {                  // one reference FAST cell (src/ORBextractor.cc:805-868), tested region only
    short level;
    short x0, y0;                  // first tested pixel (level coords) = ini + 3
    short w, h;                    // tested extent
    short relX, relY;              // j*wCell+3, i*hCell+3: add to local coords to get coords relative to minBorder
    short pad;
    int outOffset;                 // slot range in cand
    int cap;
};

struct ResizeTap { int ofs; short c0, c1; };   // source index + 11-bit coefficients

// parameters passed by value to kernels
struct DevGeom {
    int nlevels;
    int nCells;
    int candCapFrame;              // sum of cell caps
    size_t qtCapFrame;
    int kpCapFrame;
    int outCap;
    int iniTh, minTh;
    LevelGeom lv[kMaxLevels];
    uint8_t* lvl[kMaxLevels];      // base of B planes
    uint8_t* blur[kMaxLevels];
    uint8_t* score[kMaxLevels];    // FAST strength of every pixel (k_fast_score), same pitch as the level
};

// cudaFuncAttributeMaxDynamicSharedMemorySize belongs to a (function, device) pair, not to a handle: two extractors with
// different nfeatures on one GPU (Tracking's 5x initialisation extractor beside the regular one, src/Tracking.cc:591-606), or
// agents sharing a GPU, would lower each other's limit if every handle wrote its own size.  The limit is therefore only
// ever RAISED, per device, under a lock; the device must be current.
inline cudaError_t raise_dyn_smem(const void* kern, int device, size_t bytes) {
    static std::mutex mu;
    static std::map<std::pair<const void*, int>, size_t> cur;
    std::lock_guard<std::mutex> lk(mu);
    size_t& c = cur[std::make_pair(kern, device)];
    if (bytes <= c) return cudaSuccess;
    const cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
    if (e == cudaSuccess) c = bytes;
    return e;
}

// Launch with a scheduling priority (cudaLaunchAttributePriority; captured into graph kernel nodes as well).  The short,
// latency-bound kernels (quadtree, finalize, the tracking step) are given the highest priority so that their CTAs are
// dispatched ahead of the tens of thousands of queued CTAs of the throughput kernels (FAST, blur) of ANOTHER handle working
// on the same GPU; prio = 0 is a plain launch.
template <class... KArgs, class... Args>
inline cudaError_t launch_prio(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, int prio, Args&&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributePriority;
    at[0].val.priority = prio;
    cfg.attrs = at; cfg.numAttrs = prio ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
}

}  // namespace orbx
