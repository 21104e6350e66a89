// E2, second generation: grid-cell FAST as two passes (reference behaviour: src/ORBextractor.cc:781-872; pixel arithmetic
// = cv::FAST 9/16 with 3x3 non-max suppression, SURVEY.md A3, restated in oracle/cpp/orc_primitives.h).
//
//   k_fast_score   image-wide, regular: the FAST strength of EVERY pixel of a level's tested region, written as one byte
//                  per pixel into a score plane, plus one "strength > minThFAST" bit per pixel.
//                  The tested regions of the reference's cells tile the level without overlap (cell j tests the columns
//                  [16 + j*wCell + 3, 16 + (j+1)*wCell + 3)), and a pixel's strength does not depend on its cell, so this
//                  pass ignores the cell grid.  What IS per cell - neighbours outside the cell count as 0 in the 3x3
//                  non-max suppression, the iniThFAST -> minThFAST fallback, the raster output order - happens in
//   k_fast_select  one warp per reference cell: gathers the cell's corner bits, runs the (sparse) non-max suppression
//                  on the score plane with the cell's own border masks, applies the ini / min rule and writes the
//                  survivors in raster order (same output format as k_fast_cells: cand + cellCount).
//
// k_fast_score is the hot kernel (73-instruction u16x2 min/max network per pixel pair, fast_strength_x2).  Design:
//   * a WARP owns a 64-px x 28-row patch: lanes 0-15 take the upper 14-row strip, lanes 16-31 the lower one, each lane
//     4 adjacent pixels.  Its 96 x (28+6) input tile is staged by ONE 3-D TMA tile load (cp.async.bulk.tensor,
//     tensor = {x, y, frame}, zero fill outside the level) signalled on a per-warp mbarrier; no block barrier anywhere.
//     Rows are 96 bytes = 24 words, so the two half-warps (14 rows apart = 336 words = 16 mod 32) read disjoint banks:
//     every LDS is conflict free.
//   * a lane walks DOWN its 14 rows with a 7-row sliding window held in registers.  A new row costs 3 LDS + 10 ALU ops:
//     the 10 bytes x-3 .. x+6 are cut into the eight u16x2 operands G(d) = (p[x+d], p[x+d+2]), d = -3..4.  With pixel
//     lanes (x, x+2) and (x+1, x+3) every ring operand of the first pair is a G(d) and of the second pair G(d+1): no
//     per-pixel address arithmetic, no shared-memory expansion pass, no ring loads (the first generation spent 17 LDS
//     + ~55 bookkeeping instructions per pixel pair around the 73-instruction network).
//   * the strengths of 4 pixels leave as one aligned 32-bit store; the corner bits accumulate in two registers per lane
//     and leave as one 64-bit store per strip.
// Layout of the corner bits: cbits[frame][level][strip s][column group q] (uint2), column group q = pixels 16+4q .. +3,
// strip s = rows 19+14s .. +13; .x bit r = pixel 0 row r, bit 16+r = pixel 2; .y bit r = pixel 1, bit 16+r = pixel 3.
#pragma once
#include "orbx_kernels_desc.cuh"
#include <cuda.h>          // CUtensorMap (type only; the encoder is fetched through cudaGetDriverEntryPoint)

namespace orbx {

constexpr int kFsRows = 14;                      // rows per strip (one half-warp)
constexpr int kFsCols = 64;                      // pixels per strip row (16 lanes x 4)
constexpr int kFsHaloX = 16;                     // the tile starts 16 columns left of the patch: TMA wants the box start 16-byte aligned
                                                 // in the innermost dimension (x0 - 16 = 64 * cx), the ring needs only 3
constexpr int kFsBoxW = 96;                      // staged bytes per tile row: columns x0-16 .. x0+79 (x0-3 .. x0+66 used)
constexpr int kFsBoxH = 2 * kFsRows + 6;         // 34 tile rows: two strips + the 3-row ring halo on both sides
constexpr int kFsTileBytes = kFsBoxW * kFsBoxH;  // 3264
constexpr int kFsTileStride = 3328;              // 128-byte multiple (TMA destination alignment)
constexpr int kFsWarps = 4;
#ifndef ORBX_FS_MINBLOCKS
#define ORBX_FS_MINBLOCKS 5            // CTAs per SM the strength kernel is compiled for (96 registers; 4 -> 127 registers, no spill)
#endif
constexpr int kFsX0 = 16, kFsY0 = 19;            // first column group starts at x = 16 (aligned stores); first tested row

struct FsLevel { int jobBegin, nCx, nSy, bitsOffset; };          // bitsOffset: in uint2 units inside a frame's cbits
struct FsPlan { int nJobs, bitsPerFrame; FsLevel lv[kMaxLevels]; };
struct alignas(64) FsMaps { CUtensorMap m[kMaxLevels]; int tmaOk[kMaxLevels]; };

__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// the eight u16x2 operands of one tile row for a lane: o[i] = (p[x+d], p[x+d+2]), d = i-3, from the three words that hold
// the bytes x-4 .. x+7
__device__ __forceinline__ void fs_expand(uint32_t (&o)[8], const uint32_t* __restrict__ rp) {
    const uint32_t w0 = rp[0], w1 = rp[1], w2 = rp[2];
    const uint32_t t = __funnelshift_r(w0, w1, 16), t2 = __funnelshift_r(w1, w2, 16);
    o[0] = __byte_perm(w0, 0, 0x4341);
    o[1] = t & 0x00ff00ffu;
    o[2] = __byte_perm(t, 0, 0x4341);
    o[3] = w1 & 0x00ff00ffu;
    o[4] = __byte_perm(w1, 0, 0x4341);
    o[5] = t2 & 0x00ff00ffu;
    o[6] = __byte_perm(t2, 0, 0x4341);
    o[7] = w2 & 0x00ff00ffu;
}

// Upper bound of the strength from the four compass pixels only: every 9-arc of the ring contains r0 or r8 AND r4 or
// r12, so min over an arc <= min(max(r0,r8), max(r4,r12)) (bright side; mirrored for the dark side).  A pixel whose bound
// is <= minThFAST is no corner at either threshold and can never win a non-max comparison against one, so its stored
// strength may be anything <= minThFAST: 0 is written and the 73-instruction network is skipped (exact).
__device__ __forceinline__ uint32_t fs_compass_bound(uint32_t v2, uint32_t r0, uint32_t r4, uint32_t r8, uint32_t r12) {
    const uint32_t A = __vminu2(__vmaxu2(r0, r8), __vmaxu2(r4, r12));
    const uint32_t B = __vmaxu2(__vminu2(r0, r8), __vminu2(r4, r12));
    return __vimax3_u16x2(A + (0x01000100u - v2), (v2 + 0x01000100u) - B, 0x01000100u);      // biased by 256 per lane
}

// ring position k = (dx, dy), index 0 = (0, +3), in the order cv::FAST walks the circle
#define FS_RING_DX(k) ((k) == 0 ? 0 : (k) == 1 ? 1 : (k) == 2 ? 2 : (k) == 3 ? 3 : (k) == 4 ? 3 : (k) == 5 ? 3 : (k) == 6 ? 2 : (k) == 7 ? 1 : \
                       (k) == 8 ? 0 : (k) == 9 ? -1 : (k) == 10 ? -2 : (k) == 11 ? -3 : (k) == 12 ? -3 : (k) == 13 ? -3 : (k) == 14 ? -2 : -1)
#define FS_RING_DY(k) ((k) == 0 ? 3 : (k) == 1 ? 3 : (k) == 2 ? 2 : (k) == 3 ? 1 : (k) == 4 ? 0 : (k) == 5 ? -1 : (k) == 6 ? -2 : (k) == 7 ? -3 : \
                       (k) == 8 ? -3 : (k) == 9 ? -3 : (k) == 10 ? -2 : (k) == 11 ? -1 : (k) == 12 ? 0 : (k) == 13 ? 1 : (k) == 14 ? 2 : 3)

// grid (ceil(nJobs / 4), B), block 128 = 4 independent warps.  kEarly: warp-uniform compass early-out.
template <bool kEarly>
__global__ void __launch_bounds__(kFsWarps * 32, ORBX_FS_MINBLOCKS) k_fast_score(const __grid_constant__ FsMaps maps, const DevGeom* __restrict__ g,
                                                               Level0 l0, const FsPlan plan, int frameBase) {
    __shared__ __align__(128) uint8_t tiles[kFsWarps][kFsTileStride];
    __shared__ __align__(8) uint64_t mbar[kFsWarps];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int job = blockIdx.x * kFsWarps + warp;
    if (job >= plan.nJobs) return;                      // warps are independent: no block-wide barrier below
    int l = 0;
    while (l + 1 < g->nlevels && job >= plan.lv[l + 1].jobBegin) ++l;
    const FsLevel fl = plan.lv[l];
    const int jl = job - fl.jobBegin;
    const int sy = jl / fl.nCx, cx = jl - sy * fl.nCx;
    const int frame = frameBase + blockIdx.y;
    const LevelGeom& L = g->lv[l];
    const int x0 = kFsX0 + kFsCols * cx, y0 = kFsY0 + 2 * kFsRows * sy;
    uint8_t* tile = tiles[warp];

    if (maps.tmaOk[l]) {
        const uint32_t bar = smem_addr(&mbar[warp]);
        if (lane == 0) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar));
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(kFsTileBytes) : "memory");
            asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
                         :: "r"(smem_addr(tile)), "l"(reinterpret_cast<uint64_t>(&maps.m[l])), "r"(bar), "r"(x0 - kFsHaloX), "r"(y0 - 3), "r"(frame)
                         : "memory");
        }
        __syncwarp();
        uint32_t done = 0;
        while (!done)
            asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                         : "=r"(done) : "r"(bar), "r"(0) : "memory");
    } else {
        // rows that TMA cannot address (base, pitch or frame size not a multiple of 16 bytes: only a caller's level 0):
        // the same tile, byte by byte, zero outside the level
        int pitch;
        const uint8_t* img = level_ptr(g, l0, l, frame, &pitch);
        for (int i = lane; i < kFsBoxH * (kFsBoxW / 4); i += 32) {
            const int ry = i / (kFsBoxW / 4), rw = i - ry * (kFsBoxW / 4);
            const int y = y0 - 3 + ry, x = x0 - kFsHaloX + 4 * rw;
            uint32_t v = 0;
            if (y < L.h) {
                const uint8_t* rowp = img + (size_t)y * pitch;
#pragma unroll
                for (int k = 0; k < 4; ++k) if (x + k < L.w) v |= (uint32_t)rowp[x + k] << (8 * k);
            }
            reinterpret_cast<uint32_t*>(tile)[i] = v;
        }
        __syncwarp();
    }

    const int half = lane >> 4, j = lane & 15;
    const uint32_t* trow = reinterpret_cast<const uint32_t*>(tile) + half * kFsRows * (kFsBoxW / 4) + (kFsHaloX / 4 - 1) + j;
    const int x = x0 + 4 * j;
    const int pitch = L.pitch;
    const int ys = y0 + half * kFsRows;
    uint8_t* sp = g->score[l] + (size_t)frame * L.planeBytes + (size_t)ys * pitch + x;
    const int rowsOk = x < pitch ? L.h - ys : 0;        // rows of this lane that lie inside the plane
    // strengths stay biased by 256 per u16 lane (s + 256 in [256, 511]): the low byte IS the strength, and
    // s > minTh  <=>  bit 15 of (s + 256) + (0x7eff - minTh)
    const uint32_t cmp = (0x7effu - (uint32_t)g->minTh) * 0x00010001u;
    (void)cmp;

    uint32_t G[7][8];
#pragma unroll
    for (int k = 0; k < 6; ++k) fs_expand(G[k], trow + k * (kFsBoxW / 4));
#pragma unroll 1
    for (int it = 0; it < kFsRows / 7; ++it) {
#pragma unroll
        for (int k = 0; k < 7; ++k) {
            // output row c = 7*it + k uses tile rows c .. c+6, held in window slots (k + dy + 3) % 7
            fs_expand(G[(k + 6) % 7], trow + (7 * it + k + 6) * (kFsBoxW / 4));
            const uint32_t vA = G[(k + 3) % 7][3], vB = G[(k + 3) % 7][4];
            bool run = true;
            if (kEarly) {
                const uint32_t bA = fs_compass_bound(vA, G[(k + 6) % 7][3], G[(k + 3) % 7][6], G[k % 7][3], G[(k + 3) % 7][0]);
                const uint32_t bB = fs_compass_bound(vB, G[(k + 6) % 7][4], G[(k + 3) % 7][7], G[k % 7][4], G[(k + 3) % 7][1]);
                run = __any_sync(0xffffffffu, (((bA + cmp) | (bB + cmp)) & 0x80008000u) != 0);
            }
            uint32_t sA = 0x01000100u, sB = 0x01000100u;
            if (run) {
                uint32_t rA[16], rB[16];
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    rA[i] = G[(k + 3 + FS_RING_DY(i)) % 7][FS_RING_DX(i) + 3];
                    rB[i] = G[(k + 3 + FS_RING_DY(i)) % 7][FS_RING_DX(i) + 4];
                }
                sA = fast_strength_biased_x2(vA, rA);
                sB = fast_strength_biased_x2(vB, rB);
            }
            if (7 * it + k < rowsOk) *reinterpret_cast<uint32_t*>(sp) = __byte_perm(sA, sB, 0x6240);      // s(x), s(x+1), s(x+2), s(x+3)
            sp += pitch;
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------
// 3x3 non-max suppression, dense and SIMD, image-wide (same warp patches and lane layout as k_fast_score).
// cv::FAST keeps a corner iff its score exceeds all 8 neighbours' scores INSIDE the sub-image it was called on
// (src/ORBextractor.cc:826: one call per cell), so a neighbour that lies in another cell's tested interior counts as 0.
// The cells of a level form a regular grid, hence "same cell" separates into a column test and a row test: colTab[x] says
// whether x-1 / x+1 belong to the same cell column as x (and whether x is in any cell), rowTab[y] the same for rows, as
// ready-made AND masks.  On the synthetic stream ~80 % of the pixels exceed minThFAST, so the suppression is done for every
// pixel in u16x2 lanes (~16 instructions per pixel pair) rather than per corner; its output - one bit per SURVIVOR - is
// sparse (<= 1/4 of the pixels by construction, ~3 % here).  A survivor at minThFAST that also exceeds iniThFAST is a
// survivor at iniThFAST (the comparison is on raw strengths), so one bit plane serves both thresholds.
// Score tile: 96 x 30 bytes by one TMA tile load per warp (rows y0-1 .. y0+28).
constexpr int kNmsBoxH = 2 * kFsRows + 2;
constexpr int kNmsTileBytes = kFsBoxW * kNmsBoxH;     // 2880
constexpr int kNmsTileStride = 2944;                  // 128-byte multiple

struct NmsTabs { const uint8_t* col; const uint4* row; int colOff[kMaxLevels], rowOff[kMaxLevels]; };

__global__ void __launch_bounds__(kFsWarps * 32) k_fast_nms(const __grid_constant__ FsMaps smaps, const DevGeom* __restrict__ g,
                                                            const FsPlan plan, const NmsTabs tabs, uint2* __restrict__ sbits, int frameBase) {
    __shared__ __align__(128) uint8_t tiles[kFsWarps][kNmsTileStride];
    __shared__ __align__(8) uint64_t mbar[kFsWarps];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int job = blockIdx.x * kFsWarps + warp;
    if (job >= plan.nJobs) return;
    int l = 0;
    while (l + 1 < g->nlevels && job >= plan.lv[l + 1].jobBegin) ++l;
    const FsLevel fl = plan.lv[l];
    const int jl = job - fl.jobBegin;
    const int sy = jl / fl.nCx, cx = jl - sy * fl.nCx;
    const int frame = frameBase + blockIdx.y;
    const LevelGeom& L = g->lv[l];
    const int x0 = kFsX0 + kFsCols * cx, y0 = kFsY0 + 2 * kFsRows * sy;
    uint8_t* tile = tiles[warp];
    if (smaps.tmaOk[l]) {
        const uint32_t bar = smem_addr(&mbar[warp]);
        if (lane == 0) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar));
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(kNmsTileBytes) : "memory");
            asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
                         :: "r"(smem_addr(tile)), "l"(reinterpret_cast<uint64_t>(&smaps.m[l])), "r"(bar), "r"(x0 - kFsHaloX), "r"(y0 - 1), "r"(frame)
                         : "memory");
        }
        __syncwarp();
        uint32_t done = 0;
        while (!done)
            asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                         : "=r"(done) : "r"(bar), "r"(0) : "memory");
    } else {
        const uint8_t* S = g->score[l] + (size_t)frame * L.planeBytes;
        for (int i = lane; i < kNmsBoxH * (kFsBoxW / 4); i += 32) {
            const int ry = i / (kFsBoxW / 4), rw = i - ry * (kFsBoxW / 4);
            const int y = y0 - 1 + ry, x = x0 - kFsHaloX + 4 * rw;
            uint32_t v = 0;
            if (y < L.h && x + 3 < L.pitch) v = *reinterpret_cast<const uint32_t*>(S + (size_t)y * L.pitch + x);
            if (y < L.h && x + 3 >= L.w) {                       // the plane's padding columns are not part of the level
#pragma unroll
                for (int k = 0; k < 4; ++k) if (x + k >= L.w) v &= ~(0xffu << (8 * k));
            }
            reinterpret_cast<uint32_t*>(tile)[i] = v;
        }
        __syncwarp();
    }

    const int half = lane >> 4, j = lane & 15;
    const uint32_t* trow = reinterpret_cast<const uint32_t*>(tile) + half * kFsRows * (kFsBoxW / 4) + (kFsHaloX / 4 - 1) + j;
    const int x = x0 + 4 * j;
    const int ys = y0 + half * kFsRows;
    // column masks of this lane's four pixels: pair A = (x, x+2), pair B = (x+1, x+3)
    const uint32_t cf = *reinterpret_cast<const uint32_t*>(tabs.col + tabs.colOff[l] + x);
    const uint32_t f0 = cf & 255u, f1 = (cf >> 8) & 255u, f2 = (cf >> 16) & 255u, f3 = cf >> 24;
    const uint32_t al = (f0 & 1u ? 0x000000ffu : 0u) | (f2 & 1u ? 0x00ff0000u : 0u);
    const uint32_t ar = (f0 & 2u ? 0x000000ffu : 0u) | (f2 & 2u ? 0x00ff0000u : 0u);
    const uint32_t bl = (f1 & 1u ? 0x000000ffu : 0u) | (f3 & 1u ? 0x00ff0000u : 0u);
    const uint32_t br = (f1 & 2u ? 0x000000ffu : 0u) | (f3 & 2u ? 0x00ff0000u : 0u);
    const uint32_t inA = (f0 & 4u ? 0x00008000u : 0u) | (f2 & 4u ? 0x80000000u : 0u);
    const uint32_t inB = (f1 & 4u ? 0x00008000u : 0u) | (f3 & 4u ? 0x80000000u : 0u);
    const uint32_t cmp = (0x7fffu - (uint32_t)g->minTh) * 0x00010001u;       // s > minTh  <=>  bit 15 of s + cmp
    const uint4* rtab = tabs.row + tabs.rowOff[l] + ys;

    uint32_t hA[3], hB[3], cA[2], cB[2], lA[2], lB[2];      // rolling: row maxima of 3 rows, centre + left/right maxima of 2
    uint32_t accA = 0, accB = 0;
#pragma unroll
    for (int r = 0; r < kFsRows + 2; ++r) {
        const uint32_t* rp = trow + r * (kFsBoxW / 4);
        const uint32_t w0 = rp[0], w1 = rp[1], w2 = rp[2];
        const uint32_t Al = __funnelshift_r(w0, w1, 24) & al;          // (x-1, x+1) where they are same-cell neighbours
        const uint32_t A = w1 & 0x00ff00ffu;                            // (x, x+2)
        const uint32_t Bc = __byte_perm(w1, 0, 0x4341);                 // (x+1, x+3)
        const uint32_t Ar = Bc & ar, Bl = A & bl;
        const uint32_t Br = __funnelshift_r(w1, w2, 16) & br;           // (x+2, x+4)
        hA[r % 3] = __vimax3_u16x2(Al, A, Ar);
        hB[r % 3] = __vimax3_u16x2(Bl, Bc, Br);
        cA[r % 2] = A; cB[r % 2] = Bc;
        lA[r % 2] = __vmaxu2(Al, Ar); lB[r % 2] = __vmaxu2(Bl, Br);
        if (r >= 2) {
            // output row r-2 of the strip: centre = tile row r-1, above = r-2, below = r
            const uint4 rm = rtab[r - 2];                               // {above counts, below counts, row is inside a cell}
            const uint32_t nbA = __vimax3_u16x2(hA[(r - 2) % 3] & rm.x, hA[r % 3] & rm.y, lA[(r - 1) % 2]);
            const uint32_t nbB = __vimax3_u16x2(hB[(r - 2) % 3] & rm.x, hB[r % 3] & rm.y, lB[(r - 1) % 2]);
            const uint32_t sA = cA[(r - 1) % 2], sB = cB[(r - 1) % 2];
            const uint32_t tA = ((sA + 0x7fff7fffu) - nbA) & (sA + cmp) & (inA & rm.z);      // s > nb  &&  s > minTh  &&  inside
            const uint32_t tB = ((sB + 0x7fff7fffu) - nbB) & (sB + cmp) & (inB & rm.z);
            accA = ((accA >> 1) & 0x7fff7fffu) | tA;
            accB = ((accB >> 1) & 0x7fff7fffu) | tB;
        }
    }
    accA = (accA >> (16 - kFsRows)) & 0x3fff3fffu;                       // row c sits at bit c + 16 - kFsRows of each half
    accB = (accB >> (16 - kFsRows)) & 0x3fff3fffu;
    sbits[(size_t)frame * plan.bitsPerFrame + fl.bitsOffset + (size_t)(2 * sy + half) * (16 * fl.nCx) + 16 * cx + j] = make_uint2(accA, accB);
}

// ---------------------------------------------------------------------------------------------------------------
// One warp per reference FAST cell (src/ORBextractor.cc:805-868): gathers the cell's survivor bits (sparse), applies the
// iniThFAST / minThFAST rule and writes the survivors in raster order.  grid (ceil(nCells / 4), B), block 128.
constexpr int kSelWarps = 4;
constexpr int kSelMaxWords = 256;                // raster bitmap of a cell: <= 8192 pixel slots (checked on the host)

__global__ void __launch_bounds__(kSelWarps * 32) k_fast_select(const DevGeom* __restrict__ g, const FastCell* __restrict__ cells,
                                                                const FsPlan plan, const uint2* __restrict__ cbits,
                                                                uint32_t* __restrict__ cand, int* __restrict__ cellCount, int frameBase) {
    __shared__ uint32_t sKeep[kSelWarps][2][kSelMaxWords];       // survivors of the 3x3 NMS at minThFAST / iniThFAST, raster order
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int cellId = blockIdx.x * kSelWarps + warp;
    if (cellId >= g->nCells) return;
    const FastCell c = cells[cellId];
    const int frame = frameBase + blockIdx.y;
    const LevelGeom& L = g->lv[c.level];
    const FsLevel fl = plan.lv[c.level];
    const int pitch = L.pitch;
    const uint8_t* S = g->score[c.level] + (size_t)frame * L.planeBytes + (size_t)c.y0 * pitch + c.x0;    // cell origin
    const uint32_t* bits = reinterpret_cast<const uint32_t*>(cbits + (size_t)frame * plan.bitsPerFrame + fl.bitsOffset);
    const int w = c.w, h = c.h;
    const int nG = 16 * fl.nCx;
    const int sLo = (c.y0 - kFsY0) / kFsRows, sHi = (c.y0 + h - 1 - kFsY0) / kFsRows;
    const int gLo = (c.x0 - kFsX0) >> 2, gHi = (c.x0 + w - 1 - kFsX0) >> 2;
    const int nGc = gHi - gLo + 1;
    const int nW = (sHi - sLo + 1) * nGc * 2;
    const int nwr = (w * h + 31) >> 5;
    uint32_t* keepMin = sKeep[warp][0];
    uint32_t* keepIni = sKeep[warp][1];
    for (int i = lane; i < nwr; i += 32) { keepMin[i] = 0; keepIni[i] = 0; }
    __syncwarp();
    const int iniTh = g->iniTh;

    // survivors are sparse (~1 per 28-pixel word): every lane walks the set bits of its own words and files each survivor in
    // the cell's raster-order bitmaps, at both thresholds (no staging list, no prefix sum: the order comes from the bitmaps)
    for (int i = lane; i < nW; i += 32) {
        const int k = i & 1, t = i >> 1;
        const int s = t / nGc, q = t - s * nGc;
        uint32_t word = bits[((size_t)(sLo + s) * nG + gLo + q) * 2 + k];
        if (!word) continue;
        const int xs = kFsX0 + 4 * (gLo + q) + k - c.x0;         // local x of the low half's pixel; the high half's is xs + 2
        const int ys = kFsY0 + kFsRows * (sLo + s) - c.y0;       // local y of bit 0
        const int rlo = max(-ys, 0), rhi = min(h - 1 - ys, kFsRows - 1);
        const uint32_t rowmask = ((2u << rhi) - 1u) & ~((1u << rlo) - 1u);
        uint32_t m = 0;
        if (xs >= 0 && xs < w) m |= rowmask;
        if (xs + 2 >= 0 && xs + 2 < w) m |= rowmask << 16;
        word &= m;
        while (word) {
            const int b = __ffs(word) - 1;
            word &= word - 1;
            const int lx = xs + 2 * (b >> 4), ly = ys + (b & 15);
            const int sc = S[(size_t)ly * pitch + lx];
            const int p = ly * w + lx;
            atomicOr(&keepMin[p >> 5], 1u << (p & 31));
            if (sc > iniTh) atomicOr(&keepIni[p >> 5], 1u << (p & 31));
        }
    }
    __syncwarp();

    // cells keep the iniThFAST survivors, or the minThFAST ones when there are none (:826-846); raster order
    const int per = (nwr + 31) >> 5;
    const int wb = lane * per, we = min(wb + per, nwr);
    uint32_t anyI = 0;
    for (int i = wb; i < we; ++i) anyI |= keepIni[i];
    const uint32_t* keep = __any_sync(0xffffffffu, anyI != 0) ? keepIni : keepMin;
    int cnt = 0;
    for (int i = wb; i < we; ++i) cnt += __popc(keep[i]);
    int incl = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
    const int total = __shfl_sync(0xffffffffu, incl, 31);
    int rank = incl - cnt;
    uint32_t* out = cand + (size_t)frame * g->candCapFrame + c.outOffset;
    const uint32_t inv = w > 1 ? 0xffffffffu / (uint32_t)w + 1u : 0u;          // p / w = umulhi(p, inv) for p < 2^16
    for (int i = wb; i < we; ++i) {
        uint32_t m = keep[i];
        while (m) {
            const int b = __ffs(m) - 1;
            m &= m - 1;
            const int p = i * 32 + b;
            const int py = w > 1 ? (int)__umulhi((uint32_t)p, inv) : p, px = p - py * w;
            const int s = S[(size_t)py * pitch + px];
            out[rank++] = (uint32_t)px | ((uint32_t)py << 8) | ((uint32_t)(s - 1) << 16);
        }
    }
    if (lane == 0) cellCount[(size_t)frame * g->nCells + cellId] = total;
}

// ---------------------------------------------------------------------------------------------------------------
// E6, TMA-staged variant of k_blur7 (same arithmetic: separable fixed-point [18,34,48,56,48,34,18]/256, (sum + 2^15) >> 16,
// BORDER_REFLECT_101; src/ORBextractor.cc:1131-1133).  k_blur7 waits on DRAM once per strip row (issue-active 51 % for 75 M
// warp-instructions).  Here a warp's 120 x 30 strip arrives as ONE 144 x 36-byte tile by a 3-D TMA load (zero fill outside the
// level) into shared memory; the register pipeline of k_blur7 (horizontal taps through neighbour shuffles, vertical taps on
// packed row pairs) then reads its rows with conflict-free LDS.  TMA wants the box start 16-byte aligned in x: the tile starts
// at gx = floor16(x0 - 4), and the strip's words sit r = (x0 - 4 - gx) / 4 words into each tile row.
// REFLECT_101 at the level's edges is patched in shared memory after the load: mirrored columns first (rows inside the level),
// then whole mirrored rows, so that corners are mirrored in both directions.
constexpr int kBtBoxW = 144, kBtBoxH = kBlurStripH + 6;       // 36 rows
constexpr int kBtRowW = kBtBoxW / 4;                          // 36 words per tile row
constexpr int kBtTileBytes = kBtBoxW * kBtBoxH;               // 5184
constexpr int kBtTileStride = 5248;                           // 128-byte multiple

__global__ void __launch_bounds__(256) k_blur7_tma(const __grid_constant__ FsMaps maps, const DevGeom* __restrict__ g, BlurTiles bt,
                                                   int frameBase) {
    __shared__ __align__(128) uint8_t tiles[8][kBtTileStride];
    __shared__ __align__(8) uint64_t mbar[8];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int job = blockIdx.x * 8 + warp;
    const int nl = g->nlevels;
    if (job >= bt.start[nl]) return;                           // warps are independent: no block barrier below
    const int frame = frameBase + blockIdx.y;
    int l = 0;
    while (l + 1 < nl && job >= bt.start[l + 1]) ++l;
    const int t = job - bt.start[l];
    const int ty = t / bt.tilesX[l], tx = t - ty * bt.tilesX[l];
    const LevelGeom& L = g->lv[l];
    const int w = L.w, h = L.h;
    const int x0 = tx * kBlurStripW, y0 = ty * kBlurStripH;
    const int gx = (x0 - 4) & ~15, gy = y0 - 3;                // tile origin in level coordinates (may be negative: zero fill)
    uint8_t* tile = tiles[warp];
    {
        const uint32_t bar = smem_addr(&mbar[warp]);
        if (lane == 0) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar));
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(kBtTileBytes) : "memory");
            asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
                         :: "r"(smem_addr(tile)), "l"(reinterpret_cast<uint64_t>(&maps.m[l])), "r"(bar), "r"(gx), "r"(gy), "r"(frame)
                         : "memory");
        }
        __syncwarp();
        uint32_t done = 0;
        while (!done)
            asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                         : "=r"(done) : "r"(bar), "r"(0) : "memory");
    }
    // BORDER_REFLECT_101 patches (every level is >= 35 px wide and high, so a mirror source is always inside the tile)
    const bool edgeL = x0 == 0, edgeR = x0 + kBlurStripW + 3 > w, edgeT = y0 == 0, edgeB = y0 + kBlurStripH + 3 > h;
    if (edgeL | edgeR | edgeT | edgeB) {
        if (edgeL | edgeR) {
            for (int ry = lane; ry < kBtBoxH; ry += 32) {
                const int y = gy + ry;
                if (y < 0 || y >= h) continue;
                uint8_t* row = tile + ry * kBtBoxW;
                if (edgeL) { row[-1 - gx] = row[1 - gx]; row[-2 - gx] = row[2 - gx]; row[-3 - gx] = row[3 - gx]; }
                if (edgeR) {
#pragma unroll
                    for (int j = 0; j < 3; ++j) if (w + j - gx < kBtBoxW) row[w + j - gx] = row[w - 2 - j - gx];
                }
            }
            __syncwarp();
        }
        if (edgeT | edgeB) {
            uint32_t* tw = reinterpret_cast<uint32_t*>(tile);
            for (int i = lane; i < 3 * kBtRowW; i += 32) {
                const int j = i / kBtRowW, c = i - j * kBtRowW;
                if (edgeT) tw[(-1 - j - gy) * kBtRowW + c] = tw[(1 + j - gy) * kBtRowW + c];          // row -(j+1) <- row j+1
                if (edgeB && h + j - gy < kBtBoxH) tw[(h + j - gy) * kBtRowW + c] = tw[(h - 2 - j - gy) * kBtRowW + c];
            }
            __syncwarp();
        }
    }

    const int x = x0 + 4 * (lane - 1);                         // first column of this lane's word
    const uint32_t* tcol = reinterpret_cast<const uint32_t*>(tile) + ((x0 - 4 - gx) >> 2) + lane;
    uint8_t* dst = g->blur[l] + (size_t)frame * L.planeBytes;
    const int dp = L.pitch;
    const bool writer = lane >= 1 && lane <= 30 && x < w;
    const bool fullWord = x + 3 < w;

    uint32_t ring[6][4];            // ring[r % 6][j] = (h[r-1], h[r]) of pixel j as u16 pair
    uint32_t hprev[4] = {0, 0, 0, 0};
#pragma unroll
    for (int i = 0; i < 6; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) ring[i][j] = 0;

    const int rowsNeeded = min(kBlurStripH, h - y0) + 6;
    for (int r0 = 0; r0 < rowsNeeded; r0 += 6) {
#pragma unroll
        for (int k = 0; k < 6; ++k) {
            const int r = r0 + k;                               // tile row r = level row y0 + r - 3
            const int yo = y0 + r - 6;                          // output row completed by this input row
            const uint32_t C = tcol[min(r, kBtBoxH - 1) * kBtRowW];
            const uint32_t Lw = __shfl_up_sync(0xffffffffu, C, 1);
            const uint32_t Rw = __shfl_down_sync(0xffffffffu, C, 1);
            uint32_t hc[4];
            hc[0] = __dp4a(__byte_perm(Lw, C, 0x4321), 0x38302212u, __dp4a(__byte_perm(C, Rw, 0x4321), 0x00122230u, 0u));
            hc[1] = __dp4a(__byte_perm(Lw, C, 0x5432), 0x38302212u, __dp4a(__byte_perm(C, Rw, 0x5432), 0x00122230u, 0u));
            hc[2] = __dp4a(__byte_perm(Lw, C, 0x6543), 0x38302212u, __dp4a(__byte_perm(C, Rw, 0x6543), 0x00122230u, 0u));
            hc[3] = __dp4a(C, 0x38302212u, __dp4a(Rw, 0x00122230u, 0u));
            uint32_t v[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const uint32_t pr = __byte_perm(hprev[j], hc[j], 0x5410);      // (h[r-1], h[r])
                uint32_t acc = hc[j] * 18u + 32768u;
                acc = __dp2a_lo(ring[(k + 1) % 6][j], 0x2212u, acc);           // pair r-5 = (h[r-6], h[r-5]) . (18, 34)
                acc = __dp2a_lo(ring[(k + 3) % 6][j], 0x3830u, acc);           // pair r-3 = (h[r-4], h[r-3]) . (48, 56)
                acc = __dp2a_lo(ring[(k + 5) % 6][j], 0x2230u, acc);           // pair r-1 = (h[r-2], h[r-1]) . (48, 34)
                v[j] = acc;
                ring[k][j] = pr;
                hprev[j] = hc[j];
            }
            if (r >= 6 && yo < h && writer) {
                const uint32_t o = __byte_perm(__byte_perm(v[0], v[1], 0x0062), __byte_perm(v[2], v[3], 0x0062), 0x5410);
                uint8_t* d = dst + (size_t)yo * dp + x;
                if (fullWord) *reinterpret_cast<uint32_t*>(d) = o;
                else for (int j = 0; j < 4 && x + j < w; ++j) d[j] = (uint8_t)(o >> (8 * j));
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------
// E5 + E7, TMA-staged variant of k_orient_brief (same arithmetic, same outputs; src/ORBextractor.cc:76-146).  In k_orient_brief
// each of a keypoint's 16 warp-wide byte gathers on the blurred plane touches ~23 L1 sectors (149 MB of DRAM traffic per batch
// for 84 MB algorithmic).  Here lane 0 requests, before anything else, the blurred 37-row window of each keypoint of the warp's
// group as a 64 x 37-byte 3-D TMA tile (box start floor16(x - 18)): the loads land while the orientation moments are computed
// on the unblurred level, and the 16 gathers become shared-memory byte loads.  Rotated pattern coordinates are at most
// round(13 * sqrt(2)) = 18 in either direction, and keypoints lie >= 19 px inside their level, so the window never leaves it.
// The orientation window gets the same treatment (ncu: its 31 row loads with run-time pitch arithmetic were 44 % of the stall
// samples and ~450 of the kernel's 884 warp-instructions per keypoint): a 48 x 31-byte tile of the UNBLURRED level (box start
// floor16(x - 15)), read with one LDS per row at an immediate offset.
#ifndef ORBX_DT_WARPS
#define ORBX_DT_WARPS 2
#endif
constexpr int kDtWarps = ORBX_DT_WARPS;
constexpr int kDtBoxW = 64, kDtBoxH = 37, kDtTileBytes = kDtBoxW * kDtBoxH, kDtTileStride = 2432;
constexpr int kOtBoxW = 48, kOtBoxH = 31, kOtTileBytes = kOtBoxW * kOtBoxH, kOtTileStride = 1536;
constexpr int kDtSmemBytes = kDtWarps * kDescGroup * (kDtTileStride + kOtTileStride);      // 15,872 B with 2 warps x 2 keypoints (63,488 B with 4 x 4 pinned most of an SM's shared memory)

__global__ void __launch_bounds__(kDtWarps * 32) k_orient_brief_tma(const __grid_constant__ FsMaps bmaps, const __grid_constant__ FsMaps omaps,
                                                                     const DevGeom* __restrict__ g, const WorkItem* __restrict__ work,
                                                                     const int* __restrict__ outCnt, orbx_keypoint* outKp,
                                                                     uint8_t* outDesc, int frameBase) {
    extern __shared__ __align__(128) uint8_t dsm[];
    typedef uint8_t (*BlurTilesT)[kDescGroup][kDtTileStride];
    typedef uint8_t (*OriTilesT)[kDescGroup][kOtTileStride];
    BlurTilesT tiles = reinterpret_cast<BlurTilesT>(dsm);                                            // [warp][keypoint][2432]
    OriTilesT otiles = reinterpret_cast<OriTilesT>(dsm + kDtWarps * kDescGroup * kDtTileStride);    // [warp][keypoint][1536]
    __shared__ __align__(8) uint64_t mbar[kDtWarps][2];
    const int frame = frameBase + blockIdx.y;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int o0 = (blockIdx.x * kDtWarps + warp) * kDescGroup;
    const int total = outCnt[2 * frame];
    if (o0 >= total) return;                                   // warps are independent: no block barrier below
    const int ng = min(kDescGroup, total - o0);
    WorkItem mine; mine.level = 0; mine.pad = 0; mine.x = 0; mine.y = 0; mine.dst = 0;
    if (lane < ng) mine = work[(size_t)frame * g->outCap + o0 + lane];
    const uint32_t bar = smem_addr(&mbar[warp][0]), obar = smem_addr(&mbar[warp][1]);
    if (lane == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar));
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(obar));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(obar), "r"(ng * kOtTileBytes) : "memory");
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(ng * kDtTileBytes) : "memory");
    }
    __syncwarp();
    if (lane < ng) {                                           // two tile requests per keypoint of the group: orientation window first
        asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
                     :: "r"(smem_addr(otiles[warp][lane])), "l"(reinterpret_cast<uint64_t>(&omaps.m[mine.level])), "r"(obar),
                        "r"((mine.x - kHalfPatch) & ~15), "r"(mine.y - kHalfPatch), "r"(frame)
                     : "memory");
        asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
                     :: "r"(smem_addr(tiles[warp][lane])), "l"(reinterpret_cast<uint64_t>(&bmaps.m[mine.level])), "r"(bar),
                        "r"((mine.x - 18) & ~15), "r"(mine.y - 18), "r"(frame)
                     : "memory");
    }
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
    {
        uint32_t done = 0;
        while (!done)
            asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                         : "=r"(done) : "r"(obar), "r"(0) : "memory");
    }
    for (int j = 0; j < ng; ++j) {
        const int kx = __shfl_sync(0xffffffffu, mine.x, j);
        // centre of the disc inside the tile: row 15, column kx - floor16(kx - 15); lane 31 (u = 16) reads column <= 47 and is masked
        const uint8_t* ctr = otiles[warp][j] + kHalfPatch * kOtBoxW + (kx - ((kx - kHalfPatch) & ~15));
        int m10 = 0, m01 = 0;
#pragma unroll
        for (int v = -kHalfPatch; v <= kHalfPatch; ++v) {
            const int val = ((v < 0 ? -v : v) <= vm) ? (int)ctr[v * kOtBoxW + u] : 0;
            m10 += u * val;
            m01 += v * val;
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
    {
        uint32_t done = 0;
        while (!done)
            asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                         : "=r"(done) : "r"(bar), "r"(0) : "memory");
    }
    for (int j = 0; j < ng; ++j) {
        const float a = __shfl_sync(0xffffffffu, ca, j), b = __shfl_sync(0xffffffffu, sb, j);
        const int kx = __shfl_sync(0xffffffffu, mine.x, j);
        const int dst = __shfl_sync(0xffffffffu, mine.dst, j);
        // centre of the window inside the tile: row 18, column kx - floor16(kx - 18)
        const uint8_t* bc = tiles[warp][j] + 18 * kDtBoxW + (kx - ((kx - 18) & ~15));
        int byte = 0;
#pragma unroll
        for (int bit = 0; bit < 8; ++bit) {
            const int r0 = __float2int_rn(__fadd_rn(__fmul_rn(px0[bit], b), __fmul_rn(py0[bit], a)));
            const int c0 = __float2int_rn(__fsub_rn(__fmul_rn(px0[bit], a), __fmul_rn(py0[bit], b)));
            const int r1 = __float2int_rn(__fadd_rn(__fmul_rn(px1[bit], b), __fmul_rn(py1[bit], a)));
            const int c1 = __float2int_rn(__fsub_rn(__fmul_rn(px1[bit], a), __fmul_rn(py1[bit], b)));
            const int v0 = bc[r0 * kDtBoxW + c0], v1 = bc[r1 * kDtBoxW + c1];
            byte |= (v0 < v1 ? 1 : 0) << bit;
        }
        outDesc[((size_t)frame * g->outCap + dst) * 32 + lane] = (uint8_t)byte;
    }
}

// ---------------------------------------------------------------------------------------------------------------
// E1, TMA-staged pyramid level (same arithmetic as k_pyr_level_fast: cv::resize INTER_LINEAR in 11-bit fixed point, see there;
// src/ORBextractor.cc:1170-1195).  A CTA produces a 128 x 64 tile of level l; the source region its taps touch (known exactly
// from the tap tables: the host sizes the box per level, PyrBox) arrives as ONE 3-D TMA tile of level l-1.  A tap never
// selects a byte beyond the source row (the tables clamp), so the zero fill / neighbouring bytes that stand where the clamped
// duplicate words would be are never selected: bit-identical.
// A warp owns 128 columns x 8 output rows and walks the SOURCE rows it needs top to bottom: each source row gets its
// horizontal pass exactly once (3 LDS, 2 SHF and per pixel PRMT + IDP.2A + 2 bit ops), the previous row's result stays in
// registers, and an output row is emitted whenever the walk reaches its lower tap row - a warp-uniform test, so there is no
// per-row reuse bookkeeping; the eight row taps of the warp sit in lanes 0-7 and are broadcast by shuffles (the first TMA
// version spent 44 instructions per pixel, most of them on bookkeeping and tap reloads; this one ~12).
// grid (ceil(w/128), ceil(h/64), B), block 256; dynamic smem = boxW * boxH.
struct PyrBox { int boxW, boxH; };
constexpr int kPrTileH = 64, kPrRowsPerWarp = 8;

// one 128 x 64 tile of level l (tile column bx, tile row by) of `frame`; srcMap = tensor map of level l-1; whole CTA (256 threads)
__device__ __forceinline__ void pyr_level_tile(const CUtensorMap* srcMap, const DevGeom* __restrict__ g, int l, const ResizeTap* __restrict__ taps,
                                               PyrBox pb, int frame, int bx, int by, uint8_t* ptile, uint64_t& mbar) {
    const LevelGeom& D = g->lv[l];
    const LevelGeom& S = g->lv[l - 1];
    const int X0 = bx * 128, Y0 = by * kPrTileH;
    // tile origin in the source level: first tap of the CTA's first column (16-byte aligned) / first row (clamped like the taps)
    const int gx = taps[D.tabX + X0].ofs & ~15;
    const int gy = min(max(taps[D.tabY + Y0].ofs, 0), S.h - 1);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t bar = smem_addr(&mbar);
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(pb.boxW * pb.boxH) : "memory");
        asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
                     :: "r"(smem_addr(ptile)), "l"(reinterpret_cast<uint64_t>(srcMap)), "r"(bar), "r"(gx), "r"(gy), "r"(frame)
                     : "memory");
    }
    __syncthreads();                                           // the barrier object is initialised before anyone polls it
    // lanes right of the level keep walking with the last column group (the row taps are exchanged by full-warp shuffles) and store nothing
    const int x4 = min(X0 + lane * 4, (D.w - 1) & ~3);
    const int yb = Y0 + warp * kPrRowsPerWarp;
    const int yEnd = min(yb + kPrRowsPerWarp, D.h);
    const bool active = yb < D.h;                               // warp-uniform
    uint32_t coef[4], sel[4];
    int wofs = 0, sh = 0;
    if (active) {                                              // per-thread column setup overlaps the tile's flight
        const int a0 = taps[D.tabX + x4].ofs;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const ResizeTap t = taps[D.tabX + min(x4 + k, D.w - 1)];
            const int r0 = t.ofs - a0, r1 = min(t.ofs + 1, S.w - 1) - a0;
            sel[k] = (uint32_t)r0 | ((uint32_t)r1 << 4);
            coef[k] = (uint32_t)(uint16_t)t.c0 | ((uint32_t)(uint16_t)t.c1 << 16);
        }
        wofs = ((a0 & ~3) - gx) >> 2;                          // word of the tile row that holds byte a0
        sh = (a0 & 3) * 8;
    }
    {
        uint32_t done = 0;
        while (!done)
            asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                         : "=r"(done) : "r"(bar), "r"(0) : "memory");
    }
    if (!active) return;
    const int rowW = pb.boxW >> 2;
    const int nrows = yEnd - yb;
    // lane j keeps the vertical tap of output row yb + j (the host admits only levels whose row taps never clamp: s1 = s0 + 1)
    const ResizeTap myT = taps[D.tabY + min(yb + (lane & (kPrRowsPerWarp - 1)), D.h - 1)];
    const uint32_t myC = (uint32_t)(uint16_t)myT.c0 | ((uint32_t)(uint16_t)myT.c1 << 16);
    const int pitch = D.pitch;
    int fullI, wLeft;                                          // pinned in registers: the compiler otherwise recomputes them from tid / ctaid per row
    asm volatile("mov.u32 %0, %1;" : "=r"(fullI) : "r"((int)(X0 + lane * 4 + 3 < D.w)));
    asm volatile("mov.u32 %0, %1;" : "=r"(wLeft) : "r"(X0 + lane * 4 < D.w ? D.w - x4 : 0));
    uint8_t* d = g->lvl[l] + (size_t)frame * D.planeBytes + (size_t)yb * pitch + x4;
    int r = __shfl_sync(0xffffffffu, myT.ofs, 0) - 1;          // the walk computes row r + 1 next
    const uint32_t* rowp = reinterpret_cast<const uint32_t*>(ptile) + wofs + (r - gy) * rowW;
    uint32_t h[4] = {0, 0, 0, 0}, hPrev[4] = {0, 0, 0, 0};
    for (int j = 0; j < nrows; ++j) {
        const int s1 = __shfl_sync(0xffffffffu, myT.ofs, j) + 1;
        const uint32_t cc = __shfl_sync(0xffffffffu, myC, j);
#pragma unroll 1
        while (r < s1) {                                        // horizontal pass of the next source row: (p[a]*c0 + p[a1]*c1) & ~15
            ++r; rowp += rowW;
            const uint32_t W0 = rowp[0], W1 = rowp[1], W2 = rowp[2];
            const uint32_t lo = __funnelshift_r(W0, W1, sh), hi = __funnelshift_r(W1, W2, sh);
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                hPrev[k] = h[k];
                h[k] = __dp2a_lo(coef[k], __byte_perm(lo, hi, sel[k]), 0u) & 0xfffffff0u;
            }
        }
        // ((b0 * (hA >> 4)) >> 16) + ((b1 * (hB >> 4)) >> 16) with h = 16 * (h >> 4) and b << 12: one IMAD.HI each
        const uint32_t b0 = (cc & 0xffffu) << 12, b1 = (cc >> 16) << 12;
        uint32_t v[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) v[k] = (__umulhi(h[k], b1) + (__umulhi(hPrev[k], b0) + 2u)) >> 2;
        const uint32_t packed = __byte_perm(__byte_perm(v[0], v[1], 0x0040), __byte_perm(v[2], v[3], 0x0040), 0x5410);
        if (fullI) *reinterpret_cast<uint32_t*>(d) = packed;
        else for (int k = 0; k < 4 && k < wLeft; ++k) d[k] = (uint8_t)(packed >> (8 * k));
        d += pitch;
    }
}

__global__ void __launch_bounds__(256) k_pyr_level_tma(const __grid_constant__ CUtensorMap srcMap, const DevGeom* __restrict__ g, int l,
                                                       const ResizeTap* __restrict__ taps, PyrBox pb, int frameBase) {
    extern __shared__ __align__(128) uint8_t ptile[];
    __shared__ __align__(8) uint64_t mbar;
    pyr_level_tile(&srcMap, g, l, taps, pb, frameBase + blockIdx.z, blockIdx.x, blockIdx.y, ptile, mbar);
}

// The whole pyramid of a FEW frames in ONE launch (single-frame calls: seven dependent launches cost ~7 us each for ~2 us of work).
// CTA = one tile of one level; the tiles are numbered level by level, so every tile a CTA depends on has a lower block index and
// the host launches this only when ALL tiles are co-resident (no CTA can wait for one that has not been scheduled).  A tile of
// level l > 1 waits until the arrival counter of level l-1 of its frame has reached that level's tile count; the producers bump
// it after their stores (fence + atomic), the consumer acquires it and crosses to the async proxy before its TMA load.  The last
// tile of the last level puts the frame's counters back to zero for the next pass.
struct PyrChain {
    int tileBegin[kMaxLevels + 1];       // first tile index of level l (l = 1 .. nlevels-1), tileBegin[nlevels] = total
    int tilesX[kMaxLevels];
    PyrBox box[kMaxLevels];
};
__global__ void __launch_bounds__(256) k_pyr_chain_tma(const __grid_constant__ FsMaps maps, const DevGeom* __restrict__ g,
                                                       const ResizeTap* __restrict__ taps, const PyrChain pc, int frameBase,
                                                       unsigned int* __restrict__ arrive /* [frame][kMaxLevels] */) {
    extern __shared__ __align__(128) uint8_t ptile[];
    __shared__ __align__(8) uint64_t mbar;
    const int L = g->nlevels;
    int l = 1;
    while (l + 1 < L && (int)blockIdx.x >= pc.tileBegin[l + 1]) ++l;
    const int t = (int)blockIdx.x - pc.tileBegin[l];
    const int by = t / pc.tilesX[l], bx = t - by * pc.tilesX[l];
    const int frame = frameBase + blockIdx.y;
    unsigned int* cnt = arrive + (size_t)blockIdx.y * kMaxLevels;
    if (l > 1) {
        if (threadIdx.x == 0) {
            const unsigned int need = (unsigned int)(pc.tileBegin[l] - pc.tileBegin[l - 1]);
            unsigned int v;
            // bounded: the tiles below have lower block indices and are dispatched first, so this cannot wait for an unscheduled
            // CTA; should that ever not hold (it would be a scheduler that dispatches out of order), give up after ~0.1 s instead of
            // hanging the GPU
            const long long t0 = clock64();
            do { asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(cnt + l - 1) : "memory"); } while (v < need && clock64() - t0 < 200000000LL);
            asm volatile("fence.proxy.async;" ::: "memory");          // the level below was written through the generic proxy; TMA reads through the async one
        }
        __syncthreads();
    }
    pyr_level_tile(&maps.m[l], g, l, taps, pc.box[l], frame, bx, by, ptile, mbar);
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned int done = atomicAdd(cnt + l, 1u) + 1u;
        if (l == L - 1 && done == (unsigned int)(pc.tileBegin[L] - pc.tileBegin[L - 1]))
            for (int i = 1; i < L; ++i) cnt[i] = 0;                    // every tile of this frame has read what it needed
    }
}

}  // namespace orbx
