// Image-plane kernels of the extractor: pyramid level (E1), per-cell FAST (E2), 7x7 Gaussian (E6).
// Reference behaviour: src/ORBextractor.cc:1170-1195 (ComputePyramid), :781-872 (FAST grid), :1131-1133 (blur);
// pixel arithmetic = OpenCV 4.x 8-bit paths, restated independently in oracle/cpp.
#pragma once
#include "orbx_internal.h"

namespace orbx {

struct Level0 {                    // where level 0 (the input image) lives for this pass
    const uint8_t* base;
    size_t frameBytes;
    int pitch;
};

__device__ __forceinline__ const uint8_t* level_ptr(const DevGeom* g, const Level0& l0, int l, int frame, int* pitch) {
    if (l == 0) { *pitch = l0.pitch; return l0.base + (size_t)frame * l0.frameBytes; }
    *pitch = g->lv[l].pitch;
    return g->lvl[l] + (size_t)frame * g->lv[l].planeBytes;
}

// ---------------------------------------------------------------------------------------------------------------
// E1: one pyramid level from the previous one.  cv::resize INTER_LINEAR on 8U: 11-bit fixed-point taps, horizontal
// pass in int32, vertical pass ((b0*(r0>>4))>>16 + (b1*(r1>>4))>>16 + 2)>>2.  Each thread produces 4 adjacent
// pixels of one row and stores them as one 32-bit word (rows are 128-B aligned).  HBM-bound: reads ~1.44 source
// bytes and writes 1 byte per output pixel; taps come from a table built once per shape on the host.
// grid (ceil(w/128), ceil(h/8), B), block (32, 8).
__global__ void __launch_bounds__(256) k_pyr_level(const DevGeom* __restrict__ g, Level0 l0, int l,
                                                   const ResizeTap* __restrict__ taps, int frameBase) {
    const LevelGeom& D = g->lv[l];
    const LevelGeom& S = g->lv[l - 1];
    const int frame = frameBase + blockIdx.z;
    const int y = blockIdx.y * blockDim.y + threadIdx.y;
    const int x4 = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (y >= D.h || x4 >= D.w) return;
    int sp;
    const uint8_t* src = level_ptr(g, l0, l - 1, frame, &sp);
    uint8_t* dst = g->lvl[l] + (size_t)frame * D.planeBytes + (size_t)y * D.pitch;
    uint32_t packed = 0;
    if (D.areaFast) {
        const uint8_t* r0 = src + (size_t)(2 * y) * sp;
        const uint8_t* r1 = r0 + sp;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            int x = x4 + k;
            if (x < D.w) {
                int v = (r0[2 * x] + r0[2 * x + 1] + r1[2 * x] + r1[2 * x + 1] + 2) >> 2;
                packed |= (uint32_t)v << (8 * k);
            }
        }
    } else {
        const ResizeTap ty = taps[D.tabY + y];
        const int s0 = min(max(ty.ofs, 0), S.h - 1), s1 = min(max(ty.ofs + 1, 0), S.h - 1);
        const uint8_t* r0 = src + (size_t)s0 * sp;
        const uint8_t* r1 = src + (size_t)s1 * sp;
        const int b0 = ty.c0, b1 = ty.c1;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            int x = x4 + k;
            if (x < D.w) {
                const ResizeTap tx = taps[D.tabX + x];
                const int a = tx.ofs, a1 = min(tx.ofs + 1, S.w - 1);
                const int h0 = r0[a] * tx.c0 + r0[a1] * tx.c1;
                const int h1 = r1[a] * tx.c0 + r1[a1] * tx.c1;
                const int v = (((b0 * (h0 >> 4)) >> 16) + ((b1 * (h1 >> 4)) >> 16) + 2) >> 2;
                packed |= (uint32_t)v << (8 * k);
            }
        }
    }
    if (x4 + 3 < D.w) {
        *reinterpret_cast<uint32_t*>(dst + x4) = packed;
    } else {
        for (int k = 0; k < 4 && x4 + k < D.w; ++k) dst[x4 + k] = (uint8_t)(packed >> (8 * k));
    }
}

// E1, fast path (scale factor <= 2, word-aligned source rows).  A thread produces 4 adjacent pixels of 4 consecutive rows.
// Horizontal taps: the (at most 7) source bytes its 4 pixels read are cut out of three aligned words with two funnel
// shifts, each pixel's pair (p[a], p[a+1]) is picked with one PRMT (selector precomputed from the tap table) and weighted
// with one IDP.2A against its (c0, c1) pair.  Source rows are shared by consecutive output rows (1.2 source rows per output
// row), so the horizontal results of the last two source rows stay in registers.  Vertical pass: two IMAD.HI per pixel on
// operands pre-shifted so that hi32((h>>4 << 16) * b) == (b * (h>>4)) >> 16.  ~13 instructions per pixel instead of ~65.
// grid (ceil(w/128), ceil(h/32), B), block (32, 8).
constexpr int kPyrRows = 4;

__global__ void __launch_bounds__(256) k_pyr_level_fast(const DevGeom* __restrict__ g, Level0 l0, int l,
                                                        const ResizeTap* __restrict__ taps, int frameBase) {
    const LevelGeom& D = g->lv[l];
    const LevelGeom& S = g->lv[l - 1];
    const int frame = frameBase + blockIdx.z;
    const int x4 = (blockIdx.x * 32 + threadIdx.x) * 4;
    const int yb = (blockIdx.y * 8 + threadIdx.y) * kPyrRows;
    if (x4 >= D.w || yb >= D.h) return;
    int sp;
    const uint8_t* src = level_ptr(g, l0, l - 1, frame, &sp);
    uint8_t* dst = g->lvl[l] + (size_t)frame * D.planeBytes;
    // per-thread column setup
    uint32_t coef[4], sel[4];
    const ResizeTap t0 = taps[D.tabX + x4];
    const int a0 = t0.ofs;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const ResizeTap t = taps[D.tabX + min(x4 + k, D.w - 1)];
        const int r0 = t.ofs - a0, r1 = min(t.ofs + 1, S.w - 1) - a0;
        sel[k] = (uint32_t)r0 | ((uint32_t)r1 << 4);
        coef[k] = (uint32_t)(uint16_t)t.c0 | ((uint32_t)(uint16_t)t.c1 << 16);
    }
    const int maxWord = (S.w - 1) >> 2;
    const int w0i = a0 >> 2, w1i = min(w0i + 1, maxWord), w2i = min(w0i + 2, maxWord);
    const int sh = (a0 & 3) * 8;
    uint32_t hA[4], hB[4];                  // pre-shifted horizontal results of source rows rA, rB
    int rA = -1, rB = -1;
    auto hrow = [&](int r, uint32_t (&h)[4]) {
        const uint32_t* rowp = reinterpret_cast<const uint32_t*>(src + (size_t)r * sp);
        const uint32_t W0 = rowp[w0i], W1 = rowp[w1i], W2 = rowp[w2i];
        const uint32_t lo = __funnelshift_r(W0, W1, sh), hi = __funnelshift_r(W1, W2, sh);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const uint32_t hk = __dp2a_lo(coef[k], __byte_perm(lo, hi, sel[k]), 0u);     // p[a]*c0 + p[a1]*c1
            h[k] = (hk << 12) & 0xffff0000u;                                             // (hk >> 4) << 16
        }
    };
#pragma unroll
    for (int j = 0; j < kPyrRows; ++j) {
        const int y = yb + j;
        if (y >= D.h) break;
        const ResizeTap ty = taps[D.tabY + y];
        const int s0 = min(max(ty.ofs, 0), S.h - 1), s1 = min(max(ty.ofs + 1, 0), S.h - 1);
        // bring rows s0, s1 into (hA, hB); consecutive output rows usually need only one new source row
        if (s0 == rB) {
#pragma unroll
            for (int k = 0; k < 4; ++k) hA[k] = hB[k];
            rA = rB; rB = -1;
        } else if (s0 != rA) { hrow(s0, hA); rA = s0; }
        if (s1 == rA) {
#pragma unroll
            for (int k = 0; k < 4; ++k) hB[k] = hA[k];
            rB = rA;
        } else if (s1 != rB) { hrow(s1, hB); rB = s1; }
        const uint32_t b0 = (uint32_t)ty.c0, b1 = (uint32_t)ty.c1;
        uint32_t v[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) v[k] = (__umulhi(hA[k], b0) + __umulhi(hB[k], b1) + 2u) >> 2;
        const uint32_t packed = __byte_perm(__byte_perm(v[0], v[1], 0x0040), __byte_perm(v[2], v[3], 0x0040), 0x5410);
        uint8_t* d = dst + (size_t)y * D.pitch + x4;
        if (x4 + 3 < D.w) *reinterpret_cast<uint32_t*>(d) = packed;
        else for (int k = 0; k < 4 && x4 + k < D.w; ++k) d[k] = (uint8_t)(packed >> (8 * k));
    }
}

// ---------------------------------------------------------------------------------------------------------------
// E2: FAST-9/16 strength of one pixel: s = max over the 16 cyclic 9-arcs of max(min(v - ring), min(ring - v)).
// cv::FAST declares a corner at threshold t iff s > t and reports response s-1.
// Sliding 9-window minima by doubling (2,4,8,+1) on both d = v-ring and e = ring-v.
// NOTE (toolchain): the obvious form max(lo9, -hi9) (min and max over the same d) MISCOMPILES with nvcc 12.9 for
// sm_100a: ptxas fuses the negation into VIMNMX3 and the result is max(d) (measured on B200: 69 instead of 0 for
// v=123, ring={78,128,101,87,114,99,101,115,95,111,129,76,83,87,54,83}; tools/t_fs.cu).  Using only min/max of
// non-negated operands avoids the pattern; tests/test_extractor_gpu.py pins the vector.
__device__ __forceinline__ int fast_strength(int v, const int (&r)[16]) {
    int d[16], e[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) { d[k] = v - r[k]; e[k] = r[k] - v; }
    int d2[16], e2[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) { d2[k] = min(d[k], d[(k + 1) & 15]); e2[k] = min(e[k], e[(k + 1) & 15]); }
    int d4[16], e4[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) { d4[k] = min(d2[k], d2[(k + 2) & 15]); e4[k] = min(e2[k], e2[(k + 2) & 15]); }
    int best = 0;
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        const int d9 = min(min(d4[k], d4[(k + 4) & 15]), d[(k + 8) & 15]);
        const int e9 = min(min(e4[k], e4[(k + 4) & 15]), e[(k + 8) & 15]);
        best = max(best, max(d9, e9));
    }
    return best;
}

constexpr int kFastThreads = 256;          // widest variant; a cell may hold at most 32 * threads pixels

// FAST-9/16 strength of TWO horizontally adjacent pixels at once in u16x2 lanes:
//   s = max( max_i min(W_i) - v, v - min_i max(W_i), 0 ),  W_i = the 9 ring pixels r[i..i+8] (cyclic), i = 0..15.
// min / max commute with the per-lane offset (256 - v), so the windows run on the raw ring values and the centre is applied
// once at the end.  Cost model (tools/microbench.cu, profiles/int_peaks_b200.txt): the 2-input VIMNMX.U16x2 issues at 128
// /clk/SM, the 3-input VIMNMX3.U16x2 at 64 — a 3-input op costs two ALU-pipe slots, and this kernel is bound by that pipe.
// The windows are therefore evaluated in PAIRS: W_2k and W_2k+1 share the 8 pixels C_k = r[2k+1..2k+8], and
//   max( min(W_2k), min(W_2k+1) ) = min( min(C_k), max(r[2k], r[2k+9]) )          (lattice distributivity)
// with min(C_k) = min(qq[k], qq[k+2]), qq[k] = min(q[k], q[k+1]), q[k] = min(r[2k+1], r[2k+2]).  Per side: 8 + 8 + 8
// two-input ops, 8 three-input ops and a 4-op reduction = 36 instructions / 47 pipe slots (the plain sliding-window form,
// min3 over min3, took 40 / 79).  Pure min / max identities: the result is bit-identical.
// fast_strength_biased_x2 returns s + 256 in each u16 lane (the callers that pack bytes or test bit patterns use it as is).
__device__ __forceinline__ uint32_t fast_strength_biased_x2(uint32_t v2, const uint32_t (&r)[16]) {
    uint32_t q[8], Q[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        q[k] = __vminu2(r[2 * k + 1], r[(2 * k + 2) & 15]);
        Q[k] = __vmaxu2(r[2 * k + 1], r[(2 * k + 2) & 15]);
    }
    uint32_t qq[8], QQ[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        qq[k] = __vminu2(q[k], q[(k + 1) & 7]);
        QQ[k] = __vmaxu2(Q[k], Q[(k + 1) & 7]);
    }
    uint32_t a[8], b[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const uint32_t e = __vmaxu2(r[2 * k], r[(2 * k + 9) & 15]);
        const uint32_t E = __vminu2(r[2 * k], r[(2 * k + 9) & 15]);
        a[k] = __vimin3_u16x2(qq[k], qq[(k + 2) & 7], e);
        b[k] = __vimax3_u16x2(QQ[k], QQ[(k + 2) & 7], E);
    }
    uint32_t A = __vimax3_u16x2(a[0], a[1], a[2]), B = __vimin3_u16x2(b[0], b[1], b[2]);
    A = __vimax3_u16x2(A, a[3], a[4]); B = __vimin3_u16x2(B, b[3], b[4]);
    A = __vimax3_u16x2(A, a[5], a[6]); B = __vimin3_u16x2(B, b[5], b[6]);
    A = __vmaxu2(A, a[7]); B = __vminu2(B, b[7]);
    // per lane: A + 256 - v and v + 256 - B are in [1, 511]: no borrow between the two u16 lanes
    return __vimax3_u16x2(A + (0x01000100u - v2), (v2 + 0x01000100u) - B, 0x01000100u);
}

__device__ __forceinline__ uint32_t fast_strength_x2(uint32_t v2, const uint32_t (&r)[16]) {
    return fast_strength_biased_x2(v2, r) - 0x01000100u;
}

// One CTA per reference FAST cell.  The cell's pixels (+3 ring halo) are read with aligned 32-bit loads and expanded in
// shared memory to u16 lanes in TWO copies (pairs starting at even / odd x), so that any two horizontally adjacent pixels
// are one aligned 32-bit shared load; each thread then computes the strength of two pixels per pass (fast_strength_x2),
// once per pixel at minThFAST.  The reference's per-cell logic runs on the shared score tile: NMS at iniThFAST; if that
// leaves the cell empty, NMS at minThFAST (src/ORBextractor.cc:826-846).  Neighbours outside the cell's tested interior
// count as 0, exactly like a cv::FAST call on the sub-image.
// Corners are sparse, so the strength pass also sets one bit per corner in a raster-order bitmap; NMS then visits only
// the set bits (one bitmap word per thread), keeps its verdicts as two more bitmap words (ini / min threshold) and the
// survivors are compacted in raster order with a popcount prefix sum - no second pass over all pixels, no sorting.
// grid (nCells, B), block 256.  dynamic smem: 2 x rows x pitchW words + score tile.
template <int kThreads, int kPitch>      // kPitch: compile-time tile pitch in words (0 = use the pitchW argument)
__global__ void __launch_bounds__(kThreads) k_fast_cells(const DevGeom* __restrict__ g, Level0 l0,
                                                             const FastCell* __restrict__ cells,
                                                             uint32_t* __restrict__ cand, int* __restrict__ cellCount,
                                                             int pitchWArg, int tileRowsMax, int frameBase,
                                                             const int* __restrict__ cellOrder, int orderBase) {
    extern __shared__ __align__(16) uint8_t smem[];
    const int pitchW = kPitch ? kPitch : pitchWArg;    // a constant pitch turns the 17 ring addresses into immediates
    // candidate bitmaps by PAIR index (bit p = pair p holds a candidate in its even / odd pixel), written with ballots
    __shared__ uint32_t cmE[kThreads / 2 + 2], cmO[kThreads / 2 + 2];
    __shared__ int warpSum[kThreads / 32];
    __shared__ int total;
    // cells are launched in groups of similar height (cellOrder): the dynamic shared memory of a launch is sized for its own
    // group, so that the many ordinary cells are not held to the occupancy of the few tall ones of the coarse levels
    const int cellId = cellOrder[orderBase + blockIdx.x];
    const FastCell c = cells[cellId];
    const int frame = frameBase + blockIdx.y;
    const int tid = threadIdx.x;
    int pitch;
    const uint8_t* img = level_ptr(g, l0, c.level, frame, &pitch);
    const int w = c.w, h = c.h;
    const int tw = w + 6, th = h + 6;
    uint32_t* SA = reinterpret_cast<uint32_t*>(smem);                // pairs starting at even X: (X, X+1)
    uint32_t* SB = SA + (size_t)tileRowsMax * pitchW;                // pairs starting at odd X
    // raw strengths of ALL tested pixels, one byte each: pixel (px, py) at (py+1)*sp + px + 2, so that a pair (even px)
    // is one aligned 16-bit store; rows 0 and h+1 and the columns px = -1 and px = w stay 0 (the apron)
    uint8_t* score = reinterpret_cast<uint8_t*>(SB + (size_t)tileRowsMax * pitchW);
    const int sp = (w + 4) & ~1;
    const int gx0 = c.x0 - 3, gy0 = c.y0 - 3;
    const uint8_t* origin = img + (size_t)gy0 * pitch + gx0;
    const bool aligned = (((uintptr_t)img | (uintptr_t)pitch) & 3) == 0;
    const int a = aligned ? (int)((uintptr_t)origin & 3) : 0;        // X = tile x + a; rows start word-aligned when aligned
    const int pw = (w + 1) >> 1;                 // pairs per row; the last one is half empty when w is odd
    const int npairs = pw * h;

    if (aligned) {
        // 16 threads per tile row (typical cells need 11-12 words per row)
        const int wordsPerRow = (tw + a + 3) >> 2;
        const uint8_t* base = origin - a;
        const int lx = tid & 15;
        constexpr int RS = kThreads / 16;               // rows per pass
        for (int rw = lx; rw < wordsPerRow; rw += 16) {
            for (int ryb = tid >> 4; ryb < th; ryb += 4 * RS) {
                // four rows in flight: all loads are issued before the first expansion
                uint32_t v[4], nx[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const int ry = ryb + j * RS;
                    if (ry < th) {
                        const uint8_t* rowp = base + (size_t)ry * pitch + 4 * rw;
                        v[j] = *reinterpret_cast<const uint32_t*>(rowp);
                        nx[j] = rowp[4];
                    }
                }
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const int ry = ryb + j * RS;
                    if (ry < th) {
                        uint32_t* ra = SA + ry * pitchW + 2 * rw;
                        uint32_t* rb = SB + ry * pitchW + 2 * rw;
                        ra[0] = __byte_perm(v[j], 0, 0x4140);          // (p0, p1) as u16 lanes
                        ra[1] = __byte_perm(v[j], 0, 0x4342);          // (p2, p3)
                        rb[0] = __byte_perm(v[j], 0, 0x4241);          // (p1, p2)
                        rb[1] = __byte_perm(v[j], nx[j], 0x5453);      // (p3, next p0); nx is zero-extended, its byte 1 is 0
                    }
                }
            }
        }
    } else {
        const int pairsPerRow = (tw + 2) >> 1;
        for (int i = tid; i < pairsPerRow * th; i += kThreads) {
            const int ry = i / pairsPerRow, rp = i - ry * pairsPerRow;
            const uint8_t* src = origin + (size_t)ry * pitch + 2 * rp;
            const uint32_t p0 = src[0], p1 = src[1], p2 = src[2];
            SA[ry * pitchW + rp] = p0 | (p1 << 16);
            SB[ry * pitchW + rp] = p1 | (p2 << 16);
        }
    }
    {
        uint16_t* s16 = reinterpret_cast<uint16_t*>(score);          // apron only: every tested pixel is written below
        const int sp2 = sp >> 1;
        for (int i = tid; i < sp2; i += kThreads) { s16[i] = 0; s16[(h + 1) * sp2 + i] = 0; }
        for (int i = tid; i < h; i += kThreads) {
            s16[(i + 1) * sp2] = 0;                                   // px = -1 (and the unused byte before it)
            if (!(w & 1)) s16[(i + 1) * sp2 + ((w + 2) >> 1)] = 0;    // px = w; for odd w the last pair's store covers it
        }
    }
    __syncthreads();

    const int minTh = g->minTh, iniTh = g->iniTh;
    const int lane = tid & 31, wid = tid >> 5;
    {
        // tested pixel (px, py) = tile (px+3, py+3); pair start X = px0 + 3 + a + dx with px0 even
        const int ap = a + 3;
        const uint32_t* E = ((ap & 1) ? SB : SA) + (ap >> 1);          // even dx: index px0/2 + dx/2
        const uint32_t* O = ((ap & 1) ? SA : SB) + ((ap + 1) >> 1);    // odd dx : index px0/2 + (dx-1)/2
        const int P = pitchW;
        // (py, pc) of pair p advance by a constant step per pass: one division per thread instead of one per pair
        const int stepY = kThreads / pw, stepC = kThreads - stepY * pw;
        int py = tid / pw, pc = tid - py * pw;
        // whole warps run the loop (the ballots need every lane); lanes past the last pair only vote 'no'
        for (int p0 = wid * 32; p0 < npairs; p0 += kThreads) {
            bool c0 = false, c1 = false;
            if (p0 + lane < npairs) {
                const int b = (py + 3) * P + pc;
                const uint32_t v2 = E[b];
                uint32_t r[16];
                r[0] = E[b + 3 * P];       r[1] = O[b + 3 * P];       r[2] = E[b + 2 * P + 1];   r[3] = O[b + P + 1];
                r[4] = O[b + 1];           r[5] = O[b - P + 1];       r[6] = E[b - 2 * P + 1];   r[7] = O[b - 3 * P];
                r[8] = E[b - 3 * P];       r[9] = O[b - 3 * P - 1];   r[10] = E[b - 2 * P - 1];  r[11] = O[b - P - 2];
                r[12] = O[b - 2];          r[13] = O[b + P - 2];      r[14] = E[b + 2 * P - 1];  r[15] = O[b + 3 * P - 1];
                uint32_t s2 = fast_strength_x2(v2, r);
                if (2 * pc + 1 >= w) s2 &= 0xffffu;                    // the odd pixel of the last pair lies outside the cell
                *reinterpret_cast<uint16_t*>(score + (py + 1) * sp + 2 * pc + 2) = (uint16_t)__byte_perm(s2, 0, 0x4420);
                c0 = (int)(s2 & 0xffffu) > minTh;
                c1 = (int)(s2 >> 16) > minTh;
                py += stepY; pc += stepC;
                if (pc >= pw) { pc -= pw; ++py; }
            }
            const uint32_t bE = __ballot_sync(0xffffffffu, c0), bO = __ballot_sync(0xffffffffu, c1);
            if (lane == 0) { cmE[p0 >> 5] = bE; cmO[p0 >> 5] = bO; }
        }
        // the two words after the last one are read by the funnel shifts below
        if (tid < 2) { cmE[((npairs + 31) >> 5) + tid] = 0; cmO[((npairs + 31) >> 5) + tid] = 0; }
    }
    __syncthreads();

    // NMS, candidates only.  A corner survives iff its strength exceeds all 8 neighbours' (cv::FAST compares the stored
    // scores, strength - 1 for corners and 0 otherwise: a neighbour at or below the threshold can never win against a
    // corner above it, so comparing raw strengths is the same test at either threshold); cells keep the iniThFAST
    // survivors, or the minThFAST ones when there are none (src/ORBextractor.cc:826-846).
    // Every thread owns a contiguous run of <= 16 pairs = 32 pixel slots in padded raster order (row width w2 = 2 * pw).
    const int w2 = 2 * pw;
    const int per = (npairs + kThreads - 1) / kThreads;
    const int pb = tid * per, pe = min(pb + per, npairs);
    uint32_t keepIni = 0, keepMin = 0;
    if (pb < pe) {
        uint32_t z = (__funnelshift_r(cmE[pb >> 5], cmE[(pb >> 5) + 1], pb & 31) & 0xffffu) |
                     (__funnelshift_r(cmO[pb >> 5], cmO[(pb >> 5) + 1], pb & 31) << 16);
        if (pe - pb < 16) z &= ((1u << (pe - pb)) - 1) * 0x00010001u;
        // perfect shuffle: bit i of the low half -> bit 2i, bit i of the high half -> bit 2i+1
        uint32_t t;
        t = (z ^ (z >> 8)) & 0x0000ff00u; z ^= t ^ (t << 8);
        t = (z ^ (z >> 4)) & 0x00f000f0u; z ^= t ^ (t << 4);
        t = (z ^ (z >> 2)) & 0x0c0c0c0cu; z ^= t ^ (t << 2);
        t = (z ^ (z >> 1)) & 0x22222222u; z ^= t ^ (t << 1);
        uint32_t m = z;
        const int py = pb / pw, px = 2 * (pb - py * pw);   // chunk origin; a corner's coordinates follow by carry, no division
        while (m) {
            const int bpos = __ffs(m) - 1;
            m &= m - 1;
            int cy = py, cx = px + bpos;
            while (cx >= w2) { cx -= w2; ++cy; }
            const uint8_t* sc = score + (cy + 1) * sp + cx + 2;
            const int s = sc[0];
            const int nb = __vimax3_s32(__vimax3_s32(sc[-sp - 1], sc[-sp], sc[-sp + 1]),
                                        __vimax3_s32(sc[sp - 1], sc[sp], sc[sp + 1]), max((int)sc[-1], (int)sc[1]));
            if (s > nb) {
                keepMin |= 1u << bpos;
                if (s > iniTh) keepIni |= 1u << bpos;
            }
        }
    }
    const int anyIni = __syncthreads_or(keepIni != 0);
    const uint32_t keep = anyIni ? keepIni : keepMin;
    const int cnt = __popc(keep);
    int incl = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { int t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
    if (lane == 31) warpSum[wid] = incl;
    __syncthreads();
    if (wid == 0) {
        int v = lane < kThreads / 32 ? warpSum[lane] : 0;
        int iv = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { int t = __shfl_up_sync(0xffffffffu, iv, o); if (lane >= o) iv += t; }
        if (lane < kThreads / 32) warpSum[lane] = iv - v;
        if (lane == 31) total = iv;
    }
    __syncthreads();
    int rank = warpSum[wid] + incl - cnt;
    uint32_t* out = cand + (size_t)frame * g->candCapFrame + c.outOffset;
    uint32_t m = keep;
    while (m) {
        const int b = __ffs(m) - 1;
        m &= m - 1;
        const int p = 2 * pb + b;
        const int py = p / w2, px = p - py * w2;
        const int s = score[(py + 1) * sp + px + 2];
        out[rank++] = (uint32_t)px | ((uint32_t)py << 8) | ((uint32_t)(s - 1) << 16);
    }
    if (tid == 0) cellCount[(size_t)frame * g->nCells + cellId] = total;
}

// ---------------------------------------------------------------------------------------------------------------
// E6: cv::GaussianBlur 7x7 sigma 2 on 8U = separable fixed-point kernel [18,34,48,56,48,34,18]/256, no intermediate
// rounding, (sum + 2^15) >> 16, BORDER_REFLECT_101 at the edge of the (unpadded) level.
// One WARP blurs a strip of 120 x 30 pixels and keeps everything in registers: lane i owns the 4 pixels of one aligned
// 32-bit word (lanes 0 and 31 only supply the horizontal halo through shuffles), and the warp slides down the strip:
//   horizontal pass  4 px = 6 PRMT (unaligned byte windows) + 8 IDP.4A  -> h <= 65280 (16 bit)
//   vertical pass    consecutive rows are packed as u16 pairs (h[r-1], h[r]) so that one IDP.2A applies two taps;
//                    4 px = 4 PRMT + 12 IDP.2A + 4 IMAD (seventh tap, carries the rounding constant)
// ~13 instructions per pixel instead of ~45 with byte-wise shared-memory passes; no shared memory, no block barrier.
// grid (ceil(jobs/8), B), block 256 (8 independent warps).
constexpr int kBlurStripW = 120, kBlurStripH = 30;

__device__ __forceinline__ int reflect101_dev(int p, int n) {
    if (n == 1) return 0;
    while (p < 0 || p >= n) p = p < 0 ? -p : 2 * (n - 1) - p;
    return p;
}

// same for -3 <= p <= n+2 and n >= 4 (every pyramid level is >= 35 px, orbx_api.cu): one reflection suffices
__device__ __forceinline__ int reflect101_near(int p, int n) {
    p = abs(p);
    return p >= n ? 2 * (n - 1) - p : p;
}

struct BlurTiles { int start[kMaxLevels + 1]; int tilesX[kMaxLevels]; };

// the 4 bytes at columns x..x+3 of a row, BORDER_REFLECT_101 outside [0, w)
__device__ __forceinline__ uint32_t blur_load_word(const uint8_t* __restrict__ row, int x, int w, bool aligned) {
    if (aligned && x >= 0 && x + 3 < w) return *reinterpret_cast<const uint32_t*>(row + x);
    uint32_t v = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const int xx = reflect101_near(min(max(x + k, -3), w + 2), w);
        v |= (uint32_t)row[xx] << (8 * k);
    }
    return v;
}

__global__ void __launch_bounds__(256) k_blur7(const DevGeom* __restrict__ g, Level0 l0, BlurTiles bt, int frameBase) {
    const int lane = threadIdx.x & 31;
    const int job = blockIdx.x * 8 + (threadIdx.x >> 5);
    const int nl = g->nlevels;
    if (job >= bt.start[nl]) return;
    const int frame = frameBase + blockIdx.y;
    int l = 0;
    while (l + 1 < nl && job >= bt.start[l + 1]) ++l;
    const int t = job - bt.start[l];
    const int ty = t / bt.tilesX[l], tx = t - ty * bt.tilesX[l];
    const LevelGeom& L = g->lv[l];
    const int w = L.w, h = L.h;
    int pitch;
    const uint8_t* src = level_ptr(g, l0, l, frame, &pitch);
    const bool aligned = (((uintptr_t)src | (uintptr_t)pitch) & 3) == 0;
    const int x = tx * kBlurStripW + 4 * (lane - 1);           // first column of this lane's word
    const int y0 = ty * kBlurStripH;
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
    // the words of the next three strip rows are in flight while the current one is worked on (the loop is latency-bound
    // otherwise: one DRAM round trip per row)
    auto rowWord = [&](int r) { return blur_load_word(src + (size_t)reflect101_near(min(y0 + r - 3, h + 2), h) * pitch, x, w, aligned); };
    uint32_t Cq[3] = {rowWord(0), rowWord(1), rowWord(2)};
    for (int r0 = 0; r0 < rowsNeeded; r0 += 6) {
#pragma unroll
        for (int k = 0; k < 6; ++k) {
            const int r = r0 + k;                               // strip row r = image row y0 + r - 3
            const int yo = y0 + r - 6;                          // output row completed by this input row
            const uint32_t C = Cq[k % 3];
            Cq[k % 3] = rowWord(r + 3);
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
                // output row y = r-3 (strip rows): taps h[r-6..r] = pairs r-5, r-3, r-1 and h[r]
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

}  // namespace orbx
