// Frame::ComputeStereoMatches (reference src/Frame.cc:811-981) on the device-resident results of
// two extractors (left / right image of a rectified pair).  Integer work throughout (Hamming distances, 11x11 SAD); the few
// float expressions are evaluated with separately rounded operations in the reference's order.
#pragma once
#include "orbx_kernels_image.cuh"

namespace orbx {

constexpr int ORBX_STEREO_TH_HIGH = 100;                            // ORBmatcher::TH_HIGH
constexpr int ORBX_STEREO_TH_ORB = (100 + 50) / 2;                   // thOrbDist = (TH_HIGH + TH_LOW) / 2 (src/Frame.cc:816)

struct StereoParams {
    float scale[kMaxLevels], invScale[kMaxLevels];
    int colsR[kMaxLevels];
    float mb, mbf;
};

__device__ __forceinline__ int hamming256(const uint32_t* __restrict__ a, const uint32_t* __restrict__ b) {
    int d = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) d += __popc(a[i] ^ b[i]);
    return d;
}

// One warp per left keypoint.  Candidate scan (:855-890): the row table vRowIndices[(int)vL] lists, in ascending iR, the
// right keypoints with floor(y - r) <= row <= ceil(y + r), r = 2 * scale[octave] (:825-838), so the lanes stride over all
// right keypoints with that predicate and the warp keeps the lexicographic minimum of (distance, iR) = the first minimum of
// the reference's strict '<' scan starting from TH_HIGH.  Sub-pixel refinement (:893-957): lane i < 11 evaluates the
// 11x11 SAD at shift i - 5.  sad[iL] = best SAD (the key of vDistIdx) or -1.
constexpr int kStereoWarps = 8;
__global__ void __launch_bounds__(kStereoWarps * 32) k_stereo_search(const DevGeom* __restrict__ gL, Level0 l0L,
        const DevGeom* __restrict__ gR, Level0 l0R, const orbx_keypoint* __restrict__ kpL, const uint32_t* __restrict__ descL,
        const int* __restrict__ cntL, const orbx_keypoint* __restrict__ kpR, const uint32_t* __restrict__ descR,
        const int* __restrict__ cntR, StereoParams P, float* __restrict__ uRight, float* __restrict__ depth, int* __restrict__ sad) {
    const int iL = blockIdx.x * kStereoWarps + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    const int N = cntL[0], Nr = cntR[0];
    if (iL >= N) return;
    if (lane == 0) { uRight[iL] = -1.0f; depth[iL] = -1.0f; sad[iL] = -1; }
    const orbx_keypoint kL = kpL[iL];
    const int levelL = kL.octave;
    const float uL = kL.x, vL = kL.y;
    const int row = (int)vL;
    const float maxD = __fdiv_rn(P.mbf, P.mb);
    const float minU = __fsub_rn(uL, maxD), maxU = uL;               // minD = 0
    if (maxU < 0) return;
    uint32_t dl[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) dl[i] = descL[(size_t)iL * 8 + i];
    unsigned best = ((unsigned)ORBX_STEREO_TH_HIGH << 20) | 0xfffffu;   // (dist, iR); a candidate must have dist < TH_HIGH
    for (int iR = lane; iR < Nr; iR += 32) {
        const orbx_keypoint kR = kpR[iR];
        const float r = __fmul_rn(2.0f, P.scale[kR.octave]);
        const int maxr = (int)ceilf(__fadd_rn(kR.y, r));
        const int minr = (int)floorf(__fsub_rn(kR.y, r));
        if (row < minr || row > maxr) continue;
        if (kR.octave < levelL - 1 || kR.octave > levelL + 1) continue;
        if (kR.x >= minU && kR.x <= maxU) {
            const unsigned key = ((unsigned)hamming256(dl, descR + (size_t)iR * 8) << 20) | (unsigned)iR;
            best = min(best, key);
        }
    }
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, s));
    const int bestDist = (int)(best >> 20);
    if (bestDist >= ORBX_STEREO_TH_ORB) return;                      // (TH_HIGH + TH_LOW) / 2 (:815,893)
    const int bestIdxR = (int)(best & 0xfffffu);
    const float uR0 = kpR[bestIdxR].x;
    const float sf = P.invScale[levelL];
    const float scaleduL = roundf(__fmul_rn(kL.x, sf));
    const float scaledvL = roundf(__fmul_rn(kL.y, sf));
    const float scaleduR0 = roundf(__fmul_rn(uR0, sf));
    const int w = 5, L = 5;
    const float iniu = scaleduR0 + L - w;
    const float endu = scaleduR0 + L + w + 1;
    if (iniu < 0 || endu >= (float)P.colsR[levelL]) return;
    int pl, pr;
    const uint8_t* imL = level_ptr(gL, l0L, levelL, 0, &pl);
    const uint8_t* imR = level_ptr(gR, l0R, levelL, 0, &pr);
    const uint8_t* IL = imL + (size_t)((int)(scaledvL - w)) * pl + (int)(scaleduL - w);
    int sum = 0x7fffffff;
    if (lane <= 2 * L) {
        const int incR = lane - L;
        const uint8_t* IR = imR + (size_t)((int)(scaledvL - w)) * pr + (int)(scaleduR0 + incR - w);
        sum = 0;
        for (int y = 0; y < 2 * w + 1; ++y)
#pragma unroll
            for (int x = 0; x < 2 * w + 1; ++x) sum += abs((int)IL[y * pl + x] - (int)IR[y * pr + x]);
    }
    // first minimum over ascending incR (:931-940): lexicographic (sum, lane)
    unsigned long long k2 = ((unsigned long long)(unsigned)sum << 8) | (unsigned)lane;
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) {
        const unsigned long long o = __shfl_xor_sync(0xffffffffu, k2, s);
        k2 = o < k2 ? o : k2;
    }
    const int bl = (int)(k2 & 0xff);
    const int bestincR = bl - L;
    const int bestSad = (int)(k2 >> 8);
    const float dist1 = (float)__shfl_sync(0xffffffffu, sum, max(bl - 1, 0));
    const float dist2 = (float)bestSad;
    const float dist3 = (float)__shfl_sync(0xffffffffu, sum, min(bl + 1, 2 * L));
    if (bestincR == -L || bestincR == L) return;
    if (lane != 0) return;
    const float den = __fmul_rn(2.0f, __fsub_rn(__fadd_rn(dist1, dist3), __fmul_rn(2.0f, dist2)));
    const float deltaR = __fdiv_rn(__fsub_rn(dist1, dist3), den);
    if (deltaR < -1 || deltaR > 1) return;
    float bestuR = __fmul_rn(P.scale[levelL], __fadd_rn(__fadd_rn(scaleduR0, (float)bestincR), deltaR));
    float disparity = __fsub_rn(uL, bestuR);
    if (disparity >= 0 && disparity < maxD) {
        if (disparity <= 0) { disparity = 0.01f; bestuR = (float)((double)uL - 0.01); }
        depth[iL] = __fdiv_rn(P.mbf, disparity);
        uRight[iL] = bestuR;
        sad[iL] = bestSad;
    }
}

// The outlier filter at the end (:964-980): median = key of the element at position size/2 of the sorted (SAD, iL) list,
// thDist = 1.5f * 1.4f * median, every match with SAD >= thDist is removed.  One CTA; the order statistic is found by rank
// counting (N <= a few thousand).  kept[0] = matches left.
__global__ void __launch_bounds__(1024) k_stereo_filter(const int* __restrict__ cntL, const int* __restrict__ sad,
                                                        float* __restrict__ uRight, float* __restrict__ depth, int* __restrict__ kept) {
    __shared__ int sTotal, sMedian, sKept;
    const int N = cntL[0];
    if (threadIdx.x == 0) { sTotal = 0; sMedian = -1; sKept = 0; }
    __syncthreads();
    int mine = 0;
    for (int i = threadIdx.x; i < N; i += blockDim.x) mine += sad[i] >= 0;
    if (mine) atomicAdd(&sTotal, mine);
    __syncthreads();
    const int total = sTotal;
    if (total == 0) { if (threadIdx.x == 0) kept[0] = 0; return; }
    const int want = total / 2;
    for (int i = threadIdx.x; i < N; i += blockDim.x) {
        const int s = sad[i];
        if (s < 0) continue;
        int rank = 0;
        for (int j = 0; j < N; ++j) {
            const int t = sad[j];
            rank += (t >= 0) && (t < s || (t == s && j < i));
        }
        if (rank == want) sMedian = s;
    }
    __syncthreads();
    const float thDist = __fmul_rn(1.5f * 1.4f, (float)sMedian);
    int k = 0;
    for (int i = threadIdx.x; i < N; i += blockDim.x) {
        const int s = sad[i];
        if (s < 0) continue;
        if ((float)s < thDist) ++k;
        else { uRight[i] = -1.0f; depth[i] = -1.0f; }
    }
    if (k) atomicAdd(&sKept, k);
    __syncthreads();
    if (threadIdx.x == 0) kept[0] = sKept;
}

}  // namespace orbx
