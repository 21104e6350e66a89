// Frame-side kernels next to the matcher: keypoint undistortion and the DBoW2 vocabulary-tree descent.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace orbm {

// ---- Frame::UndistortKeyPoints / ComputeImageBounds (src/Frame.cc:747-812) ----------------------------------------------
// cv::undistortPoints(pts, pts, K, distCoef, Mat(), K): normalise, five fixed-point iterations of the inverse Brown
// model, re-project with K.  OpenCV evaluates this in double without contraction; every operation below is an explicit
// round-to-nearest intrinsic so that ptxas cannot fuse a multiply into an add (tests pin the result bit-exact against
// cv2.undistortPoints).
struct UndistortParams {
    double k[12];
    double fx, fy, cx, cy, ifx, ify;
};

__device__ __forceinline__ double dm(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ double da(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ double ds(double a, double b) { return __dsub_rn(a, b); }

__device__ __forceinline__ float2 undistort_one(float uf, float vf, const UndistortParams& P) {
    const double u = uf, v = vf;
    double x = dm(ds(u, P.cx), P.ifx), y = dm(ds(v, P.cy), P.ify);
    const double x0 = x, y0 = y;
    const double* k = P.k;
#pragma unroll 1
    for (int j = 0; j < 5; ++j) {
        const double r2 = da(dm(x, x), dm(y, y));
        const double num = da(1.0, dm(da(dm(da(dm(k[7], r2), k[6]), r2), k[5]), r2));
        const double den = da(1.0, dm(da(dm(da(dm(k[4], r2), k[1]), r2), k[0]), r2));
        const double icdist = __ddiv_rn(num, den);
        if (icdist < 0) { x = x0; y = y0; break; }
        // 2*k2*x*y + k3*(r2 + 2*x*x) + k8*r2 + k9*r2*r2   (left to right, as the C++ expression associates)
        const double dX = da(da(da(dm(dm(dm(2.0, k[2]), x), y), dm(k[3], da(r2, dm(dm(2.0, x), x)))), dm(k[8], r2)), dm(dm(k[9], r2), r2));
        const double dY = da(da(da(dm(k[2], da(r2, dm(dm(2.0, y), y))), dm(dm(dm(2.0, k[3]), x), y)), dm(k[10], r2)), dm(dm(k[11], r2), r2));
        x = dm(ds(x0, dX), icdist);
        y = dm(ds(y0, dY), icdist);
    }
    // RR = K * I: xx = fx*x + 0*y + cx, ww = 1
    const double xx = da(da(dm(P.fx, x), dm(0.0, y)), P.cx);
    const double yy = da(da(dm(0.0, x), dm(P.fy, y)), P.cy);
    return make_float2(__double2float_rn(xx), __double2float_rn(yy));
}

// generic strided form: record r of frame f lives at in + (f*frameStride + r)*recBytes; the first 8 bytes are pt.x, pt.y.
// When copyRecord != 0 the rest of the record (a 28-byte cv::KeyPoint image) is copied as well (mvKeysUn[i] = mvKeys[i]
// with pt replaced, src/Frame.cc:772-779).
__global__ void __launch_bounds__(256) k_undistort(const uint8_t* __restrict__ in, uint8_t* __restrict__ out, int recBytes,
                                                   const int* __restrict__ counts, int countStride, int frameStride, int nfixed,
                                                   int copyRecord, UndistortParams P) {
    const int f = blockIdx.y;
    const int n = counts ? min(counts[(size_t)f * countStride], frameStride) : nfixed;
    for (int r = blockIdx.x * blockDim.x + threadIdx.x; r < n; r += gridDim.x * blockDim.x) {
        const size_t at = ((size_t)f * frameStride + r) * recBytes;
        const float* src = reinterpret_cast<const float*>(in + at);
        float* dst = reinterpret_cast<float*>(out + at);
        const float2 o = undistort_one(src[0], src[1], P);
        if (copyRecord) {
            for (int w = 2; w < recBytes / 4; ++w) dst[w] = src[w];
        }
        dst[0] = o.x; dst[1] = o.y;
    }
}

// ---- DBoW2 TemplatedVocabulary::transform, per-feature descent (TemplatedVocabulary.h:1218-1260) ---------------------------
// One thread per feature.  The node table (32-byte descriptors) is read through L2: ORBvoc (k = 10, L = 6, 1.1 M nodes) is
// 35 MB and stays resident in the 126 MB L2 after the first frames.  Ties go to the first child (strict '<').
struct VocabDev {
    const int* childOff; const int* children; const uint4* desc; const int* wordId; const double* weight;
    int nnodes, levels, maxDepth;
};

__global__ void __launch_bounds__(128) k_bow_descend(VocabDev V, const uint4* __restrict__ desc, const int* __restrict__ counts,
                                                     int countStride, int frameStride, int nfixed, int levelsup,
                                                     int* __restrict__ featWord, int* __restrict__ featNode) {
    const int f = blockIdx.y;
    const int n = counts ? min(counts[(size_t)f * countStride], frameStride) : nfixed;
    const int nidLevel = V.levels - levelsup;
    for (int r = blockIdx.x * blockDim.x + threadIdx.x; r < n; r += gridDim.x * blockDim.x) {
        const size_t row = (size_t)f * frameStride + r;
        const uint4 a = desc[2 * row], b = desc[2 * row + 1];
        int id = 0, nid = 0, level = 0;
        int cb = __ldg(V.childOff), ce = __ldg(V.childOff + 1);
        while (ce > cb && level < V.maxDepth) {
            ++level;
            int best = 0x7fffffff, bestId = 0;
            // children in batches of 4: the loads of one batch are independent, the compare order stays sequential
            for (int c = cb; c < ce; c += 4) {
                int ids[4]; uint4 p[4], q[4];
#pragma unroll
                for (int t = 0; t < 4; ++t) ids[t] = (c + t < ce) ? __ldg(V.children + c + t) : -1;
#pragma unroll
                for (int t = 0; t < 4; ++t) {
                    if (ids[t] >= 0) { p[t] = __ldg(V.desc + 2 * (size_t)ids[t]); q[t] = __ldg(V.desc + 2 * (size_t)ids[t] + 1); }
                }
#pragma unroll
                for (int t = 0; t < 4; ++t) {
                    if (ids[t] < 0) continue;
                    const int d = __popc(a.x ^ p[t].x) + __popc(a.y ^ p[t].y) + __popc(a.z ^ p[t].z) + __popc(a.w ^ p[t].w) +
                                  __popc(b.x ^ q[t].x) + __popc(b.y ^ q[t].y) + __popc(b.z ^ q[t].z) + __popc(b.w ^ q[t].w);
                    if (d < best) { best = d; bestId = ids[t]; }
                }
            }
            id = bestId;
            if (level == nidLevel) nid = id;
            cb = __ldg(V.childOff + id); ce = __ldg(V.childOff + id + 1);
        }
        const int w = __ldg(V.wordId + id);
        featWord[row] = (__ldg(V.weight + id) > 0.0) ? w : -1 - w;      // stopped words (weight 0) are reported as -(id+1)
        featNode[row] = nid;
    }
}

}  // namespace orbm
