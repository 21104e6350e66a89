// ORACLE — TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product path; only tests/,
// __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may build, load or call it.
//
// CPU restatement of the OpenCV pixel primitives that the reference's ORB front end calls
// (OpenCV itself is an un-vendored dependency: /root/reference/CMakeLists.txt:42, ">= 4.4", the Dockerfile
// resolves it to Ubuntu 22.04's 4.5.4).  The arithmetic is the published OpenCV 4.x algorithm for 8-bit
// images; parity is pinned against cv2 4.13.0 by tests/test_oracle_vs_cv2.py and by the committed golden
// vectors under tests/golden/ (generated with cv2 by tools/gen_golden.py).
//
// call sites in the reference:
//   cv::resize(INTER_LINEAR)              src/ORBextractor.cc:1183
//   cv::copyMakeBorder(REFLECT_101)       src/ORBextractor.cc:1185,1190
//   cv::FAST(th, nonmax=true)             src/ORBextractor.cc:826,845
//   cv::GaussianBlur(7x7, sigma 2)        src/ORBextractor.cc:1133
//   cv::fastAtan2                         src/ORBextractor.cc:102
//   cvRound                               src/ORBextractor.cc:80,114,118-119,441,1175
#pragma once
#include <cstdint>
#include <cmath>
#include <cfloat>
#include <vector>
#include <cstring>
#include <algorithm>

namespace orc {

// cvRound(float/double): round-half-to-even (SSE cvtss2si under the default rounding mode).
static inline int cv_round(float v) { return (int)lrintf(v); }
static inline int cv_round(double v) { return (int)lrint(v); }
static inline int cv_floor(float v) { int i = (int)v; return i - (i > v); }
static inline int cv_ceil(float v) { int i = (int)v; return i + (i < v); }

// BORDER_REFLECT_101 index map (gfedcb|abcdefgh|gfedcba), valid for |overshoot| < n.
static inline int reflect101(int p, int n) {
    if (n == 1) return 0;
    while (p < 0 || p >= n) {
        if (p < 0) p = -p;
        else p = 2 * (n - 1) - p;
    }
    return p;
}

// ---- 8U bilinear resize (cv::resize INTER_LINEAR, 11-bit fixed-point coefficients) -------------------------
struct LinearTab {
    std::vector<int> ofs;      // source index of the left/top tap
    std::vector<short> coef;   // 2 per destination index, scaled by 2048
};

static inline LinearTab make_linear_tab(int ssize, int dsize) {
    LinearTab t;
    t.ofs.resize(dsize);
    t.coef.resize(2 * (size_t)dsize);
    const double inv_scale = (double)dsize / ssize;
    const double scale = 1.0 / inv_scale;
    for (int d = 0; d < dsize; ++d) {
        float f = (float)((d + 0.5) * scale - 0.5);
        int s = cv_floor(f);
        f -= s;
        if (s < 0) { f = 0.f; s = 0; }
        if (s >= ssize - 1) { f = 0.f; s = ssize - 1; }
        t.ofs[d] = s;
        t.coef[2 * d] = (short)cv_round((1.f - f) * 2048.f);
        t.coef[2 * d + 1] = (short)cv_round(f * 2048.f);
    }
    return t;
}

// Vertical taps: OpenCV does not zero the fraction at the ends, it clamps the ROW indices instead.
static inline LinearTab make_linear_tab_rows(int ssize, int dsize) {
    LinearTab t;
    t.ofs.resize(dsize);
    t.coef.resize(2 * (size_t)dsize);
    const double inv_scale = (double)dsize / ssize;
    const double scale = 1.0 / inv_scale;
    for (int d = 0; d < dsize; ++d) {
        float f = (float)((d + 0.5) * scale - 0.5);
        int s = cv_floor(f);
        f -= s;
        t.ofs[d] = s;   // may be -1 or ssize-1; callers clamp s and s+1 separately
        t.coef[2 * d] = (short)cv_round((1.f - f) * 2048.f);
        t.coef[2 * d + 1] = (short)cv_round(f * 2048.f);
    }
    return t;
}

static inline void resize_linear_8u(const uint8_t* src, int sw, int sh, int sstride,
                                    uint8_t* dst, int dw, int dh, int dstride) {
    // exact 2x decimation is rerouted by cv::resize to the INTER_AREA fast path: rounded 2x2 mean
    if (sw == 2 * dw && sh == 2 * dh) {
        for (int y = 0; y < dh; ++y) {
            const uint8_t* r0 = src + (size_t)(2 * y) * sstride;
            const uint8_t* r1 = r0 + sstride;
            for (int x = 0; x < dw; ++x)
                dst[(size_t)y * dstride + x] = (uint8_t)((r0[2 * x] + r0[2 * x + 1] + r1[2 * x] + r1[2 * x + 1] + 2) >> 2);
        }
        return;
    }
    LinearTab tx = make_linear_tab(sw, dw), ty = make_linear_tab_rows(sh, dh);
    std::vector<int> row0(dw), row1(dw);
    int have0 = -1000000, have1 = -1000000;
    auto hpass = [&](int sy, std::vector<int>& out) {
        const uint8_t* S = src + (size_t)sy * sstride;
        for (int x = 0; x < dw; ++x) {
            int s = tx.ofs[x];
            int s1 = std::min(s + 1, sw - 1);
            out[x] = S[s] * tx.coef[2 * x] + S[s1] * tx.coef[2 * x + 1];
        }
    };
    for (int y = 0; y < dh; ++y) {
        int s0 = std::min(std::max(ty.ofs[y], 0), sh - 1);
        int s1 = std::min(std::max(ty.ofs[y] + 1, 0), sh - 1);
        if (s0 == have1) { std::swap(row0, row1); std::swap(have0, have1); }
        if (s0 != have0) { hpass(s0, row0); have0 = s0; }
        if (s1 == have0) { row1 = row0; have1 = s1; }
        else if (s1 != have1) { hpass(s1, row1); have1 = s1; }
        const int b0 = ty.coef[2 * y], b1 = ty.coef[2 * y + 1];
        uint8_t* D = dst + (size_t)y * dstride;
        for (int x = 0; x < dw; ++x)
            D[x] = (uint8_t)((((b0 * (row0[x] >> 4)) >> 16) + ((b1 * (row1[x] >> 4)) >> 16) + 2) >> 2);
    }
}

// ---- copyMakeBorder(..., BORDER_REFLECT_101) ------------------------------------------------------------------
// dst is (w+2b) x (h+2b); the interior may alias src rows already placed in dst.
static inline void make_border_101(const uint8_t* src, int w, int h, int sstride,
                                   uint8_t* dst, int dstride, int b) {
    for (int y = -b; y < h + b; ++y) {
        const uint8_t* S = src + (size_t)reflect101(y, h) * sstride;
        uint8_t* D = dst + (size_t)(y + b) * dstride;
        // interior first (memmove: dst interior can be src itself)
        if (D + b != S) std::memmove(D + b, S, w);
        for (int x = -b; x < 0; ++x) D[x + b] = S[reflect101(x, w)];
        for (int x = w; x < w + b; ++x) D[x + b] = S[reflect101(x, w)];
    }
}

// ---- GaussianBlur 7x7 sigma=2 on 8U: fixed-point 8.8 kernel, exact ------------------------------------------
static inline void gaussian7_8u(const uint8_t* src, int w, int h, int sstride, uint8_t* dst, int dstride) {
    static const int K[7] = {18, 34, 48, 56, 48, 34, 18};
    std::vector<uint16_t> tmp((size_t)w * h);
    for (int y = 0; y < h; ++y) {
        const uint8_t* S = src + (size_t)y * sstride;
        for (int x = 0; x < w; ++x) {
            int acc = 0;
            for (int k = -3; k <= 3; ++k) acc += K[k + 3] * S[reflect101(x + k, w)];
            tmp[(size_t)y * w + x] = (uint16_t)acc;          // <= 255*256
        }
    }
    for (int y = 0; y < h; ++y) {
        const uint16_t* R[7];
        for (int k = -3; k <= 3; ++k) R[k + 3] = &tmp[(size_t)reflect101(y + k, h) * w];
        uint8_t* D = dst + (size_t)y * dstride;
        for (int x = 0; x < w; ++x) {
            uint32_t acc = 0;
            for (int k = 0; k < 7; ++k) acc += (uint32_t)K[k] * R[k][x];
            D[x] = (uint8_t)((acc + 32768u) >> 16);
        }
    }
}

// ---- cv::fastAtan2 (scalar path), degrees in [0,360) ---------------------------------------------------------
static inline float fast_atan2_deg(float y, float x) {
    const float p1 = 0.9997878412794807f * (float)(180 / M_PI);
    const float p3 = -0.3258083974640975f * (float)(180 / M_PI);
    const float p5 = 0.1555786518463281f * (float)(180 / M_PI);
    const float p7 = -0.04432655554792128f * (float)(180 / M_PI);
    float ax = std::fabs(x), ay = std::fabs(y);
    float a, c, c2;
    if (ax >= ay) {
        c = ay / (ax + (float)DBL_EPSILON);
        c2 = c * c;
        a = (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
    } else {
        c = ax / (ay + (float)DBL_EPSILON);
        c2 = c * c;
        a = 90.f - (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
    }
    if (x < 0) a = 180.f - a;
    if (y < 0) a = 360.f - a;
    return a;
}

// ---- FAST-9/16 with 3x3 non-max suppression (cv::FAST(img, kps, th, true)) ------------------------------------
struct FastPt { int x, y, score; };   // score == KeyPoint::response

// ring offsets, index 0 = (0,+3), clockwise as OpenCV lays them out
static const int kRingDx[16] = {0, 1, 2, 3, 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1};
static const int kRingDy[16] = {3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1, 0, 1, 2, 3};

// s(p) = max over the 16 cyclic 9-arcs of max(min(v - ring), min(ring - v)); p is a corner at th iff s > th,
// and cv::FAST stores s-1 as the score.
static inline int fast_arc_strength(const uint8_t* p, const int* off) {
    int d[25];
    const int v = p[0];
    for (int k = 0; k < 16; ++k) d[k] = v - p[off[k]];
    for (int k = 16; k < 25; ++k) d[k] = d[k - 16];
    int best = 0;
    for (int k = 0; k < 16; ++k) {
        int mn = d[k], mx = d[k];
        for (int j = 1; j < 9; ++j) { mn = std::min(mn, d[k + j]); mx = std::max(mx, d[k + j]); }
        best = std::max(best, std::max(mn, -mx));
    }
    return best;
}

static inline void fast9_nms(const uint8_t* img, int w, int h, int stride, int th, std::vector<FastPt>& out) {
    out.clear();
    if (w < 7 || h < 7) return;
    int off[16];
    for (int k = 0; k < 16; ++k) off[k] = kRingDy[k] * stride + kRingDx[k];
    std::vector<uint8_t> sc((size_t)w * h, 0);   // stored score (s-1) where corner, else 0
    for (int y = 3; y < h - 3; ++y) {
        const uint8_t* row = img + (size_t)y * stride;
        for (int x = 3; x < w - 3; ++x) {
            const uint8_t* p = row + x;
            const int v = p[0];
            // quick reject: any 9-arc contains 2 opposite... at least two of the four compass points
            int hi = 0, lo = 0;
            for (int k = 0; k < 16; k += 4) { int r = p[off[k]]; hi += (r > v + th); lo += (r < v - th); }
            if (hi < 2 && lo < 2) continue;
            int s = fast_arc_strength(p, off);
            if (s > th) sc[(size_t)y * w + x] = (uint8_t)(s - 1);
        }
    }
    for (int y = 3; y < h - 3; ++y)
        for (int x = 3; x < w - 3; ++x) {
            const uint8_t* c = &sc[(size_t)y * w + x];
            int s = c[0];
            if (s == 0) continue;   // not a corner (or th==0 corner with stored score 0: 0 > 0 never holds)
            if (s > c[-1] && s > c[1] && s > c[-w - 1] && s > c[-w] && s > c[-w + 1] &&
                s > c[w - 1] && s > c[w] && s > c[w + 1])
                out.push_back({x, y, s});
        }
}

}  // namespace orc
