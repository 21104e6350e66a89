// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_primitives.h).  ctypes-facing C entry points of the CPU oracle.
#include "orc_extractor.h"
#include "orc_matcher.h"
#include <chrono>

using namespace orc;

extern "C" {

// ---- primitives (checked one by one against cv2) ---------------------------------------------------------------
void orc_resize_linear(const uint8_t* src, int sw, int sh, int sstride, uint8_t* dst, int dw, int dh, int dstride) {
    resize_linear_8u(src, sw, sh, sstride, dst, dw, dh, dstride);
}
void orc_border101(const uint8_t* src, int w, int h, int sstride, uint8_t* dst, int dstride, int b) {
    make_border_101(src, w, h, sstride, dst, dstride, b);
}
void orc_gaussian7(const uint8_t* src, int w, int h, int sstride, uint8_t* dst, int dstride) {
    gaussian7_8u(src, w, h, sstride, dst, dstride);
}
float orc_fast_atan2(float y, float x) { return fast_atan2_deg(y, x); }
// returns the number of keypoints; writes up to cap (x,y,score) triples
int orc_fast9_nms(const uint8_t* img, int w, int h, int stride, int th, int* xys, int cap) {
    std::vector<FastPt> v;
    fast9_nms(img, w, h, stride, th, v);
    for (int i = 0; i < (int)v.size() && i < cap; ++i) { xys[3 * i] = v[i].x; xys[3 * i + 1] = v[i].y; xys[3 * i + 2] = v[i].score; }
    return (int)v.size();
}
// std::sort with the reference's compareNodes ordering on (size, ULx) pairs; perm receives the resulting order
// of the input indices.  Used to validate the device-side emulation of libstdc++'s introsort.
void orc_sort_nodes(const int* size, const int* ulx, int n, int* perm) {
    struct E { int s, x, i; };
    std::vector<E> v(n);
    for (int i = 0; i < n; ++i) v[i] = {size[i], ulx[i], i};
    std::sort(v.begin(), v.end(), [](E& a, E& b) { return a.s != b.s ? a.s < b.s : a.x < b.x; });
    for (int i = 0; i < n; ++i) perm[i] = v[i].i;
}

// ---- extractor ---------------------------------------------------------------------------------------------------
void* orc_extractor_create(int nfeatures, float scaleFactor, int nlevels, int iniTh, int minTh) {
    return new Extractor(nfeatures, scaleFactor, nlevels, iniTh, minTh);
}
void orc_extractor_destroy(void* h) { delete (Extractor*)h; }

void orc_extractor_tables(void* h, float* scale, float* inv, float* sigma2, float* invSigma2, int* perLevel, int* umax16) {
    Extractor* e = (Extractor*)h;
    for (int i = 0; i < e->nlevels; ++i) {
        scale[i] = e->scale[i]; inv[i] = e->invScale[i]; sigma2[i] = e->sigma2[i]; invSigma2[i] = e->invSigma2[i];
        perLevel[i] = e->featuresPerLevel[i];
    }
    for (int i = 0; i < 16; ++i) umax16[i] = e->umax[i];
}

// kps: n x 6 floats (x, y, size, angle, response, octave-as-float); desc: n x 32.  Returns monoIndex (or -1),
// *n_out = number of keypoints (may exceed cap: then nothing is written).
int orc_extract(void* h, const uint8_t* img, int w, int hh, int stride, int lap0, int lap1,
                float* kps, uint8_t* desc, int cap, int* n_out) {
    Extractor* e = (Extractor*)h;
    std::vector<KeyPoint> k; std::vector<uint8_t> d;
    int mono = e->extract(img, w, hh, stride, lap0, lap1, k, d);
    *n_out = (int)k.size();
    if (mono < 0 || (int)k.size() > cap) return mono;
    for (size_t i = 0; i < k.size(); ++i) {
        float* o = kps + 6 * i;
        o[0] = k[i].x; o[1] = k[i].y; o[2] = k[i].size; o[3] = k[i].angle; o[4] = k[i].response; o[5] = (float)k[i].octave;
    }
    if (!d.empty()) std::memcpy(desc, d.data(), d.size());
    return mono;
}

// timing helper for the CPU baseline: runs extract() `iters` times on the same frame set, returns seconds
double orc_extract_timed(void* h, const uint8_t* imgs, int nimg, int w, int hh, int stride, int lap0, int lap1, int iters, long* total_kp) {
    Extractor* e = (Extractor*)h;
    std::vector<KeyPoint> k; std::vector<uint8_t> d;
    long tot = 0;
    auto t0 = std::chrono::steady_clock::now();
    for (int it = 0; it < iters; ++it) {
        const uint8_t* img = imgs + (size_t)(it % nimg) * stride * hh;
        e->extract(img, w, hh, stride, lap0, lap1, k, d);
        tot += (long)k.size();
    }
    auto t1 = std::chrono::steady_clock::now();
    if (total_kp) *total_kp = tot;
    return std::chrono::duration<double>(t1 - t0).count();
}

void orc_level_size(void* h, int level, int* w, int* hh) {
    Extractor* e = (Extractor*)h; *w = e->levels[level].w; *hh = e->levels[level].h;
}
// which: 0 = unpadded level (w x h), 1 = padded mvImagePyramid buffer ((w+38) x (h+38)), 2 = blurred level (w x h)
int orc_level_pixels(void* h, int level, int which, uint8_t* out) {
    Extractor* e = (Extractor*)h; const Extractor::Level& L = e->levels[level];
    if (which == 1) { std::memcpy(out, L.padded.data(), L.padded.size()); return (int)L.padded.size(); }
    if (which == 2) { if (L.blurred.empty()) return 0; std::memcpy(out, L.blurred.data(), L.blurred.size()); return (int)L.blurred.size(); }
    for (int y = 0; y < L.h; ++y) std::memcpy(out + (size_t)y * L.w, L.roi() + (size_t)y * L.pstride, L.w);
    return L.w * L.h;
}
// which: 0 = FAST candidates before culling (x,y relative to minBorder, response), 1 = kept keypoints of the level
// (x,y level coords incl. border offset, response, angle).  out: n x 4 floats.  Returns n.
int orc_level_points(void* h, int level, int which, float* out, int cap) {
    Extractor* e = (Extractor*)h; const Extractor::Level& L = e->levels[level];
    const std::vector<KeyPoint>& v = which ? L.kept : L.cand;
    for (int i = 0; i < (int)v.size() && i < cap; ++i) {
        out[4 * i] = v[i].x; out[4 * i + 1] = v[i].y; out[4 * i + 2] = v[i].response; out[4 * i + 3] = v[i].angle;
    }
    return (int)v.size();
}
// IC_Angle (src/ORBextractor.cc:76-103) at n integer positions of a plain image (positions must be >= 15 px inside)
void orc_ic_angles(const uint8_t* img, int w, int h, int stride, const int* xy, int n, float* out) {
    Extractor e(1000, 1.2f, 8, 20, 7);
    for (int i = 0; i < n; ++i) out[i] = e.icAngleAt(img + (size_t)xy[2 * i + 1] * stride + xy[2 * i], stride);
    (void)w; (void)h;
}
// computeOrbDescriptor (src/ORBextractor.cc:106-146) on an already blurred plain image; kps n x 3 floats (x,y,angle)
void orc_brief(const uint8_t* blurred, int w, int h, int stride, const float* kps, int n, uint8_t* out) {
    Extractor e(1000, 1.2f, 8, 20, 7);
    for (int i = 0; i < n; ++i) {
        KeyPoint k{kps[3 * i], kps[3 * i + 1], 31.f, kps[3 * i + 2], 0.f, 0};
        e.orbDescriptor(k, blurred, stride, out + 32 * (size_t)i);
    }
    (void)w; (void)h;
}
long orc_sort_tie_count(void* h) { return ((Extractor*)h)->sortTieCount; }

// standalone quadtree: keys n x 3 ints (x, y, response) relative to the min border; returns kept count, out n x 3
int orc_distribute(int nfeatures_unused, const int* keys, int n, int minX, int maxX, int minY, int maxY, int N, int* out, int cap) {
    (void)nfeatures_unused;
    Extractor e(1000, 1.2f, 8, 20, 7);
    std::vector<KeyPoint> v(n);
    for (int i = 0; i < n; ++i) v[i] = KeyPoint{(float)keys[3 * i], (float)keys[3 * i + 1], 7.f, -1.f, (float)keys[3 * i + 2], 0};
    std::vector<KeyPoint> r = e.distributeOctTree(v, minX, maxX, minY, maxY, N);
    for (int i = 0; i < (int)r.size() && i < cap; ++i) { out[3 * i] = (int)r[i].x; out[3 * i + 1] = (int)r[i].y; out[3 * i + 2] = (int)r[i].response; }
    return (int)r.size();
}

}  // extern "C"

// ---- matcher ---------------------------------------------------------------------------------------------------------
namespace {
FrameView makeView(int n, const void* kps, const uint8_t* desc, const float* bounds) {
    FrameView f;
    f.build(n, (const KP*)kps, desc, bounds[0], bounds[1], bounds[2], bounds[3]);
    return f;
}
}  // namespace

extern "C" {

int orc_descriptor_distance(const uint8_t* a, const uint8_t* b) { return descriptor_distance(a, b); }

// grid as CSR (cell = ix*48+iy): off has 64*48+1 entries
void orc_grid_build(int n, const void* kps, const float* bounds, int* off, int* idx) {
    FrameView f = makeView(n, kps, nullptr, bounds);
    int o = 0;
    for (int c = 0; c < GRID_COLS * GRID_ROWS; ++c) { off[c] = o; for (int i : f.grid[c]) idx[o++] = i; }
    off[GRID_COLS * GRID_ROWS] = o;
}
int orc_features_in_area(int n, const void* kps, const float* bounds, float x, float y, float r, int minLevel, int maxLevel, int* out) {
    FrameView f = makeView(n, kps, nullptr, bounds);
    std::vector<int> v;
    f.features_in_area(x, y, r, minLevel, maxLevel, v);
    for (size_t i = 0; i < v.size(); ++i) out[i] = v[i];
    return (int)v.size();
}
void orc_three_maxima(const int* sizes, int L, int* ind) { ind[0] = ind[1] = ind[2] = -1; three_maxima(sizes, L, ind[0], ind[1], ind[2]); }
int orc_rot_bin(float a1, float a2) { return rot_bin(a1, a2); }

int orc_search_local_map(int n, const void* kps, const uint8_t* desc, const float* bounds, uint8_t* occupied,
                         const float* scaleFactors, int nq, const uint8_t* valid, const float* projX, const float* projY,
                         const int* level, const float* viewCos, const uint8_t* qdesc, float th, float nnratio, int* kpMatch,
                         const uint8_t* qocc) {
    FrameView f = makeView(n, kps, desc, bounds);
    return search_local_map(f, occupied, scaleFactors, nq, valid, projX, projY, level, viewCos, qdesc, th, nnratio, kpMatch, qocc);
}
int orc_search_frame(int n, const void* kps, const uint8_t* desc, const float* bounds, uint8_t* occupied,
                     const float* scaleFactors, int nq, const uint8_t* valid, const float* u, const float* v,
                     const int* octave, const float* qangle, const uint8_t* qdesc, float th, int thDist, int checkOri, int* kpMatch,
                     const uint8_t* qocc) {
    FrameView f = makeView(n, kps, desc, bounds);
    return search_frame(f, occupied, scaleFactors, nq, valid, u, v, octave, qangle, qdesc, th, thDist, checkOri != 0, kpMatch, qocc);
}
int orc_search_sim3(int n, const void* kps, const uint8_t* desc, const float* bounds, uint8_t* matched,
                    const float* scaleFactors, int nq, const uint8_t* valid, const float* u, const float* v,
                    const int* level, const uint8_t* qdesc, int th, float ratioHamming, int* kpMatch) {
    FrameView f = makeView(n, kps, desc, bounds);
    return search_sim3(f, matched, scaleFactors, nq, valid, u, v, level, qdesc, th, ratioHamming, kpMatch);
}
int orc_search_initialization(int n1, const void* kps1, const uint8_t* desc1, int n2, const void* kps2, const uint8_t* desc2,
                              const float* bounds, float* prevMatched, int windowSize, float nnratio, int checkOri, int* matches12) {
    FrameView f1 = makeView(n1, kps1, desc1, bounds), f2 = makeView(n2, kps2, desc2, bounds);
    return search_initialization(f1, f2, prevMatched, windowSize, nnratio, checkOri != 0, matches12);
}
int orc_search_bow_kf_frame(int n1, const void* kps1, const uint8_t* desc1, int nn1, const int* node1, const int* off1, const int* idx1,
                            const uint8_t* elig1, int n2, const void* kps2, const uint8_t* desc2, int nn2, const int* node2,
                            const int* off2, const int* idx2, float nnratio, int checkOri, int* matchF) {
    const float b[4] = {0, 0, 1, 1};
    FrameView f1 = makeView(n1, kps1, desc1, b), f2 = makeView(n2, kps2, desc2, b);
    FeatVec a{nn1, node1, off1, idx1}, c{nn2, node2, off2, idx2};
    return search_bow_kf_frame(f1, a, elig1, f2, c, nnratio, checkOri != 0, matchF);
}
int orc_search_bow_kf_kf(int n1, const void* kps1, const uint8_t* desc1, int nn1, const int* node1, const int* off1, const int* idx1,
                         const uint8_t* elig1, int n2, const void* kps2, const uint8_t* desc2, int nn2, const int* node2,
                         const int* off2, const int* idx2, const uint8_t* elig2, float nnratio, int checkOri, int* matches12) {
    const float b[4] = {0, 0, 1, 1};
    FrameView f1 = makeView(n1, kps1, desc1, b), f2 = makeView(n2, kps2, desc2, b);
    FeatVec a{nn1, node1, off1, idx1}, c{nn2, node2, off2, idx2};
    return search_bow_kf_kf(f1, a, elig1, f2, c, elig2, nnratio, checkOri != 0, matches12);
}
int orc_search_triangulation(int n1, const void* kps1, const uint8_t* desc1, int nn1, const int* node1, const int* off1, const int* idx1,
                             const uint8_t* free1, int n2, const void* kps2, const uint8_t* desc2, int nn2, const int* node2,
                             const int* off2, const int* idx2, const uint8_t* free2, const float* ep, const float* scaleFactors2,
                             const float* F12, const float* sigma2, int coarse, int checkOri, int* matches12) {
    const float b[4] = {0, 0, 1, 1};
    FrameView f1 = makeView(n1, kps1, desc1, b), f2 = makeView(n2, kps2, desc2, b);
    FeatVec a{nn1, node1, off1, idx1}, c{nn2, node2, off2, idx2};
    return search_triangulation(f1, a, free1, f2, c, free2, ep, scaleFactors2, F12, sigma2, coarse != 0, checkOri != 0, matches12);
}
void orc_fuse_candidates(int n, const void* kps, const uint8_t* desc, const float* bounds, const float* scaleFactors,
                         const float* invSigma2, int nq, const uint8_t* valid, const float* u, const float* v, const int* level,
                         const uint8_t* qdesc, float th, int chi2Gate, int* bestIdx, int* bestDist) {
    FrameView f = makeView(n, kps, desc, bounds);
    fuse_candidates(f, scaleFactors, invSigma2, nq, valid, u, v, level, qdesc, th, chi2Gate != 0, bestIdx, bestDist);
}
int orc_search_by_sim3(int n1, const void* kps1, const uint8_t* desc1, const float* bounds1, int n2, const void* kps2,
                       const uint8_t* desc2, const float* bounds2, const float* sf1, const float* sf2, const uint8_t* valid1,
                       const float* u12, const float* v12, const int* level12, const uint8_t* qdesc1, const uint8_t* valid2,
                       const float* u21, const float* v21, const int* level21, const uint8_t* qdesc2, float th, int* matches12) {
    FrameView f1 = makeView(n1, kps1, desc1, bounds1), f2 = makeView(n2, kps2, desc2, bounds2);
    return search_by_sim3(f1, f2, sf1, sf2, valid1, u12, v12, level12, qdesc1, valid2, u21, v21, level21, qdesc2, th, matches12);
}
int orc_stereo_matches(int N, const void* kL, const uint8_t* dL, int Nr, const void* kR, const uint8_t* dR, int nlevels,
                       const uint8_t* const* pyrL, const int* strideL, const uint8_t* const* pyrR, const int* strideR,
                       const int* colsR, const float* sf, const float* isf, float mb, float mbf, float* uRight, float* depth) {
    return stereo_matches(N, (const KP*)kL, dL, Nr, (const KP*)kR, dR, nlevels, pyrL, strideL, pyrR, strideR, colsR, sf, isf, mb,
                          mbf, uRight, depth);
}
void orc_knn(const uint8_t* q, int nq, const uint8_t* db, long ndb, int k, uint16_t* dist, int32_t* idx, int threads) {
    knn(q, nq, db, ndb, k, dist, idx, threads);
}
int orc_distinctive_descriptor(const uint8_t* desc, int n) { return distinctive_descriptor(desc, n); }
void orc_undistort_points(const float* xy, int n, const float* K, const float* dist, int ndist, float* out) {
    undistort_points(xy, n, K, dist, ndist, out);
}
void orc_bow_transform(int nnodes, int L, const int* childOff, const int* children, const uint8_t* ndesc, const int* wordId,
                       const double* weight, int weighting, int norm, const uint8_t* desc, int n, int levelsup, int* featWord,
                       int* featNode, int* nBow, int* bowWord, double* bowValue, int* nFv, int* fvNode, int* fvOff, int* fvIdx) {
    VocabView V{nnodes, L, childOff, children, ndesc, wordId, weight, weighting, norm};
    bow_transform(V, desc, n, levelsup, featWord, featNode, nBow, bowWord, bowValue, nFv, fvNode, fvOff, fvIdx);
}

}  // extern "C"
