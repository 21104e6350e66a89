// ORACLE — TEST INFRASTRUCTURE ONLY.  ctypes-facing C entry points around the reference's OWN extractor class:
// this file is linked with /root/reference/src/ORBextractor.cc compiled UNMODIFIED, in place, against the header shim in
// oracle/ref_shim/ (see cv_ref_shim.h for what is the reference's and what is the shim's).  Output: oracle/_ref/liborb_ref*.so.
// The signatures mirror oracle/cpp/orc_capi.cc so that tests can run the restatement and the reference side by side.
#include "ORBextractor.h"          // /root/reference/include/ORBextractor.h
#include <chrono>
#include <cstring>

namespace {

// protected members are reached through a derived class; the reference source is not touched
struct Probe : MAM3SLAM::ORBextractor {
    using MAM3SLAM::ORBextractor::ORBextractor;
    using MAM3SLAM::ORBextractor::ComputePyramid;
    using MAM3SLAM::ORBextractor::ComputeKeyPointsOctTree;
    using MAM3SLAM::ORBextractor::DistributeOctTree;
    using MAM3SLAM::ORBextractor::mnFeaturesPerLevel;
    using MAM3SLAM::ORBextractor::umax;
    std::vector<std::vector<cv::KeyPoint>> lastLevels;     // filled by ref_level_keypoints' pass
};

}  // namespace

extern "C" {

void* ref_extractor_create(int nfeatures, float scaleFactor, int nlevels, int iniTh, int minTh) {
    return new Probe(nfeatures, scaleFactor, nlevels, iniTh, minTh);
}
void ref_extractor_destroy(void* h) { delete (Probe*)h; }

void ref_extractor_tables(void* h, float* scale, float* inv, float* sigma2, float* invSigma2, int* perLevel, int* umax16) {
    Probe* e = (Probe*)h;
    const std::vector<float> s = e->GetScaleFactors(), i = e->GetInverseScaleFactors(), g = e->GetScaleSigmaSquares(),
                             ig = e->GetInverseScaleSigmaSquares();
    for (int l = 0; l < e->GetLevels(); ++l) {
        scale[l] = s[l]; inv[l] = i[l]; sigma2[l] = g[l]; invSigma2[l] = ig[l];
        perLevel[l] = e->mnFeaturesPerLevel[l];
    }
    for (int k = 0; k < 16; ++k) umax16[k] = e->umax[k];
}

// ORBextractor::operator() exactly as Frame::ExtractORB calls it (src/Frame.cc:418-425): image Mat, empty mask, lapping.
// kps: n x 6 floats (x, y, size, angle, response, octave); desc: n x 32.  Returns monoIndex (or -1).
int ref_extract(void* h, const uint8_t* img, int w, int hh, int stride, int lap0, int lap1,
                float* kps, uint8_t* desc, int cap, int* n_out) {
    Probe* e = (Probe*)h;
    cv::Mat image(hh, w, CV_8UC1, (void*)img, (size_t)stride);
    if (w <= 0 || hh <= 0) image = cv::Mat();
    std::vector<cv::KeyPoint> k;
    cv::Mat d;
    std::vector<int> lap = {lap0, lap1};
    const int mono = (*e)(image, cv::Mat(), k, d, lap);
    *n_out = (int)k.size();
    if (mono < 0 || (int)k.size() > cap) return mono;
    for (size_t i = 0; i < k.size(); ++i) {
        float* o = kps + 6 * i;
        o[0] = k[i].pt.x; o[1] = k[i].pt.y; o[2] = k[i].size; o[3] = k[i].angle; o[4] = k[i].response; o[5] = (float)k[i].octave;
    }
    for (int r = 0; r < d.rows; ++r) std::memcpy(desc + 32 * (size_t)r, d.ptr(r), 32);
    return mono;
}

double ref_extract_timed(void* h, const uint8_t* imgs, int nimg, int w, int hh, int stride, int lap0, int lap1, int iters, long* total_kp) {
    Probe* e = (Probe*)h;
    std::vector<cv::KeyPoint> k;
    std::vector<int> lap = {lap0, lap1};
    long tot = 0;
    auto t0 = std::chrono::steady_clock::now();
    for (int it = 0; it < iters; ++it) {
        cv::Mat image(hh, w, CV_8UC1, (void*)(imgs + (size_t)(it % nimg) * stride * hh), (size_t)stride);
        cv::Mat d;
        (*e)(image, cv::Mat(), k, d, lap);
        tot += (long)k.size();
    }
    auto t1 = std::chrono::steady_clock::now();
    if (total_kp) *total_kp = tot;
    return std::chrono::duration<double>(t1 - t0).count();
}

// mvImagePyramid of the last call (public member, include/ORBextractor.h:83)
void ref_level_size(void* h, int level, int* w, int* hh) {
    Probe* e = (Probe*)h; *w = e->mvImagePyramid[level].cols; *hh = e->mvImagePyramid[level].rows;
}
// which: 0 = the level as the reference sees it (w x h), 1 = with its 19-px ring (the parent buffer `temp`, :1177-1178)
int ref_level_pixels(void* h, int level, int which, uint8_t* out) {
    Probe* e = (Probe*)h;
    const cv::Mat& m = e->mvImagePyramid[level];
    const int b = which == 1 ? 19 : 0;
    const int w = m.cols + 2 * b, hh = m.rows + 2 * b;
    const uint8_t* base = m.data - (size_t)b * m.step - b;
    for (int y = 0; y < hh; ++y) std::memcpy(out + (size_t)y * w, base + (size_t)y * m.step, (size_t)w);
    return w * hh;
}

// ComputePyramid + ComputeKeyPointsOctTree alone (:1097-1100): the kept keypoints of every level in list order, with
// border offset, octave, size and angle, before descriptors and rescaling.  out: n x 4 floats (x, y, response, angle).
int ref_level_keypoints(void* h, const uint8_t* img, int w, int hh, int stride, int level, float* out, int cap) {
    Probe* e = (Probe*)h;
    if (img) {
        cv::Mat image(hh, w, CV_8UC1, (void*)img, (size_t)stride);
        e->ComputePyramid(image);
        e->lastLevels.clear();
        e->ComputeKeyPointsOctTree(e->lastLevels);
    }
    const std::vector<cv::KeyPoint>& v = e->lastLevels[level];
    for (int i = 0; i < (int)v.size() && i < cap; ++i) {
        out[4 * i] = v[i].pt.x; out[4 * i + 1] = v[i].pt.y; out[4 * i + 2] = v[i].response; out[4 * i + 3] = v[i].angle;
    }
    return (int)v.size();
}

// DistributeOctTree alone (:555-779): keys n x 3 ints (x, y, response) relative to the min border; out n x 3
int ref_distribute(const int* keys, int n, int minX, int maxX, int minY, int maxY, int N, int* out, int cap) {
    Probe e(1000, 1.2f, 8, 20, 7);
    std::vector<cv::KeyPoint> v(n);
    for (int i = 0; i < n; ++i) v[i] = cv::KeyPoint((float)keys[3 * i], (float)keys[3 * i + 1], 7.f, -1.f, (float)keys[3 * i + 2]);
    const int level = 0;
    std::vector<cv::KeyPoint> r = e.DistributeOctTree(v, minX, maxX, minY, maxY, N, level);
    for (int i = 0; i < (int)r.size() && i < cap; ++i) { out[3 * i] = (int)r[i].pt.x; out[3 * i + 1] = (int)r[i].pt.y; out[3 * i + 2] = (int)r[i].response; }
    return (int)r.size();
}

}  // extern "C"
