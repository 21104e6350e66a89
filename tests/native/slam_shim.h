// Minimal stand-ins for the MAM3SLAM / Sophus / Eigen / DBoW2 types the ORBmatcher adapter touches, ONLY so that
// mam3slam_b200/cpp/ORBmatcher.h can be compiled and exercised without the reference tree (Eigen, Sophus, DBoW2 and OpenCV
// C++ headers are not installed here).  Member names and meanings follow the reference (include/Frame.h, KeyFrame.h,
// MapPoint.h, CameraModels/Pinhole.h); the arithmetic is plain float in the order written below, which the parity test
// (tests/test_cpp_adapter_gpu.py) repeats in numpy float32.  Test infrastructure, not product code.
#pragma once
#include <cmath>
#include <map>
#include <set>
#include <tuple>
#include <vector>
#include "cv_shim.h"

namespace shim {

struct Vec2 { float d[2]; float operator()(int i) const { return d[i]; } };
struct Vec3 {
    float d[3] = {0, 0, 0};
    Vec3() {}
    Vec3(float x, float y, float z) { d[0] = x; d[1] = y; d[2] = z; }
    float& operator()(int i) { return d[i]; }
    float operator()(int i) const { return d[i]; }
    float norm() const { return std::sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]); }
    float dot(const Vec3& o) const { return d[0] * o.d[0] + d[1] * o.d[1] + d[2] * o.d[2]; }
    Vec3 operator-(const Vec3& o) const { return Vec3(d[0] - o.d[0], d[1] - o.d[1], d[2] - o.d[2]); }
    Vec3 operator+(const Vec3& o) const { return Vec3(d[0] + o.d[0], d[1] + o.d[1], d[2] + o.d[2]); }
    Vec3 operator-() const { return Vec3(-d[0], -d[1], -d[2]); }
    Vec3 operator/(float s) const { return Vec3(d[0] / s, d[1] / s, d[2] / s); }
    Vec3 operator*(float s) const { return Vec3(d[0] * s, d[1] * s, d[2] * s); }
};
struct Mat3 {
    float m[3][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
    float& operator()(int r, int c) { return m[r][c]; }
    float operator()(int r, int c) const { return m[r][c]; }
    Mat3 transpose() const { Mat3 o; for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) o.m[r][c] = m[c][r]; return o; }
    Mat3 operator*(const Mat3& b) const {
        Mat3 o;
        for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) o.m[r][c] = m[r][0] * b.m[0][c] + m[r][1] * b.m[1][c] + m[r][2] * b.m[2][c];
        return o;
    }
    Vec3 operator*(const Vec3& v) const {
        return Vec3(m[0][0] * v(0) + m[0][1] * v(1) + m[0][2] * v(2), m[1][0] * v(0) + m[1][1] * v(1) + m[1][2] * v(2),
                    m[2][0] * v(0) + m[2][1] * v(1) + m[2][2] * v(2));
    }
    Mat3 inverse() const {                       // cofactor form
        const float a = m[0][0], b = m[0][1], c = m[0][2], d = m[1][0], e = m[1][1], f = m[1][2], g = m[2][0], h = m[2][1], i = m[2][2];
        const float det = a * (e * i - f * h) - b * (d * i - f * g) + c * (d * h - e * g);
        Mat3 o;
        o.m[0][0] = (e * i - f * h) / det; o.m[0][1] = (c * h - b * i) / det; o.m[0][2] = (b * f - c * e) / det;
        o.m[1][0] = (f * g - d * i) / det; o.m[1][1] = (a * i - c * g) / det; o.m[1][2] = (c * d - a * f) / det;
        o.m[2][0] = (d * h - e * g) / det; o.m[2][1] = (b * g - a * h) / det; o.m[2][2] = (a * e - b * d) / det;
        return o;
    }
};
struct SE3 {                                     // Sophus::SE3f
    Mat3 R; Vec3 t;
    SE3() {}
    SE3(const Mat3& R_, const Vec3& t_) : R(R_), t(t_) {}
    Vec3 operator*(const Vec3& p) const { return R * p + t; }
    SE3 operator*(const SE3& o) const { return SE3(R * o.R, R * o.t + t); }
    SE3 inverse() const { const Mat3 Rt = R.transpose(); return SE3(Rt, -(Rt * t)); }
    const Vec3& translation() const { return t; }
    Mat3 rotationMatrix() const { return R; }
};
struct Sim3 {                                    // Sophus::Sim3f
    float s = 1; Mat3 R; Vec3 t;
    Mat3 rotationMatrix() const { return R; }
    const Vec3& translation() const { return t; }
    float scale() const { return s; }
    Vec3 operator*(const Vec3& p) const { return (R * p) * s + t; }
    Sim3 inverse() const { Sim3 o; o.s = 1.f / s; o.R = R.transpose(); o.t = -(o.R * t) * o.s; return o; }
};
struct Camera {                                  // GeometricCamera / Pinhole (src/CameraModels/Pinhole.cpp:32-38, 96-100)
    float fx = 1, fy = 1, cx = 0, cy = 0;
    unsigned mnType = 0;                         // CAM_PINHOLE = 0, CAM_FISHEYE = 1 (include/CameraModels/GeometricCamera.h:93-96)
    unsigned GetType() const { return mnType; }
    Vec2 project(const Vec3& p) const { return Vec2{{fx * p(0) / p(2) + cx, fy * p(1) / p(2) + cy}}; }
    Mat3 toK_() const { Mat3 K; K(0, 0) = fx; K(0, 2) = cx; K(1, 1) = fy; K(1, 2) = cy; return K; }
    // Stand-in for a non-Pinhole GeometricCamera::epipolarConstrain (KannalaBrandt8's triangulates through an SVD): any pure
    // predicate of its arguments will do to test that the adapter calls the CALLER'S camera, on the reference's candidates and
    // in the reference's order.  It logs its calls: count and an order-sensitive hash of (kp1.x, kp2.y) bit patterns.
    long calls = 0; uint32_t callHash = 0;
    bool epipolarConstrain(Camera* /*other*/, const cv::KeyPoint& kp1, const cv::KeyPoint& kp2, const Mat3& /*R12*/, const Vec3& t12,
                           const float sigmaLevel, const float unc) {
        uint32_t a, b;
        std::memcpy(&a, &kp1.pt.x, 4); std::memcpy(&b, &kp2.pt.y, 4);
        callHash = callHash * 1000003u + (a ^ (b * 31u));
        ++calls;
        const float v = (kp1.pt.y - kp2.pt.y) + t12(1);
        const float w = kp1.pt.x - kp2.pt.x;
        return v * v < 9.0f * unc && w * w < 40000.0f * sigmaLevel;
    }
};

struct KeyFrame;
struct Frame;
struct Event { char kind; int a, b; };           // 'A' AddObservation(mp a, idx b), 'R' a->Replace(b), 'M' KF AddMapPoint(mp a, idx b)
static std::vector<Event> g_events;

struct MapPoint {
    int id = -1;
    Vec3 pos, normal;
    int level = 0;
    cv::Mat desc;
    bool bad = false;
    int nObs = 1;
    std::map<KeyFrame*, int> obs;
    bool mbTrackInView = false;
    float mTrackProjX = 0, mTrackProjY = 0, mTrackDepth = 1, mTrackViewCos = 1;
    int mnTrackScaleLevel = 0;
    bool isBad() const { return bad; }
    Vec3 GetWorldPos() const { return pos; }
    Vec3 GetNormal() const { return normal; }
    float GetMaxDistanceInvariance() const { return 1e9f; }
    float GetMinDistanceInvariance() const { return 0.f; }
    int PredictScale(float, KeyFrame*) const { return level; }
    int PredictScale(float, Frame*) const { return level; }
    cv::Mat GetDescriptor() const { return desc; }
    int Observations() const { return nObs; }
    bool IsInKeyFrame(KeyFrame* k) const { return obs.count(k) != 0; }
    std::tuple<int, int> GetIndexInKeyFrame(KeyFrame* k) const {
        auto it = obs.find(k);
        return it == obs.end() ? std::make_tuple(-1, -1) : std::make_tuple(it->second, -1);
    }
    void AddObservation(KeyFrame* k, int idx) { obs[k] = idx; nObs++; g_events.push_back(Event{'A', id, idx}); }
    void Replace(MapPoint* o);
};

struct FrameBase {
    int N = 0, Nleft = -1;
    std::vector<cv::KeyPoint> mvKeys, mvKeysUn;
    cv::Mat mDescriptors;
    float mnMinX = 0, mnMinY = 0, mnMaxX = 0, mnMaxY = 0;
    std::vector<float> mvScaleFactors, mvLevelSigma2, mvInvLevelSigma2;
    std::map<unsigned, std::vector<unsigned> > mFeatVec;     // DBoW2::FeatureVector
    Camera* mpCamera = nullptr;
    SE3 Tcw;
    SE3 GetPose() const { return Tcw; }
};
struct Frame : FrameBase {
    std::vector<MapPoint*> mvpMapPoints;
    std::vector<bool> mvbOutlier;
};
struct KeyFrame : FrameBase {
    float fx = 1, fy = 1, cx = 0, cy = 0;
    std::vector<MapPoint*> mps;
    std::vector<MapPoint*> GetMapPointMatches() const { return mps; }
    MapPoint* GetMapPoint(size_t i) const { return mps[i]; }
    std::set<MapPoint*> GetMapPoints() const { std::set<MapPoint*> s; for (MapPoint* p : mps) if (p && !p->isBad()) s.insert(p); return s; }
    void AddMapPoint(MapPoint* p, size_t idx) { mps[idx] = p; g_events.push_back(Event{'M', p->id, (int)idx}); }
    bool IsInImage(float x, float y) const { return x >= mnMinX && x < mnMaxX && y >= mnMinY && y < mnMaxY; }
    SE3 GetPoseInverse() const { return Tcw.inverse(); }
    Vec3 GetCameraCenter() const { return Tcw.inverse().translation(); }
};
// MapPoint::Replace (src/MapPoint.cc:228-279) reduced to what the Fuse loop can observe: this point goes bad, its
// observations move to the survivor (which therefore IsInKeyFrame of them) and the keyframe slots are re-pointed
inline void MapPoint::Replace(MapPoint* o) {
    g_events.push_back(Event{'R', id, o->id});
    bad = true;
    for (auto& kv : obs) {
        if (!o->IsInKeyFrame(kv.first)) { kv.first->mps[kv.second] = o; o->obs[kv.first] = kv.second; o->nObs++; }
        else kv.first->mps[kv.second] = nullptr;
    }
    obs.clear();
    nObs = 0;
}

}  // namespace shim
