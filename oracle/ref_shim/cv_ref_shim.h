// ORACLE — TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product path.
//
// Minimal `cv::` surface that lets the reference's OWN, UNMODIFIED src/ORBextractor.cc (and include/ORBextractor.h)
// compile in this image, where OpenCV's C++ headers and libraries are not installed (only the cv2 Python wheel is).
// oracle/Makefile compiles /root/reference/src/ORBextractor.cc in place against this header into oracle/_ref/.
//
// What is the reference's and what is ours in the resulting library:
//   * reference source, byte for byte: the constructor tables and quotas (:409-469), ComputePyramid's buffer / ROI / border
//     choreography (:1170-1195), the FAST grid loop with its 3-vs-6 asymmetry and ini/min fallback (:781-872),
//     DistributeOctTree incl. std::list / std::sort(compareNodes) (:555-779), DivideNode (:480-533), IC_Angle (:76-103),
//     computeOrbDescriptor incl. its float expression (:106-146), operator()'s lapping / monoIndex assembly (:1086-1168);
//   * ours: the containers below (Mat views with OpenCV's ROI semantics) and the five pixel primitives, which forward to
//     oracle/cpp/orc_primitives.h — the restatement of OpenCV's 8-bit paths that tests/golden/*.npz and
//     tests/test_oracle_vs_cv2.py pin byte for byte to the real OpenCV (cv2 4.13).
// Semantics kept because the reference relies on them (OpenCV 4.x behaviour):
//   * Mat::operator()(Rect) / rowRange / colRange / row return views that share the parent's buffer and remember the
//     parent's size and their offset (locateROI);
//   * OutputArray::create() on a Mat that already has the requested size and type keeps its buffer (resize and
//     copyMakeBorder at :1183-1191 write INTO the ROI of `temp` that way);
//   * copyMakeBorder without BORDER_ISOLATED takes border pixels from the parent of a sub-matrix where they exist;
//   * cv::FAST reports KeyPoint(x, y, size 7, angle -1, response = score), row-major order;
//   * cvRound = round-half-even (lrint), cvFloor / cvCeil as OpenCV's inline versions.
#pragma once
#include <cassert>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <algorithm>
#include <memory>
#include <vector>
#include "../cpp/orc_primitives.h"

typedef unsigned char uchar;

#define CV_8U 0
#define CV_8UC1 0
#define CV_PI 3.1415926535897932384626433832795
#define CV_Assert(expr) assert(expr)

static inline int cvRound(double v) { return (int)lrint(v); }
static inline int cvRound(float v) { return (int)lrintf(v); }
static inline int cvRound(int v) { return v; }
static inline int cvFloor(double v) { int i = (int)v; return i - (i > v); }
static inline int cvFloor(float v) { int i = (int)v; return i - (i > v); }
static inline int cvCeil(double v) { int i = (int)v; return i + (i < v); }
static inline int cvCeil(float v) { int i = (int)v; return i + (i < v); }

namespace cv {

using ::cvRound; using ::cvFloor; using ::cvCeil;

template <class T> struct Point_ {
    T x, y;
    Point_() : x(0), y(0) {}
    Point_(T x_, T y_) : x(x_), y(y_) {}
    template <class U> Point_(const Point_<U>& p) : x((T)p.x), y((T)p.y) {}
    Point_& operator*=(float s) { x = (T)(x * s); y = (T)(y * s); return *this; }   // saturate_cast<float> is the identity
    Point_& operator*=(double s) { x = (T)(x * s); y = (T)(y * s); return *this; }
    Point_& operator+=(const Point_& o) { x += o.x; y += o.y; return *this; }
};
typedef Point_<int> Point2i;
typedef Point_<int> Point;
typedef Point_<float> Point2f;

template <class T> struct Size_ {
    T width, height;
    Size_() : width(0), height(0) {}
    Size_(T w, T h) : width(w), height(h) {}
};
typedef Size_<int> Size;

template <class T> struct Rect_ {
    T x, y, width, height;
    Rect_() : x(0), y(0), width(0), height(0) {}
    Rect_(T x_, T y_, T w_, T h_) : x(x_), y(y_), width(w_), height(h_) {}
};
typedef Rect_<int> Rect;

struct KeyPoint {            // field order and size of cv::KeyPoint (28 bytes)
    Point2f pt;
    float size, angle, response;
    int octave, class_id;
    KeyPoint() : pt(0, 0), size(0), angle(-1), response(0), octave(0), class_id(-1) {}
    KeyPoint(float x, float y, float size_, float angle_ = -1, float response_ = 0, int octave_ = 0, int class_id_ = -1)
        : pt(x, y), size(size_), angle(angle_), response(response_), octave(octave_), class_id(class_id_) {}
};

enum { BORDER_CONSTANT = 0, BORDER_REPLICATE = 1, BORDER_REFLECT = 2, BORDER_WRAP = 3, BORDER_REFLECT_101 = 4,
       BORDER_REFLECT101 = 4, BORDER_DEFAULT = 4, BORDER_ISOLATED = 16 };
enum { INTER_NEAREST = 0, INTER_LINEAR = 1, INTER_CUBIC = 2, INTER_AREA = 3 };

class _OutputArray;

// 8-bit single-channel matrix (the only type this file ever creates: CV_8UC1)
class Mat {
public:
    int rows, cols;
    size_t step;               // bytes per row (MatStep converts to size_t; the reference only casts it to int)
    uchar* data;

    Mat() : rows(0), cols(0), step(0), data(nullptr), wholeRows_(0), wholeCols_(0), ofsX_(0), ofsY_(0) {}
    Mat(int r, int c, int type) : Mat() { create(r, c, type); }
    Mat(Size sz, int type) : Mat() { create(sz.height, sz.width, type); }
    Mat(int r, int c, int /*type*/, void* d, size_t s = 0) : Mat() {        // user-allocated data, not owned
        rows = r; cols = c; step = s ? s : (size_t)c; data = (uchar*)d; wholeRows_ = r; wholeCols_ = c;
    }

    void create(int r, int c, int /*type*/) {
        if (data && rows == r && cols == c) return;                          // same size and type: keep the buffer
        buf_ = std::shared_ptr<uchar>(new uchar[std::max<size_t>((size_t)r * c, 1)], std::default_delete<uchar[]>());
        rows = r; cols = c; step = (size_t)c; data = buf_.get();
        wholeRows_ = r; wholeCols_ = c; ofsX_ = ofsY_ = 0;
    }
    void release() { buf_.reset(); rows = cols = 0; step = 0; data = nullptr; wholeRows_ = wholeCols_ = ofsX_ = ofsY_ = 0; }
    static Mat zeros(int r, int c, int type) { Mat m(r, c, type); for (int y = 0; y < r; ++y) memset(m.ptr(y), 0, (size_t)c); return m; }

    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
    int type() const { return CV_8UC1; }
    int channels() const { return 1; }
    size_t step1() const { return step; }
    size_t elemSize() const { return 1; }
    bool isContinuous() const { return step == (size_t)cols || rows == 1; }
    bool isSubmatrix() const { return rows != wholeRows_ || cols != wholeCols_; }
    void locateROI(Size& whole, Point& ofs) const { whole = Size(wholeCols_, wholeRows_); ofs = Point(ofsX_, ofsY_); }
    Size size() const { return Size(cols, rows); }

    uchar* ptr(int r = 0) { return data + (size_t)r * step; }
    const uchar* ptr(int r = 0) const { return data + (size_t)r * step; }
    template <class T> T* ptr(int r = 0) { return (T*)(data + (size_t)r * step); }
    template <class T> const T* ptr(int r = 0) const { return (const T*)(data + (size_t)r * step); }
    template <class T> T& at(int r, int c) { return ((T*)(data + (size_t)r * step))[c]; }
    template <class T> const T& at(int r, int c) const { return ((const T*)(data + (size_t)r * step))[c]; }

    Mat operator()(const Rect& r) const { return view(r.x, r.y, r.width, r.height); }
    Mat rowRange(int a, int b) const { return view(0, a, cols, b - a); }
    Mat colRange(int a, int b) const { return view(a, 0, b - a, rows); }
    Mat row(int r) const { return view(0, r, cols, 1); }
    Mat col(int c) const { return view(c, 0, 1, rows); }

    Mat clone() const {
        Mat m;
        if (empty()) return m;
        m.create(rows, cols, CV_8UC1);
        for (int y = 0; y < rows; ++y) memcpy(m.ptr(y), ptr(y), (size_t)cols);
        return m;
    }
    void copyTo(const _OutputArray& dst) const;

private:
    Mat view(int x, int y, int w, int h) const {
        assert(x >= 0 && y >= 0 && w >= 0 && h >= 0 && x + w <= cols && y + h <= rows);
        Mat m;
        m.rows = h; m.cols = w; m.step = step; m.data = data + (size_t)y * step + x; m.buf_ = buf_;
        m.wholeRows_ = wholeRows_; m.wholeCols_ = wholeCols_; m.ofsX_ = ofsX_ + x; m.ofsY_ = ofsY_ + y;
        return m;
    }
    std::shared_ptr<uchar> buf_;
    int wholeRows_, wholeCols_, ofsX_, ofsY_;
};

// proxies: the reference passes Mats (also temporaries) wherever OpenCV takes InputArray / OutputArray
class _InputArray {
public:
    _InputArray() {}
    _InputArray(const Mat& m) : m_(m) {}
    bool empty() const { return m_.empty(); }
    Mat getMat() const { return m_; }
protected:
    Mat m_;                    // header copy: shares the buffer
};
class _OutputArray {
public:
    _OutputArray(Mat& m) : p_(&m) {}
    _OutputArray(const Mat& m) : tmp_(m), p_(&tmp_) {}        // a temporary view (desc.row(i).copyTo(descriptors.row(k)))
    void create(int r, int c, int type) const { p_->create(r, c, type); }
    void create(Size sz, int type) const { p_->create(sz.height, sz.width, type); }
    void release() const { p_->release(); }
    bool empty() const { return p_->empty(); }
    Mat getMat() const { return *p_; }
    Mat& ref() const { return *p_; }
private:
    mutable Mat tmp_;
    Mat* p_;
};
typedef const _InputArray& InputArray;
typedef const _OutputArray& OutputArray;
typedef const _OutputArray& InputOutputArray;

inline void Mat::copyTo(const _OutputArray& dst) const {
    dst.create(rows, cols, CV_8UC1);
    Mat d = dst.getMat();
    for (int y = 0; y < rows; ++y) memmove(d.ptr(y), ptr(y), (size_t)cols);
}

// ---- the five pixel primitives: forwarded to the cv2-pinned restatements (oracle/cpp/orc_primitives.h) -----------------
inline void resize(InputArray _src, OutputArray _dst, Size dsize, double fx = 0, double fy = 0, int interpolation = INTER_LINEAR) {
    Mat src = _src.getMat();
    assert(interpolation == INTER_LINEAR && fx == 0 && fy == 0 && dsize.width > 0 && dsize.height > 0);
    (void)fx; (void)fy; (void)interpolation;
    _dst.create(dsize.height, dsize.width, CV_8UC1);
    Mat dst = _dst.getMat();
    orc::resize_linear_8u(src.data, src.cols, src.rows, (int)src.step, dst.data, dst.cols, dst.rows, (int)dst.step);
}

inline void copyMakeBorder(InputArray _src, OutputArray _dst, int top, int bottom, int left, int right, int borderType) {
    Mat src = _src.getMat();
    assert((borderType & ~BORDER_ISOLATED) == BORDER_REFLECT_101 && top == bottom && left == right && top == left);
    const uchar* sdata = src.data;
    int sw = src.cols, sh = src.rows;
    if (src.isSubmatrix() && !(borderType & BORDER_ISOLATED)) {
        // OpenCV: the part of the border that exists in the parent matrix is copied from there
        Size whole; Point ofs;
        src.locateROI(whole, ofs);
        const int dtop = std::min(ofs.y, top), dbottom = std::min(whole.height - sh - ofs.y, bottom);
        const int dleft = std::min(ofs.x, left), dright = std::min(whole.width - sw - ofs.x, right);
        // this shim supports the symmetric cases only: nothing available, or the whole ring available
        if (dtop == top && dbottom == bottom && dleft == left && dright == right) {
            _dst.create(sh + top + bottom, sw + left + right, CV_8UC1);
            Mat dst = _dst.getMat();
            const uchar* base = sdata - (size_t)top * src.step - left;
            for (int y = 0; y < dst.rows; ++y) memmove(dst.ptr(y), base + (size_t)y * src.step, (size_t)dst.cols);
            return;
        }
        assert(dtop == 0 && dbottom == 0 && dleft == 0 && dright == 0 && "partial parent border: not modelled by the shim");
    }
    _dst.create(sh + top + bottom, sw + left + right, CV_8UC1);
    Mat dst = _dst.getMat();
    orc::make_border_101(sdata, sw, sh, (int)src.step, dst.data, (int)dst.step, top);
}

inline void GaussianBlur(InputArray _src, OutputArray _dst, Size ksize, double sigmaX, double sigmaY = 0, int borderType = BORDER_DEFAULT) {
    Mat src = _src.getMat();
    assert(ksize.width == 7 && ksize.height == 7 && sigmaX == 2 && sigmaY == 2 && borderType == BORDER_REFLECT_101);
    (void)ksize; (void)sigmaX; (void)sigmaY; (void)borderType;
    // the bit-exact fixed-point path is the one OpenCV takes for a Mat that is not a sub-matrix (the reference clone()s
    // the level first, :1132); a sub-matrix input without BORDER_ISOLATED would read its parent's pixels instead
    assert(!src.isSubmatrix());
    _dst.create(src.rows, src.cols, CV_8UC1);
    Mat dst = _dst.getMat();
    orc::gaussian7_8u(src.data, src.cols, src.rows, (int)src.step, dst.data, (int)dst.step);
}

inline void FAST(InputArray _img, std::vector<KeyPoint>& keypoints, int threshold, bool nonmaxSuppression = true) {
    Mat img = _img.getMat();
    assert(nonmaxSuppression);
    (void)nonmaxSuppression;
    std::vector<orc::FastPt> pts;
    orc::fast9_nms(img.data, img.cols, img.rows, (int)img.step, threshold, pts);
    keypoints.clear();
    keypoints.reserve(pts.size());
    for (const orc::FastPt& p : pts) keypoints.push_back(KeyPoint((float)p.x, (float)p.y, 7.f, -1, (float)p.score));
}

// OpenCV library code, compiled apart from the reference with OpenCV's own flags: keep the cv2-pinned evaluation order
// (no FMA contraction) whatever flags the reference translation unit is built with (_ref/liborb_ref_fma.so)
__attribute__((noinline, optimize("fp-contract=off"))) inline float fastAtan2(float y, float x) { return orc::fast_atan2_deg(y, x); }

// Only ComputeKeyPointsOld (dead code: its call is commented out at src/ORBextractor.cc:1101) uses this; kept so that the
// file compiles unmodified.  Same contract as OpenCV: keep the n strongest, plus all that tie with the n-th.
struct KeyPointsFilter {
    static void retainBest(std::vector<KeyPoint>& kps, int n) {
        if (n >= 0 && kps.size() > (size_t)n) {
            if (n == 0) { kps.clear(); return; }
            std::nth_element(kps.begin(), kps.begin() + n - 1, kps.end(),
                             [](const KeyPoint& a, const KeyPoint& b) { return a.response > b.response; });
            const float amb = kps[n - 1].response;
            auto endIt = std::partition(kps.begin() + n, kps.end(), [amb](const KeyPoint& k) { return k.response >= amb; });
            kps.resize(endIt - kps.begin());
        }
    }
};

}  // namespace cv
