// Minimal stand-in for the handful of OpenCV types the adapters touch, ONLY so that mam3slam_b200/cpp/*.h can be
// compiled and exercised in this container (no OpenCV C++ headers are installed).  A real build includes
// <opencv2/core.hpp> instead.  Layouts follow OpenCV: cv::KeyPoint is 28 bytes.
#pragma once
#include <cstdint>
#include <cstring>
#include <memory>
#include <vector>
#define CV_8U 0
#define CV_8UC1 0
namespace cv {
struct Point2f { float x = 0, y = 0; };
struct Rect { int x, y, width, height; Rect(int x_, int y_, int w_, int h_) : x(x_), y(y_), width(w_), height(h_) {} };
struct KeyPoint { Point2f pt; float size = 0, angle = -1, response = 0; int octave = 0, class_id = -1; };
class Mat {
public:
    int rows = 0, cols = 0; size_t step = 0; uint8_t* data = nullptr;
    Mat() {}
    Mat(int r, int c, int /*type*/) { create(r, c, 0); }
    Mat(int r, int c, int, void* d, size_t s) : rows(r), cols(c), step(s), data((uint8_t*)d) {}
    void create(int r, int c, int) { buf_ = std::shared_ptr<uint8_t>(new uint8_t[(size_t)r * c], std::default_delete<uint8_t[]>()); rows = r; cols = c; step = c; data = buf_.get(); }
    void release() { buf_.reset(); rows = cols = 0; step = 0; data = nullptr; }
    bool empty() const { return rows == 0 || cols == 0; }
    int type() const { return CV_8UC1; }
    uint8_t* ptr(int r) { return data + (size_t)r * step; }
    const uint8_t* ptr(int r) const { return data + (size_t)r * step; }
    template <class T> const T* ptr(int r = 0) const { return (const T*)(data + (size_t)r * step); }
    Mat row(int r) const { Mat m; m.rows = 1; m.cols = cols; m.step = step; m.data = data + (size_t)r * step; m.buf_ = buf_; return m; }
    Mat operator()(const Rect& r) const { Mat m; m.rows = r.height; m.cols = r.width; m.step = step; m.data = data + (size_t)r.y * step + r.x; m.buf_ = buf_; return m; }
    Mat getMat() const { return *this; }
private:
    std::shared_ptr<uint8_t> buf_;
};
struct InputArray_ { const Mat* m; InputArray_(const Mat& mm) : m(&mm) {} bool empty() const { return m->empty(); } Mat getMat() const { return *m; } };
struct OutputArray_ { Mat* m; OutputArray_(Mat& mm) : m(&mm) {} void create(int r, int c, int t) const { m->create(r, c, t); } void release() const { m->release(); } Mat getMat() const { return *m; } };
typedef const InputArray_& InputArray;
typedef const OutputArray_& OutputArray;
}  // namespace cv
