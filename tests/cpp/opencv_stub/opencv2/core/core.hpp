// TEST INFRASTRUCTURE: the slice of OpenCV's core API that include/gforb/orb_slam.hpp touches when it is compiled with
// -DGFORB_WITH_OPENCV (cv::Mat, cv::KeyPoint, cv::Point2f with OpenCV's member names, signatures and layouts).  OpenCV's C++ headers
// are not in this image; tests/test_abi.py compiles the shim's OpenCV branch against this stand-in so that the branch a maintainer
// of the reference would use is at least syntax- and type-checked against the OpenCV interface.  Not a reimplementation of OpenCV.
#pragma once
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <vector>

#define CV_8U 0
#define CV_32F 5
#define CV_8UC1 0
#define CV_ELEM_SIZE_STUB(type) ((type) == CV_32F ? 4 : 1)

namespace cv {
template <typename T> struct Point_ { T x, y; Point_() : x(0), y(0) {} Point_(T x_, T y_) : x(x_), y(y_) {} };
typedef Point_<float> Point2f;

struct KeyPoint {                 // opencv2/features2d/features2d.hpp: same field order, 28 bytes
    Point2f pt; float size, angle, response; int octave, class_id;
    KeyPoint() : size(0), angle(-1), response(0), octave(0), class_id(-1) {}
};

struct MatStep { size_t v; MatStep(size_t s = 0) : v(s) {} operator size_t() const { return v; } };

class Mat {
public:
    int rows, cols;
    unsigned char* data;
    MatStep step;
    Mat() : rows(0), cols(0), data(nullptr), type_(CV_8U) {}
    Mat(int r, int c, int type) : rows(0), cols(0), data(nullptr), type_(type) { create(r, c, type); }
    Mat(int r, int c, int type, void* ext, size_t stp = 0) : rows(r), cols(c), data((unsigned char*)ext), step(stp ? stp : (size_t)c * CV_ELEM_SIZE_STUB(type)), type_(type) {}
    Mat(const Mat& o) { *this = o; }
    Mat& operator=(const Mat& o) {
        rows = o.rows; cols = o.cols; step = o.step; type_ = o.type_; own_ = o.own_;
        data = (!o.own_.empty() && o.data == o.own_.data()) ? own_.data() : o.data;
        return *this;
    }
    void create(int r, int c, int type) {
        rows = r; cols = c; type_ = type; step = MatStep((size_t)c * CV_ELEM_SIZE_STUB(type));
        own_.assign((size_t)r * step, 0); data = own_.data();
    }
    bool empty() const { return rows == 0 || cols == 0 || data == nullptr; }
    int type() const { return type_; }
    unsigned char* ptr(int r = 0) { return data + (size_t)r * step; }
    const unsigned char* ptr(int r = 0) const { return data + (size_t)r * step; }
    Mat row(int r) const { return Mat(1, cols, type_, const_cast<unsigned char*>(ptr(r)), step); }
    template <class T> T& at(int r, int c) { return reinterpret_cast<T*>(ptr(r))[c]; }
    template <class T> const T& at(int r, int c) const { return reinterpret_cast<const T*>(ptr(r))[c]; }
private:
    int type_;
    std::vector<unsigned char> own_;
};
}  // namespace cv
