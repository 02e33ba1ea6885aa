/* opencv_min.hpp — the few OpenCV types the ORB front-end's class surface mentions, for builds WITHOUT OpenCV.
 *
 * The reference's boundary is its C++ class surface (include/ORBextractor.h, ORBmatcher.h, Frame.h), which
 * speaks cv::Mat / cv::KeyPoint / cv::InputArray.  When <opencv2/core.hpp> is available the shim headers use
 * the real types; this file is the stand-in used in the build container (no OpenCV C++ headers there) so the
 * shim and its tests compile.  cv::KeyPoint is layout-identical to OpenCV's (7 x 4 bytes); cv::Mat here is a
 * minimal owning 2-D byte/float matrix with the members the hot path touches (rows, cols, step, data, ptr,
 * create, empty, type, rowRange).  InputArray/OutputArray are plain references to that Mat. */
#ifndef ORBFE_OPENCV_MIN_HPP
#define ORBFE_OPENCV_MIN_HPP
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <memory>
#include <vector>

#define CV_8U 0
#define CV_8UC1 0
#define CV_32F 5

namespace cv {

struct Point2f { float x = 0, y = 0; Point2f() {} Point2f(float x_, float y_) : x(x_), y(y_) {} };
struct Point { int x = 0, y = 0; };

struct KeyPoint {
    Point2f pt;
    float size = 0, angle = -1, response = 0;
    int octave = 0, class_id = -1;
};
static_assert(sizeof(KeyPoint) == 28, "cv::KeyPoint layout");

class Mat {
public:
    int rows = 0, cols = 0;
    size_t step = 0;
    unsigned char* data = nullptr;
    Mat() {}
    Mat(int r, int c, int type) { create(r, c, type); }
    Mat(int r, int c, int type, void* ext, size_t stp = 0) : rows(r), cols(c), step(stp ? stp : (size_t)c * esz(type)), data((unsigned char*)ext), type_(type) {}
    void create(int r, int c, int type) {
        if (store_ && data == store_.get() && r == rows && c == cols && type == type_) return;   /* like cv::Mat::create: same shape keeps the buffer */
        rows = r; cols = c; type_ = type; step = (size_t)c * esz(type);
        store_ = std::shared_ptr<unsigned char>(new unsigned char[step * (size_t)(r > 0 ? r : 1)], std::default_delete<unsigned char[]>());
        data = store_.get();
    }
    void release() { store_.reset(); data = nullptr; rows = cols = 0; step = 0; }
    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
    int type() const { return type_; }
    template <typename T> T* ptr(int r = 0) { return (T*)(data + step * (size_t)r); }
    template <typename T> const T* ptr(int r = 0) const { return (const T*)(data + step * (size_t)r); }
    unsigned char* ptr(int r = 0) { return data + step * (size_t)r; }
    const unsigned char* ptr(int r = 0) const { return data + step * (size_t)r; }
    Mat row(int r) const { Mat m = *this; m.rows = 1; m.data = data + step * (size_t)r; return m; }
    Mat rowRange(int a, int b) const { Mat m = *this; m.rows = b - a; m.data = data + step * (size_t)a; return m; }
    Mat getMat() const { return *this; }
private:
    static size_t esz(int type) { return type == CV_32F ? 4 : 1; }
    int type_ = CV_8U;
    std::shared_ptr<unsigned char> store_;
};

typedef const Mat& InputArray;
typedef Mat& OutputArray;

}  // namespace cv
#endif
