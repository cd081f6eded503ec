// TEST INFRASTRUCTURE (oracle/_ref) -- a minimal stand-in for the OpenCV C++ API surface that the
// reference's src/ORBextractor.cpp uses, so that file can be compiled WHERE IT LIES under
// /root/reference (OpenCV's C++ library is not installed on either box).
//
// Containers (Mat, KeyPoint, Point_, Size, Rect, Input/OutputArray) are written here from the
// OpenCV documentation; the five numeric primitives forward to oracle/cv_prims.cpp, the
// restatements pinned bit-exactly against cv2 4.13.0 (tests/test_oracle_prims.py, tests/golden/).
// Nothing in the product (orb-slam2-_b200/) includes this header.
#pragma once
#include <algorithm>
#include <cassert>
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <memory>
#include <vector>

#include "../../../cv_prims.h"

#define CV_PI 3.1415926535897932384626433832795
#define CV_8U 0
#define CV_8UC1 0

typedef unsigned char uchar;

inline int cvRound(double v) { return orb_oracle::cv_round(v); }
inline int cvRound(float v) { return orb_oracle::cv_round_f(v); }
inline int cvRound(int v) { return v; }
inline int cvFloor(double v) { int i = (int)v; return i - (i > v); }
inline int cvCeil(double v) { int i = (int)v; return i + (i < v); }

namespace cv {

using ::uchar;

enum { BORDER_CONSTANT = 0, BORDER_REPLICATE = 1, BORDER_REFLECT = 2, BORDER_WRAP = 3, BORDER_REFLECT_101 = 4,
       BORDER_DEFAULT = 4, BORDER_ISOLATED = 16 };
enum { INTER_NEAREST = 0, INTER_LINEAR = 1 };

template <typename T>
struct Point_ {
  T x, y;
  Point_() : x(0), y(0) {}
  Point_(T _x, T _y) : x(_x), y(_y) {}
  Point_& operator*=(float s) { x = (T)(x * s); y = (T)(y * s); return *this; }
};
typedef Point_<int> Point2i;
typedef Point2i Point;
typedef Point_<float> Point2f;

struct Size {
  int width, height;
  Size() : width(0), height(0) {}
  Size(int w, int h) : width(w), height(h) {}
};
struct Rect {
  int x, y, width, height;
  Rect() : x(0), y(0), width(0), height(0) {}
  Rect(int _x, int _y, int w, int h) : x(_x), y(_y), width(w), height(h) {}
};

struct KeyPoint {
  Point2f pt;
  float size, angle, response;
  int octave, class_id;
  KeyPoint() : pt(0, 0), size(0), angle(-1), response(0), octave(0), class_id(-1) {}
  KeyPoint(float x, float y, float _size, float _angle = -1, float _response = 0, int _octave = 0, int _class_id = -1)
      : pt(x, y), size(_size), angle(_angle), response(_response), octave(_octave), class_id(_class_id) {}
};

// cv::Mat::zeros returns a MatExpr; assigning it to a Mat of the same size and type fills that
// matrix IN PLACE (this is how the reference's computeDescriptors zeroes a row-range view of the
// output descriptors, R/src/ORBextractor.cpp:1113, without detaching it).
struct MatZerosExpr { int rows, cols, type; };

// 8-bit single-channel matrix header over a shared buffer (the only type the extractor uses).
class Mat {
 public:
  int rows = 0, cols = 0;
  size_t step = 0;
  uchar* data = nullptr;

  Mat() {}
  Mat(int r, int c, int type) { create(r, c, type); }
  Mat(Size sz, int type) { create(sz.height, sz.width, type); }
  Mat(int r, int c, int type, void* ext, size_t extStep) : rows(r), cols(c), step(extStep), data((uchar*)ext) { (void)type; }

  void create(int r, int c, int type) {
    assert(type == CV_8UC1);
    (void)type;
    if (data && r == rows && c == cols) return;  // cv::Mat::create keeps a matching allocation
    buf_ = std::make_shared<std::vector<uchar>>((size_t)r * c);
    rows = r; cols = c; step = (size_t)c; data = buf_->data();
  }
  void release() { buf_.reset(); rows = cols = 0; step = 0; data = nullptr; }
  static MatZerosExpr zeros(int r, int c, int type) { return MatZerosExpr{r, c, type}; }
  Mat(const MatZerosExpr& z) { *this = z; }
  Mat& operator=(const MatZerosExpr& z) {
    create(z.rows, z.cols, z.type);
    for (int r = 0; r < rows; r++) std::memset(ptr(r), 0, (size_t)cols);
    return *this;
  }

  int type() const { return CV_8UC1; }
  bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
  size_t step1() const { return step; }
  bool isContinuous() const { return step == (size_t)cols; }

  template <typename T> T& at(int r, int c) { return *(T*)(data + (size_t)r * step + (size_t)c * sizeof(T)); }
  template <typename T> const T& at(int r, int c) const { return *(const T*)(data + (size_t)r * step + (size_t)c * sizeof(T)); }
  uchar* ptr(int r = 0) { return data + (size_t)r * step; }
  const uchar* ptr(int r = 0) const { return data + (size_t)r * step; }

  Mat rowRange(int r0, int r1) const { Mat m(*this); m.data = data + (size_t)r0 * step; m.rows = r1 - r0; return m; }
  Mat colRange(int c0, int c1) const { Mat m(*this); m.data = data + c0; m.cols = c1 - c0; return m; }
  Mat operator()(const Rect& r) const { return rowRange(r.y, r.y + r.height).colRange(r.x, r.x + r.width); }
  Mat clone() const {
    Mat m;
    if (empty()) return m;
    m.create(rows, cols, CV_8UC1);
    for (int r = 0; r < rows; r++) std::memcpy(m.ptr(r), ptr(r), (size_t)cols);
    return m;
  }

 private:
  std::shared_ptr<std::vector<uchar>> buf_;
};

class _InputArray {
 public:
  _InputArray(const Mat& m) : m_(const_cast<Mat*>(&m)) {}
  Mat getMat() const { return *m_; }
  bool empty() const { return m_->empty(); }
 protected:
  Mat* m_;
};
class _OutputArray : public _InputArray {
 public:
  _OutputArray(Mat& m) : _InputArray(m) {}
  void create(int r, int c, int type) const { m_->create(r, c, type); }
  void create(Size sz, int type) const { m_->create(sz.height, sz.width, type); }
  void release() const { m_->release(); }
};
typedef const _InputArray& InputArray;
typedef const _OutputArray& OutputArray;

inline float fastAtan2(float y, float x) { return orb_oracle::fast_atan2(y, x); }

inline void FAST(InputArray image, std::vector<KeyPoint>& keypoints, int threshold, bool nonmaxSuppression = true) {
  const Mat img = image.getMat();
  std::vector<orb_oracle::FastPoint> pts;
  orb_oracle::fast9_16(img.data, img.cols, img.rows, (int)img.step, threshold, nonmaxSuppression, pts);
  keypoints.clear();
  for (const auto& p : pts) keypoints.push_back(KeyPoint((float)p.x, (float)p.y, 7.f, -1, (float)p.score));
}

inline void GaussianBlur(InputArray src, OutputArray dst, Size ksize, double sigmaX, double sigmaY = 0,
                         int borderType = BORDER_DEFAULT) {
  assert(ksize.width == 7 && ksize.height == 7 && sigmaX == 2 && sigmaY == 2 && borderType == BORDER_REFLECT_101);
  (void)ksize; (void)sigmaX; (void)sigmaY; (void)borderType;
  const Mat s = src.getMat().clone();  // in-place calls are allowed
  dst.create(s.rows, s.cols, CV_8UC1);
  Mat d = dst.getMat();
  orb_oracle::gaussian_blur7_u8(s.data, s.cols, s.rows, (int)s.step, d.data, (int)d.step);
}

inline void resize(InputArray src, OutputArray dst, Size dsize, double fx = 0, double fy = 0,
                   int interpolation = INTER_LINEAR) {
  assert(fx == 0 && fy == 0 && interpolation == INTER_LINEAR);
  (void)fx; (void)fy; (void)interpolation;
  const Mat s = src.getMat();
  dst.create(dsize, CV_8UC1);
  Mat d = dst.getMat();
  orb_oracle::resize_linear_u8(s.data, s.cols, s.rows, (int)s.step, d.data, d.cols, d.rows, (int)d.step);
}

inline int borderInterpolate101(int p, int len) {
  if (len == 1) return 0;
  while (p < 0 || p >= len) p = p < 0 ? -p : 2 * (len - 1) - p;
  return p;
}

// BORDER_REFLECT_101 only (with or without BORDER_ISOLATED: the shim never looks outside a ROI).
inline void copyMakeBorder(InputArray src, OutputArray dst, int top, int bottom, int left, int right, int borderType) {
  assert((borderType & ~BORDER_ISOLATED) == BORDER_REFLECT_101);
  (void)borderType;
  const Mat s = src.getMat().clone();
  dst.create(s.rows + top + bottom, s.cols + left + right, CV_8UC1);
  Mat d = dst.getMat();
  for (int r = 0; r < d.rows; r++) {
    const uchar* sr = s.ptr(borderInterpolate101(r - top, s.rows));
    uchar* dr = d.ptr(r);
    for (int c = 0; c < d.cols; c++) dr[c] = sr[borderInterpolate101(c - left, s.cols)];
  }
}

struct KeyPointsFilter {
  // keep the n_points strongest responses (only reached from the reference's dead ComputeKeyPointsOld)
  static void retainBest(std::vector<KeyPoint>& kps, int n_points) {
    if (n_points >= 0 && (int)kps.size() > n_points) {
      std::stable_sort(kps.begin(), kps.end(), [](const KeyPoint& a, const KeyPoint& b) { return a.response > b.response; });
      kps.resize(n_points);
    }
  }
};

}  // namespace cv
