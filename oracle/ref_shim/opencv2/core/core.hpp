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
#include <climits>
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <vector>

#include "../../../cv_prims.h"

#define CV_PI 3.1415926535897932384626433832795
#define CV_8U 0
#define CV_8UC1 0
#define CV_32S 4
#define CV_32F 5
#define CV_32FC1 5
#define CV_64F 6

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

// Dense 2-D matrix header over a shared buffer. Depths: CV_8U (images, descriptors), CV_32S, CV_32F (poses,
// points, SAD windows), CV_64F; `type` also carries a channel count for reshape(), which only the
// undistortion path of R/src/Frame.cpp uses (never taken with zero distortion coefficients).
class Mat {
 public:
  int rows = 0, cols = 0;
  size_t step = 0;
  uchar* data = nullptr;

  Mat() {}
  Mat(int r, int c, int type) { create(r, c, type); }
  Mat(Size sz, int type) { create(sz.height, sz.width, type); }
  Mat(int r, int c, int type, void* ext, size_t extStep = 0) : rows(r), cols(c), data((uchar*)ext), type_(type) {
    step = extStep ? extStep : (size_t)c * elemSize();
  }

  static size_t depthSize(int type) {
    switch (type & 7) { case CV_8U: return 1; case CV_32S: return 4; case CV_32F: return 4; case CV_64F: return 8; default: return 1; }
  }
  size_t elemSize() const { return depthSize(type_) * (size_t)channels(); }
  int channels() const { return (type_ >> 3) + 1; }
  int depth() const { return type_ & 7; }
  int type() const { return type_; }

  void create(int r, int c, int type) {
    if (data && r == rows && c == cols && type == type_) return;  // cv::Mat::create keeps a matching allocation
    type_ = type;
    buf_ = std::make_shared<std::vector<uchar>>((size_t)r * c * elemSize());
    rows = r; cols = c; step = (size_t)c * elemSize(); data = buf_->data();
  }
  void release() { buf_.reset(); rows = cols = 0; step = 0; data = nullptr; }
  static MatZerosExpr zeros(int r, int c, int type) { return MatZerosExpr{r, c, type}; }
  Mat(const MatZerosExpr& z) { *this = z; }
  Mat& operator=(const MatZerosExpr& z) {
    create(z.rows, z.cols, z.type);
    for (int r = 0; r < rows; r++) std::memset(ptr(r), 0, (size_t)cols * elemSize());
    return *this;
  }
  static Mat ones(int r, int c, int type) {
    Mat m(r, c, type);
    assert(type == CV_32F);
    for (int i = 0; i < r; i++) for (int j = 0; j < c; j++) m.at<float>(i, j) = 1.0f;
    return m;
  }
  static Mat eye(int r, int c, int type) {
    Mat m = zeros(r, c, type);
    assert(type == CV_32F);
    for (int i = 0; i < std::min(r, c); i++) m.at<float>(i, i) = 1.0f;
    return m;
  }

  bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
  size_t step1() const { return step / depthSize(type_); }
  bool isContinuous() const { return step == (size_t)cols * elemSize(); }
  size_t total() const { return (size_t)rows * cols; }

  template <typename T> T& at(int r, int c) { return *(T*)(data + (size_t)r * step + (size_t)c * sizeof(T)); }
  template <typename T> const T& at(int r, int c) const { return *(const T*)(data + (size_t)r * step + (size_t)c * sizeof(T)); }
  // single index: element i of a row or column vector
  template <typename T> T& at(int i) { return rows == 1 ? at<T>(0, i) : at<T>(i, 0); }
  template <typename T> const T& at(int i) const { return rows == 1 ? at<T>(0, i) : at<T>(i, 0); }
  uchar* ptr(int r = 0) { return data + (size_t)r * step; }
  const uchar* ptr(int r = 0) const { return data + (size_t)r * step; }
  template <typename T> T* ptr(int r = 0) { return (T*)(data + (size_t)r * step); }
  template <typename T> const T* ptr(int r = 0) const { return (const T*)(data + (size_t)r * step); }

  Mat rowRange(int r0, int r1) const { Mat m(*this); m.data = data + (size_t)r0 * step; m.rows = r1 - r0; return m; }
  Mat colRange(int c0, int c1) const { Mat m(*this); m.data = data + (size_t)c0 * elemSize(); m.cols = c1 - c0; return m; }
  Mat row(int r) const { return rowRange(r, r + 1); }
  Mat col(int c) const { return colRange(c, c + 1); }
  Mat operator()(const Rect& r) const { return rowRange(r.y, r.y + r.height).colRange(r.x, r.x + r.width); }
  Mat clone() const {
    Mat m;
    copyTo(m);
    return m;
  }
  void copyTo(Mat& m) const {
    if (empty()) { m.release(); return; }
    if (m.data == data && m.rows == rows && m.cols == cols) return;
    m.create(rows, cols, type_);
    for (int r = 0; r < rows; r++) std::memcpy(m.ptr(r), ptr(r), (size_t)cols * elemSize());
  }
  // copyTo into a temporary header (Twc.rowRange(0,3).col(3) as OutputArray, R/src/KeyFrame.cpp:87-88): a view of
  // matching size and type is filled in place, like cv::Mat::create on a matching allocation
  void copyTo(Mat&& view) const {
    assert(view.rows == rows && view.cols == cols && view.type_ == type_ && view.data);
    for (int r = 0; r < rows; r++) std::memcpy(view.ptr(r), ptr(r), (size_t)cols * elemSize());
  }
  // depth conversion; only what R/src/Frame.cpp:676,698 needs (8-bit window -> float), in place allowed
  void convertTo(Mat& m, int type) const {
    assert(type == CV_32F && (depth() == CV_8U || depth() == CV_32F) && channels() == 1);
    Mat out(rows, cols, CV_32F);
    for (int r = 0; r < rows; r++)
      for (int c = 0; c < cols; c++) out.at<float>(r, c) = depth() == CV_8U ? (float)at<uchar>(r, c) : at<float>(r, c);
    m = out;
  }
  // same data seen with another channel count (continuous matrices only)
  Mat reshape(int cn) const {
    assert(isContinuous());
    Mat m(*this);
    const size_t rowBytes = (size_t)cols * elemSize();
    m.type_ = (type_ & 7) | ((cn - 1) << 3);
    m.cols = (int)(rowBytes / m.elemSize());
    return m;
  }
  Mat t() const {
    assert(type_ == CV_32F);
    Mat m(cols, rows, CV_32F);
    for (int r = 0; r < rows; r++) for (int c = 0; c < cols; c++) m.at<float>(c, r) = at<float>(r, c);
    return m;
  }
  // cv::Mat::dot accumulates float products in double
  double dot(const Mat& o) const {
    assert(type_ == CV_32F && o.type_ == CV_32F && total() == o.total());
    double acc = 0;
    if (rows == o.rows) { for (int r = 0; r < rows; r++) for (int c = 0; c < cols; c++) acc += (double)at<float>(r, c) * o.at<float>(r, c); }
    else { for (size_t i = 0; i < total(); i++) acc += (double)at<float>((int)i) * o.at<float>((int)i); }
    return acc;
  }

 private:
  int type_ = CV_8UC1;
  std::shared_ptr<std::vector<uchar>> buf_;
};

// ---- float matrix arithmetic (MatExpr in OpenCV; evaluated eagerly here). cv::gemm accumulates float
// products in double for these small sizes (GEMMSingleMul<float,double>) and rounds once per element. ----
inline Mat operator*(const Mat& a, const Mat& b) {
  assert(a.type() == CV_32F && b.type() == CV_32F && a.cols == b.rows);
  Mat m(a.rows, b.cols, CV_32F);
  for (int i = 0; i < a.rows; i++)
    for (int j = 0; j < b.cols; j++) {
      double acc = 0;
      for (int k = 0; k < a.cols; k++) acc += (double)a.at<float>(i, k) * b.at<float>(k, j);
      m.at<float>(i, j) = (float)acc;
    }
  return m;
}
template <typename F>
inline Mat mat_map2(const Mat& a, const Mat& b, F f) {
  assert(a.type() == CV_32F && b.type() == CV_32F && a.rows == b.rows && a.cols == b.cols);
  Mat m(a.rows, a.cols, CV_32F);
  for (int i = 0; i < a.rows; i++) for (int j = 0; j < a.cols; j++) m.at<float>(i, j) = f(a.at<float>(i, j), b.at<float>(i, j));
  return m;
}
template <typename F>
inline Mat mat_map1(const Mat& a, F f) {
  assert(a.type() == CV_32F);
  Mat m(a.rows, a.cols, CV_32F);
  for (int i = 0; i < a.rows; i++) for (int j = 0; j < a.cols; j++) m.at<float>(i, j) = f(a.at<float>(i, j));
  return m;
}
inline Mat operator+(const Mat& a, const Mat& b) { return mat_map2(a, b, [](float x, float y) { return x + y; }); }
inline Mat operator-(const Mat& a, const Mat& b) { return mat_map2(a, b, [](float x, float y) { return x - y; }); }
inline Mat operator-(const Mat& a) { return mat_map1(a, [](float x) { return -x; }); }
inline Mat operator*(const Mat& a, double s) { return mat_map1(a, [s](float x) { return (float)(x * s); }); }
inline Mat operator*(double s, const Mat& a) { return a * s; }
inline Mat operator/(const Mat& a, double s) { return mat_map1(a, [s](float x) { return (float)(x / s); }); }

enum { NORM_INF = 1, NORM_L1 = 2, NORM_L2 = 4 };
inline double norm(const Mat& a) {  // L2, double accumulation
  double acc = 0;
  for (int i = 0; i < a.rows; i++) for (int j = 0; j < a.cols; j++) { const double v = a.at<float>(i, j); acc += v * v; }
  return std::sqrt(acc);
}
inline double norm(const Mat& a, const Mat& b, int normType) {
  assert(normType == NORM_L1 && a.rows == b.rows && a.cols == b.cols);
  (void)normType;
  double acc = 0;
  for (int i = 0; i < a.rows; i++) for (int j = 0; j < a.cols; j++) acc += std::fabs((double)a.at<float>(i, j) - (double)b.at<float>(i, j));
  return acc;
}

// cv::Mat_<float>(r, c) << a, b, c;
template <typename T>
class Mat_ : public Mat {
 public:
  Mat_(int r, int c) : Mat(r, c, CV_32F) { static_assert(sizeof(T) == 4, "float only"); }
  struct Init {
    Mat_* m; int i;
    Init& operator,(T v) { m->template at<T>(i / m->cols, i % m->cols) = v; i++; return *this; }
    operator Mat() const { return *m; }
  };
  Init operator<<(T v) { Init in{this, 0}; in, v; return in; }
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

// R/src/Frame.cpp:491,523 only call this when the distortion coefficients are non-zero; the parity harness
// always passes zero distortion (the hot path works on the image as given), so reaching this is a harness bug.
inline void undistortPoints(InputArray, OutputArray, InputArray, InputArray, InputArray, InputArray) { std::abort(); }

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
