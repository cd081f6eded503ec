// TEST INFRASTRUCTURE (oracle) -- never linked into, imported by or executed from the product path.
//
// CPU restatements of the five OpenCV primitives the reference's ORB hot path calls. OpenCV is an
// external, un-vendored dependency of the reference (find_package(OpenCV 2.4.10), R/CMakeLists.txt:20)
// and its C++ library is not installed here; the only OpenCV on either box is the Python wheel
// opencv-python-headless 4.13.0. Each function restates the *published* OpenCV algorithm and is pinned
// bit-exactly against cv2 4.13.0 by tests/test_oracle_prims.py (live, when cv2 imports) and by the
// golden vectors under tests/golden/ (made by tools/make_golden.py with cv2 as the generator).
//
// Call sites in the reference these replace:
//   cv::resize(.., INTER_LINEAR)        R/src/ORBextractor.cpp:1212
//   cv::FAST(cell, kps, th, true)       R/src/ORBextractor.cpp:875,881
//   cv::GaussianBlur(7x7, s=2, R101)    R/src/ORBextractor.cpp:1167
//   cv::fastAtan2                       R/src/ORBextractor.cpp:106
//   cvRound                             R/src/ORBextractor.cpp:84,120,127-128,466,1203
#pragma once
#include <cstdint>
#include <vector>

namespace orb_oracle {

struct FastPoint { int x, y, score; };

// cvRound(float/double): round half to even (SSE cvtss2si under the default MXCSR).
int cv_round(double v);
int cv_round_f(float v);

// cv::resize, INTER_LINEAR, 8-bit single channel (fixed-point 11-bit coefficients).
void resize_linear_u8(const uint8_t* src, int sw, int sh, int sstride,
                      uint8_t* dst, int dw, int dh, int dstride);

// cv::GaussianBlur(src, dst, Size(7,7), 2, 2, BORDER_REFLECT_101) for CV_8UC1 (fixed-point path).
void gaussian_blur7_u8(const uint8_t* src, int w, int h, int sstride, uint8_t* dst, int dstride);

// cv::FAST(img, kps, threshold, nonmaxSuppression) with TYPE_9_16. Output in raster order.
void fast9_16(const uint8_t* img, int w, int h, int stride, int threshold, bool nms,
              std::vector<FastPoint>& out);

// cv::fastAtan2(y, x): degrees in [0, 360), float polynomial, no FMA contraction.
float fast_atan2(float y, float x);

// Restatement of glibc 2.39's sincosf (__sincosf_fma variant: every a+b*c contracted), used to
// check the device port; the extractor oracle itself calls the host libm like the reference does.
void sincosf_port(float y, float* sinp, float* cosp);

}  // namespace orb_oracle
