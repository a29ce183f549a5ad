// TEST SCAFFOLDING: the OpenCV names the reference's tracker-side headers and TrackKLTGpu.h use, declared just far enough
// for a syntax check (see Eigen/Eigen in this directory).
#pragma once
#include <algorithm>
#include <cassert>
#include <map>
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <string>
#include <vector>
namespace cv {
struct Point2f { float x = 0, y = 0; Point2f() {} Point2f(float a, float b) : x(a), y(b) {} };
typedef Point2f Point_f;
struct Size { int width = 0, height = 0; Size() {} Size(int w, int h) : width(w), height(h) {} };
struct MatStep { size_t v = 0; operator size_t() const { return v; } };
#define CV_32F 5
#define CV_8UC1 0
struct Mat {
  unsigned char *data = nullptr;
  MatStep step;
  int cols = 0, rows = 0;
  Mat() {}
  Mat(int, int, int) {}
  bool empty() const { return data == nullptr; }
  Mat clone() const { return *this; }
  Mat reshape(int) const { return *this; }
  template <class T> T &at(int, int) { static T t; return t; }
};
struct KeyPoint { Point2f pt; float size = 0, angle = -1, response = 0; int octave = 0, class_id = -1; };
template <class T, int M, int N> struct Matx {
  T val[M * N];
  T &operator()(int i, int j) { return val[i * N + j]; }
  T &operator()(int i) { return val[i]; }
};
typedef Matx<double, 3, 3> Matx33d;
template <class T, int N> struct Vec { T val[N]; T &operator()(int i) { return val[i]; } };
typedef Vec<double, 4> Vec4d;
typedef Vec<float, 2> Vec2f;
struct Scalar { Scalar(double = 0, double = 0, double = 0, double = 0) {} };
typedef std::vector<Mat> InputArrayOfArrays;
inline void undistortPoints(const Mat &, Mat &, const Matx33d &, const Vec4d &) {}
namespace fisheye { inline void undistortPoints(const Mat &, Mat &, const Matx33d &, const Vec4d &) {} }
}  // namespace cv
