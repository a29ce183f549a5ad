// Undistortion to normalised coordinates and fundamental-matrix outlier rejection.
//
//   CamBase::undistort_cv -> cv::undistortPoints / cv::fisheye::undistortPoints   cam/CamRadtan.h:102, cam/CamEqui.h:121
//   cv::findFundamentalMat(pts0_n, pts1_n, FM_RANSAC, 2.0/max_focal, 0.999, mask) track/TrackKLT.cpp:873
//
// SURVEY.md App. A.7 / A.8.  One thread block per (rig, camera) job.  The sample stream is OpenCV's: a multiply-with-
// carry generator seeded with 2^64-1 on every call, seven distinct indices per sample, collinearity-rejected, so the
// hypotheses tested are the ones the CPU tracker tests, in the same order.  Samples are drawn a chunk at a time by one
// lane, solved one per lane, scored lane-parallel over all points, and a serial scan applies the adaptive iteration
// cut-off in order.  The 7-point solver follows cv::run7Point as of OpenCV 4.13: Hartley-normalised points, the two
// null-space vectors OpenCV's Jacobi SVD produces (see seven_point), cv::solveCubic.  The solutions do not depend on
// those choices but their ORDER does, and RANSAC keeps the first of two equally good models (tests/golden/
// ransac_ties.npz).  N < 15 points switches to least-median (as OpenCV does silently).
#pragma once
#include <float.h>

#include "image_body.h"
#include "simt.h"

namespace uvio {

// ---- undistortion (fp64 inside, fp32 in/out) -------------------------------------------------------------------------
struct CamModel {
  int model;        // 0 radtan, 1 equidistant
  double k[4];      // fx fy cx cy
  double d[4];      // k1 k2 p1 p2  |  k1 k2 k3 k4
};

UVIO_HD void undistort_point(const CamModel &c, float u, float v, float &xo, float &yo) {
  if (c.model == 0) {
    double x = ((double)u - c.k[2]) * (1. / c.k[0]), y = ((double)v - c.k[3]) * (1. / c.k[1]);
    double x0 = x, y0 = y;
    for (int j = 0; j < 5; j++) {
      double r2 = x * x + y * y;
      double icdist = 1. / (1 + (c.d[1] * r2 + c.d[0]) * r2);
      double dX = 2 * c.d[2] * x * y + c.d[3] * (r2 + 2 * x * x);
      double dY = c.d[2] * (r2 + 2 * y * y) + 2 * c.d[3] * x * y;
      x = (x0 - dX) * icdist;
      y = (y0 - dY) * icdist;
    }
    xo = (float)x;
    yo = (float)y;
  } else {
    const double PI_2 = 1.57079632679489661923;
    double pwx = ((double)u - c.k[2]) / c.k[0], pwy = ((double)v - c.k[3]) / c.k[1];
    double theta_d = sqrt(pwx * pwx + pwy * pwy);
    if (theta_d > PI_2) theta_d = PI_2;
    if (!(theta_d > 1e-8)) {
      xo = (float)pwx;
      yo = (float)pwy;
      return;
    }
    double theta = theta_d;
    bool converged = false;
    for (int j = 0; j < 10; j++) {
      double t2 = theta * theta, t4 = t2 * t2, t6 = t4 * t2, t8 = t6 * t2;
      double k0 = c.d[0] * t2, k1 = c.d[1] * t4, k2 = c.d[2] * t6, k3 = c.d[3] * t8;
      double fix = (theta * (1 + k0 + k1 + k2 + k3) - theta_d) / (1 + 3 * k0 + 5 * k1 + 7 * k2 + 9 * k3);
      theta = theta - fix;
      if (fabs(fix) < 1e-8) {
        converged = true;
        break;
      }
    }
    double scale = tan(theta) / theta_d;
    if (converged && !(theta < 0)) {
      xo = (float)(pwx * scale);
      yo = (float)(pwy * scale);
    } else {
      xo = -1000000.0f;
      yo = -1000000.0f;
    }
  }
}

// ---- RNG (cv::RNG: multiply-with-carry) ------------------------------------------------------------------------------
UVIO_HD unsigned rng_next(uint64_t &state) {
  state = (uint64_t)(unsigned)state * 4164903690ULL + (unsigned)(state >> 32);
  return (unsigned)state;
}

constexpr int RS_H = 32;        // most samples per chunk
#ifdef RS_H_FIRST_OVERRIDE
constexpr int RS_H_FIRST = RS_H_FIRST_OVERRIDE;
#else
constexpr int RS_H_FIRST = 8;   // samples in the first chunk
#endif
constexpr int RS_MAXROOT = 3;

struct RansacSmem {
  float *x0, *y0, *x1, *y1;                 // N normalised points of each image (dynamic shared memory on the device)
  int idx[RS_H][7];
  uint64_t state_after[RS_H];               // generator state after sample i was drawn
  int ok[RS_H];                             // 1 = passed the collinearity check
  int nmodels[RS_H];
  double F[RS_H][RS_MAXROOT][9];
  unsigned counts[RS_H * RS_MAXROOT];       // RANSAC: inlier counts
  uint8_t slots[RS_H * RS_MAXROOT];         // RANSAC: the (sample, root) slots that hold a model
  int nslots;
  float median[RS_H * RS_MAXROOT];          // LMedS: median error
  double bestF[9];
  // serial state (written by lane 0 only)
  uint64_t state;
  int iter, niters, max_good, done, have_best, attempts;
  double min_median;
};

UVIO_HD bool collinear_last(const float *xs, const float *ys, const int *idx) {
  const int i = 6;
  for (int j = 0; j < i; j++) {
    double dx1 = (double)fsub(xs[idx[j]], xs[idx[i]]), dy1 = (double)fsub(ys[idx[j]], ys[idx[i]]);
    for (int k = 0; k < j; k++) {
      double dx2 = (double)fsub(xs[idx[k]], xs[idx[i]]), dy2 = (double)fsub(ys[idx[k]], ys[idx[i]]);
      if (fabs(dx2 * dy1 - dy2 * dx1) <= (double)FLT_EPSILON * (fabs(dx1) + fabs(dy1) + fabs(dx2) + fabs(dy2))) return true;
    }
  }
  return false;
}

// cv::solveCubic for c0 x^3 + c1 x^2 + c2 x + c3 = 0; returns the root count, roots in OpenCV's order.
UVIO_HD int solve_cubic(const double *c, double *r) {
  const double PI = 3.14159265358979323846;
  double a0 = c[0], a1 = c[1], a2 = c[2], a3 = c[3];
  double x0 = 0, x1 = 0, x2 = 0;
  int n = 0;
  if (a0 == 0) {
    if (a1 == 0) {
      if (a2 == 0) n = a3 == 0 ? -1 : 0;
      else { x0 = -a3 / a2; n = 1; }
    } else {
      double d = a2 * a2 - 4 * a1 * a3;
      if (d >= 0) {
        d = sqrt(d);
        double q1 = (-a2 + d) * 0.5, q2 = (a2 + d) * -0.5;
        if (fabs(q1) > fabs(q2)) { x0 = q1 / a1; x1 = a3 / q1; }
        else { x0 = q2 / a1; x1 = a3 / q2; }
        n = d > 0 ? 2 : 1;
      }
    }
  } else {
    a0 = 1. / a0;
    a1 *= a0; a2 *= a0; a3 *= a0;
    double Q = (a1 * a1 - 3 * a2) * (1. / 9);
    double R = (2 * a1 * a1 * a1 - 9 * a1 * a2 + 27 * a3) * (1. / 54);
    double Qcubed = Q * Q * Q;
    double d = Qcubed - R * R;
    if (d > 0) {
      double theta = acos(R / sqrt(Qcubed));
      double sqrtQ = sqrt(Q);
      double t0 = -2 * sqrtQ, t1 = theta * (1. / 3), t2 = a1 * (1. / 3);
      x0 = t0 * cos(t1) - t2;
      x1 = t0 * cos(t1 + (2. * PI / 3)) - t2;
      x2 = t0 * cos(t1 + (4. * PI / 3)) - t2;
      n = 3;
    } else if (d == 0) {
      if (R >= 0) { x0 = -2 * pow(R, 1. / 3) - a1 / 3; x1 = pow(R, 1. / 3) - a1 / 3; }
      else { x0 = 2 * pow(-R, 1. / 3) - a1 / 3; x1 = -pow(-R, 1. / 3) - a1 / 3; }
      x2 = 0;
      n = x0 == x1 ? 1 : 2;
      x1 = x0 == x1 ? 0 : x1;
    } else {
      d = sqrt(-d);
      double e = pow(d + fabs(R), 1. / 3);
      if (R > 0) e = -e;
      x0 = (e + Q / e) - a1 * (1. / 3);
      n = 1;
    }
  }
  r[0] = x0; r[1] = x1; r[2] = x2;
  return n;
}

// 7-point algorithm on sample points (x0,y0) <-> (x1,y1). Writes up to 3 matrices (row-major 3x3); returns the count.
UVIO_HD int seven_point(const double *x0r, const double *y0r, const double *x1r, const double *y1r, double (*Fout)[9]) {
  // Hartley normalisation of the seven points, as cv::run7Point does (centre at the origin, mean distance sqrt(2)).  It does
  // not change the solutions, but the cubic is solved in the normalised system and its root order is the order of the models.
  double c0x = 0, c0y = 0, c1x = 0, c1y = 0;
  for (int i = 0; i < 7; i++) { c0x += x0r[i]; c0y += y0r[i]; c1x += x1r[i]; c1y += y1r[i]; }
  const double t7 = 1. / 7;
  c0x *= t7; c0y *= t7; c1x *= t7; c1y *= t7;
  double scale0 = 0, scale1 = 0;
  for (int i = 0; i < 7; i++) {
    scale0 += sqrt((x0r[i] - c0x) * (x0r[i] - c0x) + (y0r[i] - c0y) * (y0r[i] - c0y));
    scale1 += sqrt((x1r[i] - c1x) * (x1r[i] - c1x) + (y1r[i] - c1y) * (y1r[i] - c1y));
  }
  scale0 *= t7;
  scale1 *= t7;
  if (scale0 < FLT_EPSILON || scale1 < FLT_EPSILON) return 0;
  scale0 = sqrt(2.) / scale0;
  scale1 = sqrt(2.) / scale1;
  double x0[7], y0[7], x1[7], y1[7];
  for (int i = 0; i < 7; i++) {
    x0[i] = (x0r[i] - c0x) * scale0; y0[i] = (y0r[i] - c0y) * scale0;
    x1[i] = (x1r[i] - c1x) * scale1; y1[i] = (y1r[i] - c1y) * scale1;
  }
  double A[7][9];
  int perm[9];
  for (int j = 0; j < 9; j++) perm[j] = j;
  for (int i = 0; i < 7; i++) {
    A[i][0] = x1[i] * x0[i]; A[i][1] = x1[i] * y0[i]; A[i][2] = x1[i];
    A[i][3] = y1[i] * x0[i]; A[i][4] = y1[i] * y0[i]; A[i][5] = y1[i];
    A[i][6] = x0[i];         A[i][7] = y0[i];         A[i][8] = 1.0;
  }
  // Gauss-Jordan with complete pivoting -> [I | B] in permuted column order
  for (int r = 0; r < 7; r++) {
    int pi = r, pj = r;
    double best = -1;
    for (int i = r; i < 7; i++)
      for (int j = r; j < 9; j++) {
        double v = fabs(A[i][j]);
        if (v > best) { best = v; pi = i; pj = j; }
      }
    if (!(best > 0)) return 0;
    if (pi != r) for (int j = 0; j < 9; j++) { double t = A[r][j]; A[r][j] = A[pi][j]; A[pi][j] = t; }
    if (pj != r) {
      for (int i = 0; i < 7; i++) { double t = A[i][r]; A[i][r] = A[i][pj]; A[i][pj] = t; }
      int t = perm[r]; perm[r] = perm[pj]; perm[pj] = t;
    }
    double inv = 1.0 / A[r][r];
    for (int j = r; j < 9; j++) A[r][j] *= inv;
    for (int i = 0; i < 7; i++) {
      if (i == r) continue;
      double f = A[i][r];
      if (f != 0) for (int j = r; j < 9; j++) A[i][j] -= f * A[r][j];
    }
  }
  double f1[9], f2[9];
  for (int r = 0; r < 7; r++) { f1[perm[r]] = -A[r][7]; f2[perm[r]] = -A[r][8]; }
  f1[perm[7]] = 1; f1[perm[8]] = 0;
  f2[perm[7]] = 0; f2[perm[8]] = 1;
  double n1 = 0, dot = 0, n2 = 0;
  for (int i = 0; i < 9; i++) n1 += f1[i] * f1[i];
  n1 = 1.0 / sqrt(n1);
  for (int i = 0; i < 9; i++) { f1[i] *= n1; dot += f1[i] * f2[i]; }
  for (int i = 0; i < 9; i++) { f2[i] -= dot * f1[i]; n2 += f2[i] * f2[i]; }
  n2 = 1.0 / sqrt(n2);
  for (int i = 0; i < 9; i++) f2[i] *= n2;
  // (f1, f2) is now an orthonormal basis of the null space.  Which basis is used does not change the three solutions, but
  // it decides the ORDER solveCubic returns them in, and RANSAC keeps the first of two equally good models.  OpenCV takes
  // rows 7 and 8 of V from SVD::compute(FULL_UV); for a 7x9 system that is its own Jacobi SVD (below the LAPACK size
  // threshold), which completes V by projecting two fixed pseudo-random sign vectors (cv::RNG(0x12345678), +-1/9) off the
  // rows already found and normalising (JacobiSVDImpl_, modules/core/src/lapack.cpp).  Those two vectors are therefore the
  // projections of the fixed sign vectors onto the null space, whatever basis it was found in:
  {
    const double sg1[9] = {-1, -1, 1, -1, -1, -1, -1, 1, 1}, sg2[9] = {1, -1, 1, 1, 1, 1, 1, -1, 1};
    double a1 = 0, a2 = 0, b1 = 0, b2 = 0;
    for (int i = 0; i < 9; i++) {
      a1 += sg1[i] * f1[i]; a2 += sg1[i] * f2[i];
      b1 += sg2[i] * f1[i]; b2 += sg2[i] * f2[i];
    }
    double g1[9], g2[9], m1 = 0, m2 = 0, d = 0;
    for (int i = 0; i < 9; i++) { g1[i] = a1 * f1[i] + a2 * f2[i]; m1 += g1[i] * g1[i]; }
    if (m1 > 1e-24) {   // (OpenCV would draw another vector if the projection vanished; that needs a 1e-12 coincidence)
      m1 = 1.0 / sqrt(m1);
      for (int i = 0; i < 9; i++) { g1[i] *= m1; g2[i] = b1 * f1[i] + b2 * f2[i]; }
      for (int i = 0; i < 9; i++) d += g1[i] * g2[i];
      for (int i = 0; i < 9; i++) { g2[i] -= d * g1[i]; m2 += g2[i] * g2[i]; }
      if (m2 > 1e-24) {
        m2 = 1.0 / sqrt(m2);
        for (int i = 0; i < 9; i++) { f1[i] = g1[i]; f2[i] = g2[i] * m2; }
      }
    }
  }
  // F = lambda*f1 + (1-lambda)*f2 ; det F = 0 is a cubic in lambda
  for (int i = 0; i < 9; i++) f1[i] -= f2[i];
  double c[4], t0, t1, t2;
  t0 = f2[4] * f2[8] - f2[5] * f2[7];
  t1 = f2[3] * f2[8] - f2[5] * f2[6];
  t2 = f2[3] * f2[7] - f2[4] * f2[6];
  c[3] = f2[0] * t0 - f2[1] * t1 + f2[2] * t2;
  c[2] = f1[0] * t0 - f1[1] * t1 + f1[2] * t2 - f1[3] * (f2[1] * f2[8] - f2[2] * f2[7]) +
         f1[4] * (f2[0] * f2[8] - f2[2] * f2[6]) - f1[5] * (f2[0] * f2[7] - f2[1] * f2[6]) +
         f1[6] * (f2[1] * f2[5] - f2[2] * f2[4]) - f1[7] * (f2[0] * f2[5] - f2[2] * f2[3]) +
         f1[8] * (f2[0] * f2[4] - f2[1] * f2[3]);
  t0 = f1[4] * f1[8] - f1[5] * f1[7];
  t1 = f1[3] * f1[8] - f1[5] * f1[6];
  t2 = f1[3] * f1[7] - f1[4] * f1[6];
  c[0] = f1[0] * t0 - f1[1] * t1 + f1[2] * t2;
  c[1] = f2[0] * t0 - f2[1] * t1 + f2[2] * t2 - f2[3] * (f1[1] * f1[8] - f1[2] * f1[7]) +
         f2[4] * (f1[0] * f1[8] - f1[2] * f1[6]) - f2[5] * (f1[0] * f1[7] - f1[1] * f1[6]) +
         f2[6] * (f1[1] * f1[5] - f1[2] * f1[4]) - f2[7] * (f1[0] * f1[5] - f1[2] * f1[3]) +
         f2[8] * (f1[0] * f1[4] - f1[1] * f1[3]);
  double roots[3];
  int n = solve_cubic(c, roots);
  if (n < 1 || n > 3) return 0;
  for (int k = 0; k < n; k++) {
    double lambda = roots[k], mu = 1.;
    double s = f1[8] * roots[k] + f2[8];
    double *F = Fout[k];
    if (fabs(s) > DBL_EPSILON) {
      mu = 1. / s;
      lambda *= mu;
      F[8] = 1.;
    } else {
      F[8] = 0.;
    }
    for (int i = 0; i < 8; i++) F[i] = f1[i] * lambda + f2[i] * mu;
    // back to the caller's coordinates: F <- T1^T F T0 with T = [s 0 -s*cx; 0 s -s*cy; 0 0 1], then F(3,3) = 1
    double G[9];
    for (int r = 0; r < 3; r++) {   // G = F * T0
      G[3 * r + 0] = F[3 * r + 0] * scale0;
      G[3 * r + 1] = F[3 * r + 1] * scale0;
      G[3 * r + 2] = F[3 * r + 2] - (F[3 * r + 0] * c0x + F[3 * r + 1] * c0y) * scale0;
    }
    for (int c = 0; c < 3; c++) {   // F = T1^T * G
      F[0 + c] = G[0 + c] * scale1;
      F[3 + c] = G[3 + c] * scale1;
      F[6 + c] = G[6 + c] - (G[0 + c] * c1x + G[3 + c] * c1y) * scale1;
    }
    if (fabs(F[8]) > FLT_EPSILON) {
      double inv = 1. / F[8];
      for (int i = 0; i < 9; i++) F[i] *= inv;
    }
  }
  return n;
}

// symmetric epipolar distance^2 (FMEstimatorCallback::computeError), stored as fp32
UVIO_HD float fm_error(const double *F, float fx0, float fy0, float fx1, float fy1) {
  double x0 = fx0, y0 = fy0, x1 = fx1, y1 = fy1;
  double a = F[0] * x0 + F[1] * y0 + F[2], b = F[3] * x0 + F[4] * y0 + F[5], c = F[6] * x0 + F[7] * y0 + F[8];
  double s2 = 1. / (a * a + b * b);
  double d2 = x1 * a + y1 * b + c;
  a = F[0] * x1 + F[3] * y1 + F[6];
  b = F[1] * x1 + F[4] * y1 + F[7];
  c = F[2] * x1 + F[5] * y1 + F[8];
  double s1 = 1. / (a * a + b * b);
  double d1 = x0 * a + y0 * b + c;
  double e1 = d1 * d1 * s1, e2 = d2 * d2 * s2;
  return (float)(e1 > e2 ? e1 : e2);
}

// cv::RANSACUpdateNumIters
UVIO_HD int ransac_update_iters(double p, double ep, int model_points, int max_iters) {
  p = p < 0 ? 0 : (p > 1 ? 1 : p);
  ep = ep < 0 ? 0 : (ep > 1 ? 1 : ep);
  double num = 1. - p;
  if (num < DBL_MIN) num = DBL_MIN;
  double denom = 1. - pow(1. - ep, (double)model_points);
  if (denom < DBL_MIN) return 0;
  num = log(num);
  denom = log(denom);
  return denom >= 0 || -num >= max_iters * (-denom) ? max_iters : round_d2i(num / denom);
}

// Draw `chunk` (<= RS_H) samples (lane 0; integer work only), then check collinearity one sample per lane.
template <class G>
UVIO_HD void ransac_draw_chunk(G g, RansacSmem &s, int N, int chunk) {
  if (g.lane() == 0) {
    uint64_t st = s.state;
    for (int h = 0; h < chunk; h++) {
      int *idx = s.idx[h];
      for (int i = 0; i < 7; ++i) {
        int v;
        bool dup;
        do {
          v = (int)(rng_next(st) % (unsigned)N);
          dup = false;
          for (int k = 0; k < i; k++) dup |= (idx[k] == v);
        } while (dup);
        idx[i] = v;
      }
      s.state_after[h] = st;
    }
  }
  g.sync();
  for (int h = g.lane(); h < chunk; h += g.size())
    s.ok[h] = !collinear_last(s.x0, s.y0, s.idx[h]) && !collinear_last(s.x1, s.y1, s.idx[h]);
  g.sync();
}

// One (rig, camera) job. p0/p1: N normalised points (x,y interleaved). thr: inlier threshold (RANSAC). mask_out: N bytes.
// Returns (in every lane) 1 if a model was found.  Callers with N < 7 must not call (mask is all zero by definition).
template <class G>
UVIO_HD int ransac_job(G g, RansacSmem &s, const float *p0, const float *p1, int N, double thr, double conf, int max_iters,
                       uint8_t *mask_out) {
  const bool lmeds = N < 15;                    // cv::findFundamentalMat: FM_RANSAC needs >= 15 points, else LMedS
  const int max_attempts = lmeds ? 1000 : 10000;
  for (int i = g.lane(); i < N; i += g.size()) {
    s.x0[i] = p0[2 * i]; s.y0[i] = p0[2 * i + 1];
    s.x1[i] = p1[2 * i]; s.y1[i] = p1[2 * i + 1];
    mask_out[i] = 0;
  }
  if (g.lane() == 0) {
    s.state = 0xFFFFFFFFFFFFFFFFULL;
    s.iter = 0;
    s.max_good = 0;
    s.done = 0;
    s.have_best = 0;
    s.attempts = 0;
    s.min_median = DBL_MAX;
    s.niters = max_iters;
    if (lmeds) {
      int n = ransac_update_iters(conf, 0.45, 7, max_iters);
      s.niters = n < 3 ? 3 : n;
    }
  }
  g.sync();
  if (N == 7) {   // direct solve, every point is an inlier when a model exists
    if (g.lane() == 0) {
      double x0[7], y0[7], x1[7], y1[7];
      for (int i = 0; i < 7; i++) { x0[i] = s.x0[i]; y0[i] = s.y0[i]; x1[i] = s.x1[i]; y1[i] = s.y1[i]; }
      s.have_best = seven_point(x0, y0, x1, y1, s.F[0]) > 0;
    }
    g.sync();
    int ok = s.have_best;
    for (int i = g.lane(); i < N; i += g.size()) mask_out[i] = (uint8_t)ok;
    return ok;
  }
  const float thr_sq = (float)(thr * thr);
  // Hypotheses are evaluated speculatively, a chunk at a time, and then consumed in order with OpenCV's adaptive
  // iteration count; whatever the count cuts off was wasted.  The first good model usually drops the count to a handful,
  // so the first chunk is small and later chunks are sized by what is still allowed.
  int chunk = RS_H_FIRST;
  while (true) {
    ransac_draw_chunk(g, s, N, chunk);
    // a collinear sample ends the chunk: everything after it was drawn from a generator state OpenCV never reaches
    int nvalid = 0;
    while (nvalid < chunk && s.ok[nvalid]) nvalid++;
    for (int h = g.lane(); h < nvalid; h += g.size()) {
      double x0[7], y0[7], x1[7], y1[7];
      for (int i = 0; i < 7; i++) {
        int k = s.idx[h][i];
        x0[i] = s.x0[k]; y0[i] = s.y0[k]; x1[i] = s.x1[k]; y1[i] = s.y1[k];
      }
      s.nmodels[h] = seven_point(x0, y0, x1, y1, s.F[h]);
    }
    for (int i = g.lane(); i < RS_H * RS_MAXROOT; i += g.size()) s.counts[i] = 0;
    g.sync();
    if (!lmeds) {
      // list the models that exist (1 or 3 per sample) so that no lane scores an empty slot
      if (g.lane() == 0) {
        int m = 0;
        for (int h = 0; h < nvalid; h++)
          for (int k = 0; k < s.nmodels[h]; k++) s.slots[m++] = (uint8_t)(h * RS_MAXROOT + k);
        s.nslots = m;
      }
      g.sync();
      int total = s.nslots * N;
      for (int t = g.lane(); t < total; t += g.size()) {
        int m = t / N, p = t - m * N;
        int slot = s.slots[m];
        if (fm_error(s.F[0][0] + 9 * slot, s.x0[p], s.y0[p], s.x1[p], s.y1[p]) <= thr_sq) smem_inc(&s.counts[slot]);
      }
    } else {
      for (int slot = g.lane(); slot < nvalid * RS_MAXROOT; slot += g.size()) {
        int h = slot / RS_MAXROOT, k = slot - h * RS_MAXROOT;
        if (k >= s.nmodels[h]) continue;
        float e[16];
        for (int p = 0; p < N; p++) {       // insertion sort (N < 15)
          float v = fm_error(s.F[h][k], s.x0[p], s.y0[p], s.x1[p], s.y1[p]);
          int q = p;
          while (q > 0 && e[q - 1] > v) { e[q] = e[q - 1]; q--; }
          e[q] = v;
        }
        s.median[slot] = e[N / 2];
      }
    }
    g.sync();
    if (g.lane() == 0) {
      int h = 0;
      for (; h < nvalid; h++) {
        if (s.iter >= s.niters) { s.done = 1; break; }
        for (int k = 0; k < s.nmodels[h]; k++) {
          int slot = h * RS_MAXROOT + k;
          if (!lmeds) {
            int good = (int)s.counts[slot];
            if (good > (s.max_good > 6 ? s.max_good : 6)) {
              s.max_good = good;
              s.have_best = 1;
              for (int i = 0; i < 9; i++) s.bestF[i] = s.F[h][k][i];
              s.niters = ransac_update_iters(conf, (double)(N - good) / N, 7, s.niters);
            }
          } else {
            double med = (double)s.median[slot];
            if (med < s.min_median) {
              s.min_median = med;
              s.have_best = 1;
              for (int i = 0; i < 9; i++) s.bestF[i] = s.F[h][k][i];
            }
          }
        }
        s.iter++;
        s.attempts = 0;
      }
      if (!s.done) {
        if (s.iter >= s.niters) s.done = 1;
        else if (nvalid < chunk) {
          // sample `nvalid` was rejected: OpenCV redraws it from the generator state right after the rejected draw
          s.state = s.state_after[nvalid];
          if (++s.attempts >= max_attempts) s.done = 1;   // getSubset gave up (iter == 0: no model at all)
        } else {
          s.state = s.state_after[chunk - 1];
        }
      }
    }
    g.sync();
    if (s.done) break;
    int left = s.niters - s.iter;
    chunk = left < 1 ? 1 : (left > RS_H ? RS_H : left);
  }
  int ok = s.have_best;
  if (ok) {
    float t = thr_sq;
    if (lmeds) {
      double sigma = 2.5 * 1.4826 * (1 + 5. / (N - 7)) * sqrt(s.min_median);
      if (sigma < 0.001) sigma = 0.001;
      t = (float)(sigma * sigma);
    }
    for (int i = g.lane(); i < N; i += g.size())
      mask_out[i] = fm_error(s.bestF, s.x0[i], s.y0[i], s.x1[i], s.y1[i]) <= t ? 1 : 0;
  }
  g.sync();
  return ok;
}

}  // namespace uvio
