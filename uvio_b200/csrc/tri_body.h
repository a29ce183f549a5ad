// Batched feature triangulation: the per-feature body (one GPU thread per feature; the same text runs on the host in
// tests/hostsim).  Restates ov_core/src/feat/FeatureInitializer.cpp:
//   single_triangulation     :30-112    tri_linear
//   single_triangulation_1d  :114-195   tri_linear_1d
//   single_gaussnewton       :197-360   tri_gaussnewton
//   compute_error            :362-421   tri_cost
// fp64 throughout, sums in the reference's order (cameras, then measurements in time order), including its float casts
// (z and the residual are Eigen::Matrix<float,2,1>, :283-285, :413-415).  The three 3x3 solves Eigen does with
// colPivHouseholderQr / JacobiSVD / HouseholderQR are done here by Gaussian elimination with partial pivoting, a cyclic
// Jacobi eigenvalue iteration (A is symmetric positive semi-definite, so singular values = eigenvalues) and the closed form
// of "component of p perpendicular to x" -- same quantities, different last bits (tolerance stated in tests/test_tri.py).
#pragma once
#include <math.h>

#include "simt.h"

namespace uvio {

struct TriClone { unsigned long long cam; double t; double R[9]; double p[3]; };     // == UvioClonePose
struct TriOptions {                                                                  // == UvioTriOptions
  int triangulate_1d, refine_features, max_runs;
  double init_lamda, max_lamda, min_dx, min_dcost, lam_mult, min_dist, max_dist, max_baseline, max_cond_number;
};
struct TriResult {                                                                   // == UvioTriResult
  int ok_triangulation, ok_refine, anchor_cam_id, n_meas;
  double anchor_clone_timestamp;
  double p_FinA[3], p_FinG[3];
};
struct TriFeatures {            // CSR of uvio_featdb_fetch
  int n_features;
  const int *row0, *row_cam, *off;
  const double *t;
  const float *uvn;
};

UVIO_HD const TriClone *tri_find(const TriClone *c, int n, int cam, double t) {
  for (int i = 0; i < n; i++)
    if ((int)c[i].cam == cam && c[i].t == t) return &c[i];
  return nullptr;
}

// x = A^-1 b for a 3x3 system (partial pivoting); returns false when a pivot vanishes (x gets NaN, as a QR solve would)
UVIO_HD bool tri_solve3(const double A[9], const double b[3], double x[3]) {
  double M[3][4];
  for (int r = 0; r < 3; r++) {
    for (int c = 0; c < 3; c++) M[r][c] = A[3 * r + c];
    M[r][3] = b[r];
  }
  for (int k = 0; k < 3; k++) {
    int piv = k;
    for (int r = k + 1; r < 3; r++)
      if (fabs(M[r][k]) > fabs(M[piv][k])) piv = r;
    if (piv != k)
      for (int c = 0; c < 4; c++) { double tmp = M[k][c]; M[k][c] = M[piv][c]; M[piv][c] = tmp; }
    if (M[k][k] == 0.0 || M[k][k] != M[k][k]) { x[0] = x[1] = x[2] = NAN; return false; }
    for (int r = k + 1; r < 3; r++) {
      const double f = M[r][k] / M[k][k];
      for (int c = k; c < 4; c++) M[r][c] -= f * M[k][c];
    }
  }
  for (int r = 2; r >= 0; r--) {
    double s = M[r][3];
    for (int c = r + 1; c < 3; c++) s -= M[r][c] * x[c];
    x[r] = s / M[r][r];
  }
  return true;
}

// largest / smallest eigenvalue of a symmetric 3x3 matrix (cyclic Jacobi)
UVIO_HD double tri_cond_sym3(const double Ain[9]) {
  double a[3][3];
  for (int r = 0; r < 3; r++)
    for (int c = 0; c < 3; c++) a[r][c] = Ain[3 * r + c];
  for (int sweep = 0; sweep < 30; sweep++) {
    const double offd = fabs(a[0][1]) + fabs(a[0][2]) + fabs(a[1][2]);
    if (offd < 1e-300 || offd <= 1e-18 * (fabs(a[0][0]) + fabs(a[1][1]) + fabs(a[2][2]))) break;
    for (int p = 0; p < 2; p++)
      for (int q = p + 1; q < 3; q++) {
        if (a[p][q] == 0.0) continue;
        const double theta = (a[q][q] - a[p][p]) / (2.0 * a[p][q]);
        const double tt = (theta >= 0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1.0));
        const double c = 1.0 / sqrt(tt * tt + 1.0), s = tt * c;
        for (int k = 0; k < 3; k++) {                      // A <- A J
          const double akp = a[k][p], akq = a[k][q];
          a[k][p] = c * akp - s * akq;
          a[k][q] = s * akp + c * akq;
        }
        for (int k = 0; k < 3; k++) {                      // A <- J^T A
          const double apk = a[p][k], aqk = a[q][k];
          a[p][k] = c * apk - s * aqk;
          a[q][k] = s * apk + c * aqk;
        }
      }
  }
  double e0 = fabs(a[0][0]), e1 = fabs(a[1][1]), e2 = fabs(a[2][2]);
  const double hi = fmax(e0, fmax(e1, e2)), lo = fmin(e0, fmin(e1, e2));
  return hi / lo;
}

struct TriRel { double R[9]; double pCinA[3]; };   // R_AtoCi and p_CiinA of one measurement
UVIO_HD void tri_relative(const TriClone &A, const TriClone &C, TriRel &o) {
  for (int r = 0; r < 3; r++)                               // R_AtoCi = R_GtoCi * R_GtoA^T  (:68-69)
    for (int c = 0; c < 3; c++) o.R[3 * r + c] = C.R[3 * r] * A.R[3 * c] + C.R[3 * r + 1] * A.R[3 * c + 1] + C.R[3 * r + 2] * A.R[3 * c + 2];
  const double d[3] = {C.p[0] - A.p[0], C.p[1] - A.p[1], C.p[2] - A.p[2]};
  for (int r = 0; r < 3; r++) o.pCinA[r] = A.R[3 * r] * d[0] + A.R[3 * r + 1] * d[1] + A.R[3 * r + 2] * d[2];   // :70-71
}

// cameras are visited from the highest id down (std::unordered_map order of libstdc++ for a feature first seen in camera 0)
#define TRI_FOR_EACH_MEAS(F, f, r, m) \
  for (int r = (F).row0[(f) + 1] - 1; r >= (F).row0[f]; r--) \
    for (int m = (F).off[r]; m < (F).off[r + 1]; m++)

UVIO_HD bool tri_anchor(const TriFeatures &F, int f, const TriClone *clones, int nc, TriResult &out, const TriClone *&anchor, int &anchor_meas) {
  int most = 0, arow = -1, total = 0;
  for (int r = F.row0[f + 1] - 1; r >= F.row0[f]; r--) {    // :36-46
    const int n = F.off[r + 1] - F.off[r];
    total += n;
    if (n > most) { most = n; arow = r; }
  }
  out.n_meas = total;
  if (arow < 0) return false;
  out.anchor_cam_id = F.row_cam[arow];
  anchor_meas = F.off[arow + 1] - 1;
  out.anchor_clone_timestamp = F.t[anchor_meas];            // timestamps.at(anchor_cam_id).back()
  anchor = tri_find(clones, nc, out.anchor_cam_id, out.anchor_clone_timestamp);
  return anchor != nullptr;
}

UVIO_HD void tri_to_global(const TriClone &A, TriResult &out) {   // p_FinG = R_GtoA^T p_FinA + p_AinG
  for (int r = 0; r < 3; r++) out.p_FinG[r] = A.R[r] * out.p_FinA[0] + A.R[3 + r] * out.p_FinA[1] + A.R[6 + r] * out.p_FinA[2] + A.p[r];
}

// single_triangulation (:30-112).  Returns 1 / 0, or -1 when a clone pose is missing.
UVIO_HD int tri_linear(const TriFeatures &F, int f, const TriClone *clones, int nc, const TriOptions &o, TriResult &out, const TriClone *&anchor) {
  int am;
  if (!tri_anchor(F, f, clones, nc, out, anchor, am)) return -1;
  double A[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0}, b[3] = {0, 0, 0};
  TRI_FOR_EACH_MEAS(F, f, r, m) {
    const TriClone *C = tri_find(clones, nc, F.row_cam[r], F.t[m]);
    if (!C) return -1;
    TriRel rel;
    tri_relative(*anchor, *C, rel);
    const double u = (double)F.uvn[2 * m], v = (double)F.uvn[2 * m + 1];
    double bi[3];                                           // b_i = R_AtoCi^T [u v 1]^T, normalised (:74-77)
    for (int k = 0; k < 3; k++) bi[k] = rel.R[k] * u + rel.R[3 + k] * v + rel.R[6 + k];
    const double nrm = sqrt(bi[0] * bi[0] + bi[1] * bi[1] + bi[2] * bi[2]);
    for (int k = 0; k < 3; k++) bi[k] = bi[k] / nrm;
    // Bperp = skew(b_i); Ai = Bperp^T Bperp  (:78-81)
    const double B[9] = {0, -bi[2], bi[1], bi[2], 0, -bi[0], -bi[1], bi[0], 0};
    double Ai[9];
    for (int rr = 0; rr < 3; rr++)
      for (int cc = 0; cc < 3; cc++) Ai[3 * rr + cc] = B[rr] * B[cc] + B[3 + rr] * B[3 + cc] + B[6 + rr] * B[6 + cc];
    for (int k = 0; k < 9; k++) A[k] += Ai[k];
    for (int rr = 0; rr < 3; rr++) b[rr] += Ai[3 * rr] * rel.pCinA[0] + Ai[3 * rr + 1] * rel.pCinA[1] + Ai[3 * rr + 2] * rel.pCinA[2];
  }
  double p[3];
  tri_solve3(A, b, p);                                      // :87
  const double cond = tri_cond_sym3(A);                     // :90-94
  const double nrm = sqrt(p[0] * p[0] + p[1] * p[1] + p[2] * p[2]);
  if (fabs(cond) > o.max_cond_number || p[2] < o.min_dist || p[2] > o.max_dist || nrm != nrm || !(fabs(cond) == fabs(cond))) return 0;   // :102-105
  for (int k = 0; k < 3; k++) out.p_FinA[k] = p[k];
  tri_to_global(*anchor, out);
  return 1;
}

// single_triangulation_1d (:114-195)
UVIO_HD int tri_linear_1d(const TriFeatures &F, int f, const TriClone *clones, int nc, const TriOptions &o, TriResult &out, const TriClone *&anchor) {
  int am;
  if (!tri_anchor(F, f, clones, nc, out, anchor, am)) return -1;
  double A = 0.0, b = 0.0;
  double ba[3] = {(double)F.uvn[2 * am], (double)F.uvn[2 * am + 1], 1.0};   // :139-142
  const double bn = sqrt(ba[0] * ba[0] + ba[1] * ba[1] + ba[2] * ba[2]);
  for (int k = 0; k < 3; k++) ba[k] = ba[k] / bn;
  TRI_FOR_EACH_MEAS(F, f, r, m) {
    if (m == am) continue;                                  // :151-152 the anchor bearing itself
    const TriClone *C = tri_find(clones, nc, F.row_cam[r], F.t[m]);
    if (!C) return -1;
    TriRel rel;
    tri_relative(*anchor, *C, rel);
    const double u = (double)F.uvn[2 * m], v = (double)F.uvn[2 * m + 1];
    double bi[3];
    for (int k = 0; k < 3; k++) bi[k] = rel.R[k] * u + rel.R[3 + k] * v + rel.R[6 + k];
    const double nrm = sqrt(bi[0] * bi[0] + bi[1] * bi[1] + bi[2] * bi[2]);
    for (int k = 0; k < 3; k++) bi[k] = bi[k] / nrm;
    const double B[9] = {0, -bi[2], bi[1], bi[2], 0, -bi[0], -bi[1], bi[0], 0};
    double Ba[3], Bp[3];                                    // Bperp * bearing_inA, Bperp * p_CiinA  (:172-174)
    for (int rr = 0; rr < 3; rr++) {
      Ba[rr] = B[3 * rr] * ba[0] + B[3 * rr + 1] * ba[1] + B[3 * rr + 2] * ba[2];
      Bp[rr] = B[3 * rr] * rel.pCinA[0] + B[3 * rr + 1] * rel.pCinA[1] + B[3 * rr + 2] * rel.pCinA[2];
    }
    A += Ba[0] * Ba[0] + Ba[1] * Ba[1] + Ba[2] * Ba[2];
    b += Ba[0] * Bp[0] + Ba[1] * Bp[1] + Ba[2] * Bp[2];
  }
  const double depth = b / A;                               // :179-180
  const double p[3] = {depth * ba[0], depth * ba[1], depth * ba[2]};
  const double nrm = sqrt(p[0] * p[0] + p[1] * p[1] + p[2] * p[2]);
  if (p[2] < o.min_dist || p[2] > o.max_dist || nrm != nrm) return 0;   // :183-185
  for (int k = 0; k < 3; k++) out.p_FinA[k] = p[k];
  tri_to_global(*anchor, out);
  return 1;
}

// compute_error (:362-421) and, with H != null, the Jacobian / gradient / Hessian sums of single_gaussnewton (:233-292)
UVIO_HD double tri_cost(const TriFeatures &F, int f, const TriClone *clones, int nc, const TriClone &anchor, double alpha, double beta, double rho,
                        double *Hess, double *grad, bool &missing) {
  double err = 0;
  if (Hess)
    for (int k = 0; k < 9; k++) Hess[k] = 0;
  if (grad) grad[0] = grad[1] = grad[2] = 0;
  TRI_FOR_EACH_MEAS(F, f, r, m) {
    const TriClone *C = tri_find(clones, nc, F.row_cam[r], F.t[m]);
    if (!C) { missing = true; return err; }
    TriRel rel;
    tri_relative(anchor, *C, rel);
    double pA[3];                                           // p_AinCi = -R_AtoCi p_CiinA
    for (int rr = 0; rr < 3; rr++) pA[rr] = -(rel.R[3 * rr] * rel.pCinA[0] + rel.R[3 * rr + 1] * rel.pCinA[1] + rel.R[3 * rr + 2] * rel.pCinA[2]);
    const double hi1 = rel.R[0] * alpha + rel.R[1] * beta + rel.R[2] + rho * pA[0];
    const double hi2 = rel.R[3] * alpha + rel.R[4] * beta + rel.R[5] + rho * pA[1];
    const double hi3 = rel.R[6] * alpha + rel.R[7] * beta + rel.R[8] + rho * pA[2];
    const float z0 = (float)(hi1 / hi3), z1 = (float)(hi2 / hi3);           // Eigen::Matrix<float,2,1> z
    const float r0 = F.uvn[2 * m] - z0, r1 = F.uvn[2 * m + 1] - z1;         // float residual
    const float rn = sqrtf(r0 * r0 + r1 * r1);                              // res.norm() in float
    err += (double)rn * (double)rn;                                          // pow(res.norm(), 2)
    if (Hess) {
      const double h32 = hi3 * hi3;
      const double H[6] = {(rel.R[0] * hi3 - hi1 * rel.R[6]) / h32, (rel.R[1] * hi3 - hi1 * rel.R[7]) / h32, (pA[0] * hi3 - hi1 * pA[2]) / h32,
                           (rel.R[3] * hi3 - hi2 * rel.R[6]) / h32, (rel.R[4] * hi3 - hi2 * rel.R[7]) / h32, (pA[1] * hi3 - hi2 * pA[2]) / h32};
      for (int k = 0; k < 3; k++) grad[k] += H[k] * (double)r0 + H[3 + k] * (double)r1;
      for (int a = 0; a < 3; a++)
        for (int c = 0; c < 3; c++) Hess[3 * a + c] += H[a] * H[c] + H[3 + a] * H[3 + c];
    }
  }
  return err;
}

// single_gaussnewton (:197-360); out.p_FinA holds the linear solution on entry
UVIO_HD int tri_gaussnewton(const TriFeatures &F, int f, const TriClone *clones, int nc, const TriOptions &o, const TriClone &anchor, TriResult &out) {
  double rho = 1 / out.p_FinA[2], alpha = out.p_FinA[0] / out.p_FinA[2], beta = out.p_FinA[1] / out.p_FinA[2];
  double lam = o.init_lamda, eps = 10000;
  int runs = 0;
  bool recompute = true, missing = false;
  double Hess[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0}, grad[3] = {0, 0, 0};
  double cost_old = tri_cost(F, f, clones, nc, anchor, alpha, beta, rho, nullptr, nullptr, missing);
  int guard = 0;
  while (runs < o.max_runs && lam < o.max_lamda && eps > o.min_dx && guard++ < 10000) {
    if (recompute) tri_cost(F, f, clones, nc, anchor, alpha, beta, rho, Hess, grad, missing);
    double Hl[9];
    for (int k = 0; k < 9; k++) Hl[k] = Hess[k];
    for (int k = 0; k < 3; k++) Hl[4 * k] *= (1.0 + lam);
    double dx[3];
    tri_solve3(Hl, grad, dx);
    const double cost = tri_cost(F, f, clones, nc, anchor, alpha + dx[0], beta + dx[1], rho + dx[2], nullptr, nullptr, missing);
    if (cost <= cost_old && (cost_old - cost) / cost_old < o.min_dcost) {   // :313-319
      alpha += dx[0]; beta += dx[1]; rho += dx[2];
      eps = 0;
      break;
    }
    if (cost <= cost_old) {                                                  // :323-331
      recompute = true;
      cost_old = cost;
      alpha += dx[0]; beta += dx[1]; rho += dx[2];
      runs++;
      lam = lam / o.lam_mult;
      eps = sqrt(dx[0] * dx[0] + dx[1] * dx[1] + dx[2] * dx[2]);
    } else {
      recompute = false;
      lam = lam * o.lam_mult;
    }
  }
  if (missing) return -1;
  out.p_FinA[0] = alpha / rho;                                               // :338-340
  out.p_FinA[1] = beta / rho;
  out.p_FinA[2] = 1 / rho;
  const double pn = sqrt(out.p_FinA[0] * out.p_FinA[0] + out.p_FinA[1] * out.p_FinA[1] + out.p_FinA[2] * out.p_FinA[2]);
  // max baseline perpendicular to the bearing (:343-361): |N^T p| with N an orthonormal basis of the plane normal to p_FinA
  double base_max = 0.0;
  TRI_FOR_EACH_MEAS(F, f, r, m) {
    const TriClone *C = tri_find(clones, nc, F.row_cam[r], F.t[m]);
    if (!C) return -1;
    TriRel rel;
    tri_relative(anchor, *C, rel);
    const double dot = (rel.pCinA[0] * out.p_FinA[0] + rel.pCinA[1] * out.p_FinA[1] + rel.pCinA[2] * out.p_FinA[2]) / pn;
    const double sq = rel.pCinA[0] * rel.pCinA[0] + rel.pCinA[1] * rel.pCinA[1] + rel.pCinA[2] * rel.pCinA[2] - dot * dot;
    const double base = sq > 0 ? sqrt(sq) : 0.0;
    if (base > base_max) base_max = base;
  }
  if (out.p_FinA[2] < o.min_dist || out.p_FinA[2] > o.max_dist || (pn / base_max) > o.max_baseline || pn != pn) return 0;   // :369-372
  tri_to_global(anchor, out);
  return 1;
}

// what the updaters do per feature (UpdaterMSCKF.cpp:104-121): triangulate, then refine if asked to
UVIO_HD void tri_feature(const TriFeatures &F, int f, const TriClone *clones, int nc, const TriOptions &o, TriResult &out) {
  out.ok_triangulation = out.ok_refine = 0;
  out.anchor_cam_id = -1;
  out.n_meas = 0;
  out.anchor_clone_timestamp = 0;
  for (int k = 0; k < 3; k++) out.p_FinA[k] = out.p_FinG[k] = 0;
  const TriClone *anchor = nullptr;
  out.ok_triangulation = o.triangulate_1d ? tri_linear_1d(F, f, clones, nc, o, out, anchor) : tri_linear(F, f, clones, nc, o, out, anchor);
  if (out.ok_triangulation == 1 && o.refine_features) out.ok_refine = tri_gaussnewton(F, f, clones, nc, o, *anchor, out);
}

}  // namespace uvio
