// Sub-pixel corner refinement, one execution group (a warp on the device) per point.
//
// Replaces cv::cornerSubPix(img, pts, Size(5,5), Size(-1,-1), {COUNT+EPS, 20, 0.001}) at
// ov_core/src/track/Grider_GRID.h:163-174.  The 13x13 bilinear patch (cv::getRectSubPix 8u->32f)
// is computed lane-parallel; the five fp64 normal-equation sums run as five ordered chains (row-major over the 11x11
// window, exactly the CPU accumulation order) so the refined positions equal the CPU tracker's bit for bit.
#pragma once
#include <float.h>

#include "simt.h"

namespace uvio {

constexpr int SP_WIN = 5;                       // Grider_GRID.h:163 win_size (5,5)
constexpr int SP_W = 2 * SP_WIN + 1;            // 11
constexpr int SP_P = SP_W + 2;                  // 13
constexpr int SP_N = SP_W * SP_W;               // 121

struct SubpixSmem {
  float patch[SP_P * SP_P];
  double gxx[SP_N], gxy[SP_N], gyy[SP_N], t1[SP_N], t2[SP_N];
  double red[5];
};

// cv::getRectSubPix, 8u -> 32f, patch 13x13 centred on (cx, cy); BORDER_REPLICATE outside the image.
template <class G>
UVIO_HD void subpix_patch(G g, SubpixSmem &s, const ImgView &I, float cx, float cy) {
  cx = fsub(cx, (float)(SP_P - 1) * 0.5f);
  cy = fsub(cy, (float)(SP_P - 1) * 0.5f);
  int ipx = floor_f2i(cx), ipy = floor_f2i(cy);
  float a = fsub(cx, (float)ipx), b = fsub(cy, (float)ipy);
  if (0 <= ipx && ipx + SP_P < I.w && 0 <= ipy && ipy + SP_P < I.h) {
    if (a < 0.0001f) a = 0.0001f;
    float a12 = fmul(a, fsub(1.f, b)), a22 = fmul(a, b), b1 = fsub(1.f, b), b2 = b;
    double sd = (1. - (double)a) / (double)a;
    float a0 = fsub(1.f, a);
    for (int i = g.lane(); i < SP_P * SP_P; i += g.size()) {
      int r = i / SP_P, j = i - r * SP_P;
      const uint8_t *p = I.p + (size_t)(ipy + r) * I.pitch + ipx;
      float prev;
      if (j == 0) {
        prev = fmul(a0, fadd(fmul(b1, (float)p[0]), fmul(b2, (float)p[I.pitch])));
      } else {
        float tp = fadd(fmul(a12, (float)p[j]), fmul(a22, (float)p[j + I.pitch]));
        prev = (float)dmul((double)tp, sd);
      }
      float t = fadd(fmul(a12, (float)p[j + 1]), fmul(a22, (float)p[j + 1 + I.pitch]));
      s.patch[i] = fadd(prev, t);
    }
  } else {
    // generic path: four fp32 weights, coordinates clamped to the image (replicated border)
    float a11 = fmul(fsub(1.f, a), fsub(1.f, b)), a12 = fmul(a, fsub(1.f, b)), a21 = fmul(fsub(1.f, a), b), a22 = fmul(a, b);
    float b1 = fsub(1.f, b), b2 = b;
    // horizontal run [rx, rw) is interpolated; columns left of rx / right of rw replicate the edge column
    int rx, rw, ry, rh;
    int x_base = ipx, y_base = ipy;      // source coordinate of patch column/row 0 before clamping
    if (ipx >= 0) rx = 0; else { rx = -ipx; if (rx > SP_P) rx = SP_P; }
    if (ipx < I.w - SP_P) rw = SP_P; else { rw = I.w - ipx - 1; if (rw < 0) rw = 0; }
    if (ipy >= 0) ry = 0; else ry = -ipy;
    if (ipy < I.h - SP_P) rh = SP_P; else { rh = I.h - ipy - 1; if (rh < 0) rh = 0; }
    for (int i = g.lane(); i < SP_P * SP_P; i += g.size()) {
      int r = i / SP_P, j = i - r * SP_P;
      // rows: OpenCV walks a row pointer that only advances while i < rh and sits at row 0 until i reaches ry
      int y0 = clampi(y_base + r, 0, I.h - 1);
      int y1 = (r < ry || r >= rh) ? y0 : clampi(y_base + r + 1, 0, I.h - 1);
      const uint8_t *p = I.p + (size_t)y0 * I.pitch, *p2 = I.p + (size_t)y1 * I.pitch;
      float v;
      if (j < rx) {
        int x = clampi(x_base + rx, 0, I.w - 1);
        v = fadd(fmul((float)p[x], b1), fmul((float)p2[x], b2));
      } else if (j >= rw) {
        int x = clampi(x_base + rw, 0, I.w - 1);
        v = fadd(fmul((float)p[x], b1), fmul((float)p2[x], b2));
      } else {
        int x = x_base + j;
        v = fadd(fadd(fadd(fmul((float)p[x], a11), fmul((float)p[x + 1], a12)), fmul((float)p2[x], a21)),
                 fmul((float)p2[x + 1], a22));
      }
      s.patch[i] = v;
    }
  }
  g.sync();
}

// One point. `mask` = the 11x11 Gaussian weight table (fp32, computed on the host with the CPU's expf).
template <class G>
UVIO_HD void subpix_point(G g, SubpixSmem &s, const ImgView &I, const float *mask, int max_iters, double eps_sq,
                          float &x, float &y) {
  const float cTx = x, cTy = y;
  float cIx = cTx, cIy = cTy;
  int iter = 0;
  double err = 0;
  do {
    subpix_patch(g, s, I, cIx, cIy);
    for (int k = g.lane(); k < SP_N; k += g.size()) {
      int i = k / SP_W, j = k - i * SP_W;
      const float *sp = s.patch + (i + 1) * SP_P + (j + 1);
      double m = (double)mask[k];
      double tgx = (double)fsub(sp[1], sp[-1]);
      double tgy = (double)fsub(sp[SP_P], sp[-SP_P]);
      double gxx = dmul(dmul(tgx, tgx), m), gxy = dmul(dmul(tgx, tgy), m), gyy = dmul(dmul(tgy, tgy), m);
      double px = (double)(j - SP_WIN), py = (double)(i - SP_WIN);
      s.gxx[k] = gxx;
      s.gxy[k] = gxy;
      s.gyy[k] = gyy;
      s.t1[k] = dadd(dmul(gxx, px), dmul(gxy, py));
      s.t2[k] = dadd(dmul(gxy, px), dmul(gyy, py));
    }
    g.sync();
    for (int c = g.lane(); c < 5; c += g.size()) {
      const double *src = c == 0 ? s.gxx : c == 1 ? s.gxy : c == 2 ? s.gyy : c == 3 ? s.t1 : s.t2;
      double acc = 0;
      for (int k = 0; k < SP_N; k++) acc = dadd(acc, src[k]);
      s.red[c] = acc;
    }
    g.sync();
    double a = s.red[0], b = s.red[1], c = s.red[2], bb1 = s.red[3], bb2 = s.red[4];
    g.sync();
    double det = dsub(dmul(a, c), dmul(b, b));
    if (fabs(det) <= DBL_EPSILON * DBL_EPSILON) break;
    double scale = 1.0 / det;
    float c2x = (float)dsub(dadd((double)cIx, dmul(dmul(c, scale), bb1)), dmul(dmul(b, scale), bb2));
    float c2y = (float)dadd(dsub((double)cIy, dmul(dmul(b, scale), bb1)), dmul(dmul(a, scale), bb2));
    float ex = fsub(c2x, cIx), ey = fsub(c2y, cIy);
    err = (double)fadd(fmul(ex, ex), fmul(ey, ey));
    // OpenCV 4.13: an update that leaves the image is discarded (the last in-image iterate is kept)
    if (c2x < 0 || c2x >= (float)I.w || c2y < 0 || c2y >= (float)I.h) break;
    cIx = c2x;
    cIy = c2y;
  } while (++iter < max_iters && err > eps_sq);
  if (fabsf(fsub(cIx, cTx)) > (float)SP_WIN || fabsf(fsub(cIy, cTy)) > (float)SP_WIN) {
    cIx = cTx;
    cIy = cTy;
  }
  x = cIx;
  y = cIy;
}

}  // namespace uvio
