/* ORACLE — TEST INFRASTRUCTURE ONLY.
 *
 * Plain-C restatement of the OpenCV routines the reference's TrackKLT hot path calls (the arithmetic lives in OpenCV,
 * an un-vendored dependency of ov_core: ov_core/CMakeLists.txt:7-10).  Each function names the reference call site it
 * stands for.  Validated against cv2 4.13.0 (IPP off) in tests/test_oracle_cstages.py: integer stages byte-exact,
 * float stages to the tolerances stated there.  Parity is unpinned by the reference's own tests (it has none for this
 * path); it is pinned to OpenCV's behaviour at the reference's call sites.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may load this file's library.
 * Build: make -C oracle  (gcc -O2 -ffp-contract=off: no FMA contraction, like OpenCV's SSE3 baseline build).
 */
#include <float.h>
#include <math.h>
#include <stddef.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define EXPORT __attribute__((visibility("default")))

static inline int reflect101(int p, int len) {
  /* BORDER_REFLECT_101: gfedcb|abcdefgh|gfedcba */
  if (len == 1) return 0;
  while (p < 0 || p >= len) {
    if (p < 0) p = -p;
    else p = 2 * (len - 1) - p;
  }
  return p;
}

static inline int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

/* cvRound: round half to even (SSE cvtsd2si under the default rounding mode) */
static inline int cv_round_f(float v) { return (int)lrintf(v); }
static inline int cv_round_d(double v) { return (int)lrint(v); }
static inline uint8_t sat_u8(int v) { return (uint8_t)(v < 0 ? 0 : (v > 255 ? 255 : v)); }

/* --------------------------------------------------------------------------------------------------------------------
 * cv::equalizeHist — TrackKLT.cpp:59
 * ------------------------------------------------------------------------------------------------------------------*/
EXPORT void oracle_equalize_hist(const uint8_t *src, int w, int h, int sstride, uint8_t *dst, int dstride) {
  int hist[256];
  memset(hist, 0, sizeof(hist));
  for (int y = 0; y < h; y++)
    for (int x = 0; x < w; x++) hist[src[(size_t)y * sstride + x]]++;
  int total = w * h;
  int i = 0;
  while (!hist[i]) ++i;
  if (hist[i] == total) {
    for (int y = 0; y < h; y++) memset(dst + (size_t)y * dstride, i, (size_t)w);
    return;
  }
  float scale = 255.f / (float)(total - hist[i]);
  uint8_t lut[256];
  memset(lut, 0, sizeof(lut));
  int sum = 0;
  for (lut[i++] = 0; i < 256; ++i) {
    sum += hist[i];
    lut[i] = sat_u8(cv_round_f((float)sum * scale));
  }
  for (int y = 0; y < h; y++)
    for (int x = 0; x < w; x++) dst[(size_t)y * dstride + x] = lut[src[(size_t)y * sstride + x]];
}

/* --------------------------------------------------------------------------------------------------------------------
 * cv::createCLAHE(10.0, Size(8,8))->apply — TrackKLT.cpp:61-64
 * ------------------------------------------------------------------------------------------------------------------*/
EXPORT void oracle_clahe(const uint8_t *src, int w, int h, int sstride, uint8_t *dst, int dstride, double clip_limit,
                         int tiles_x, int tiles_y) {
  int wp = w, hp = h;
  if (w % tiles_x != 0 || h % tiles_y != 0) {
    wp = w + (tiles_x - w % tiles_x);
    hp = h + (tiles_y - h % tiles_y);
  }
  /* OpenCV pads BOTH directions as soon as either is not divisible: copyMakeBorder(src, 0, ty-(h%ty), 0, tx-(w%tx)),
   * so a divisible direction then grows by a whole tile count. */
  int tw = wp / tiles_x, th = hp / tiles_y;
  int area = tw * th;
  int clip = 0;
  if (clip_limit > 0.0) {
    clip = (int)(clip_limit * area / 256);
    if (clip < 1) clip = 1;
  }
  float lut_scale = 255.f / (float)area;
  uint8_t *luts = (uint8_t *)malloc((size_t)tiles_x * tiles_y * 256);
  for (int ty = 0; ty < tiles_y; ty++)
    for (int tx = 0; tx < tiles_x; tx++) {
      int hist[256];
      memset(hist, 0, sizeof(hist));
      for (int y = ty * th; y < (ty + 1) * th; y++)
        for (int x = tx * tw; x < (tx + 1) * tw; x++) {
          int sy = y < h ? y : reflect101(y, h), sx = x < w ? x : reflect101(x, w);
          hist[src[(size_t)sy * sstride + sx]]++;
        }
      if (clip > 0) {
        int clipped = 0;
        for (int i = 0; i < 256; i++)
          if (hist[i] > clip) { clipped += hist[i] - clip; hist[i] = clip; }
        int redist = clipped / 256, residual = clipped - redist * 256;
        for (int i = 0; i < 256; i++) hist[i] += redist;
        if (residual != 0) {
          int step = 256 / residual;
          if (step < 1) step = 1;
          for (int i = 0; i < 256 && residual > 0; i += step, residual--) hist[i]++;
        }
      }
      uint8_t *lut = luts + ((size_t)ty * tiles_x + tx) * 256;
      int sum = 0;
      for (int i = 0; i < 256; i++) {
        sum += hist[i];
        lut[i] = sat_u8(cv_round_f((float)sum * lut_scale));
      }
    }
  float inv_tw = 1.0f / (float)tw, inv_th = 1.0f / (float)th;
  for (int y = 0; y < h; y++) {
    float tyf = (float)y * inv_th - 0.5f;
    int ty1 = (int)floorf(tyf), ty2 = ty1 + 1;
    float ya = tyf - (float)ty1, ya1 = 1.0f - ya;
    ty1 = clampi(ty1, 0, tiles_y - 1);
    ty2 = clampi(ty2, 0, tiles_y - 1);
    for (int x = 0; x < w; x++) {
      float txf = (float)x * inv_tw - 0.5f;
      int tx1 = (int)floorf(txf), tx2 = tx1 + 1;
      float xa = txf - (float)tx1, xa1 = 1.0f - xa;
      tx1 = clampi(tx1, 0, tiles_x - 1);
      tx2 = clampi(tx2, 0, tiles_x - 1);
      int v = src[(size_t)y * sstride + x];
      const uint8_t *l11 = luts + ((size_t)ty1 * tiles_x + tx1) * 256, *l12 = luts + ((size_t)ty1 * tiles_x + tx2) * 256;
      const uint8_t *l21 = luts + ((size_t)ty2 * tiles_x + tx1) * 256, *l22 = luts + ((size_t)ty2 * tiles_x + tx2) * 256;
      float res = ((float)l11[v] * xa1 + (float)l12[v] * xa) * ya1 + ((float)l21[v] * xa1 + (float)l22[v] * xa) * ya;
      dst[(size_t)y * dstride + x] = sat_u8(cv_round_f(res));
    }
  }
  free(luts);
}

/* --------------------------------------------------------------------------------------------------------------------
 * cv::pyrDown (one level of cv::buildOpticalFlowPyramid) — TrackKLT.cpp:71
 * out = (w+1)/2 x (h+1)/2 ; 5x5 [1 4 6 4 1]^2 / 256 with (s+128)>>8 ; BORDER_REFLECT_101
 * ------------------------------------------------------------------------------------------------------------------*/
EXPORT void oracle_pyr_down(const uint8_t *src, int w, int h, int sstride, uint8_t *dst, int dstride) {
  static const int k[5] = {1, 4, 6, 4, 1};
  int ow = (w + 1) / 2, oh = (h + 1) / 2;
  for (int y = 0; y < oh; y++)
    for (int x = 0; x < ow; x++) {
      int s = 0;
      for (int j = -2; j <= 2; j++) {
        int sy = reflect101(2 * y + j, h);
        int r = 0;
        for (int i = -2; i <= 2; i++) r += k[i + 2] * src[(size_t)sy * sstride + reflect101(2 * x + i, w)];
        s += k[j + 2] * r;
      }
      dst[(size_t)y * dstride + x] = (uint8_t)((s + 128) >> 8);
    }
}

/* Number of the last pyramid level cv::buildOpticalFlowPyramid creates: stops before a level <= the window. */
EXPORT int oracle_pyr_max_level(int w, int h, int requested, int win) {
  int level = 0;
  while (level < requested) {
    int nw = (w + 1) / 2, nh = (h + 1) / 2;
    if (nw <= win || nh <= win) break;
    w = nw; h = nh; level++;
  }
  return level;
}

/* Scharr derivatives as calcScharrDeriv: int16 interleaved (dx,dy), BORDER_REFLECT_101 */
EXPORT void oracle_scharr(const uint8_t *src, int w, int h, int sstride, int16_t *dst /* h*w*2 */) {
  for (int y = 0; y < h; y++) {
    int ym = reflect101(y - 1, h), yp = reflect101(y + 1, h);
    for (int x = 0; x < w; x++) {
      int xm = reflect101(x - 1, w), xp = reflect101(x + 1, w);
#define P(yy, xx) ((int)src[(size_t)(yy) * sstride + (xx)])
      int dx = 3 * (P(ym, xp) - P(ym, xm)) + 10 * (P(y, xp) - P(y, xm)) + 3 * (P(yp, xp) - P(yp, xm));
      int dy = 3 * (P(yp, xm) - P(ym, xm)) + 10 * (P(yp, x) - P(ym, x)) + 3 * (P(yp, xp) - P(ym, xp));
#undef P
      dst[((size_t)y * w + x) * 2 + 0] = (int16_t)dx;
      dst[((size_t)y * w + x) * 2 + 1] = (int16_t)dy;
    }
  }
}

/* --------------------------------------------------------------------------------------------------------------------
 * cv::FAST(roi, kps, threshold, nonmaxSuppression=true), TYPE_9_16 — Grider_GRID.h:125
 * Emits keypoints row-major; returns the count (at most max_out are written). xy are ROI-relative.
 * ------------------------------------------------------------------------------------------------------------------*/
static const int fast_dx[16] = {0, 1, 2, 3, 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1};
static const int fast_dy[16] = {3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1, 0, 1, 2, 3};

/* largest arc-min over the 16 cyclic arcs of 9, both polarities (0 if none positive) */
static int fast_max_arc(const uint8_t *p, int stride) {
  int d[25];
  int v = p[0];
  for (int k = 0; k < 16; k++) d[k] = v - (int)p[fast_dy[k] * stride + fast_dx[k]];
  for (int k = 16; k < 25; k++) d[k] = d[k - 16];
  int best = -1000;
  for (int k = 0; k < 16; k++) {
    int mn = d[k], mx = d[k];
    for (int j = 1; j < 9; j++) {
      if (d[k + j] < mn) mn = d[k + j];
      if (d[k + j] > mx) mx = d[k + j];
    }
    if (mn > best) best = mn;      /* centre brighter than the whole arc by at least mn */
    if (-mx > best) best = -mx;    /* centre darker than the whole arc by at least -mx */
  }
  return best;
}

EXPORT int oracle_fast(const uint8_t *img, int w, int h, int stride, int threshold, int nonmax, int32_t *xy_out,
                       float *resp_out, int max_out) {
  uint8_t *score = (uint8_t *)calloc((size_t)w * h, 1);
  for (int y = 3; y < h - 3; y++)
    for (int x = 3; x < w - 3; x++) {
      int b = fast_max_arc(img + (size_t)y * stride + x, stride);
      if (b > threshold) score[(size_t)y * w + x] = (uint8_t)(b - 1);   /* cornerScore: largest t that keeps it a corner */
    }
  int n = 0;
  for (int y = 3; y < h - 3; y++)
    for (int x = 3; x < w - 3; x++) {
      int s = score[(size_t)y * w + x];
      if (!s) continue;
      if (nonmax) {
        const uint8_t *c = score + (size_t)y * w + x;
        if (!(s > c[-1] && s > c[1] && s > c[-w - 1] && s > c[-w] && s > c[-w + 1] && s > c[w - 1] && s > c[w] &&
              s > c[w + 1]))
          continue;
      }
      if (n < max_out) {
        xy_out[2 * n] = x;
        xy_out[2 * n + 1] = y;
        resp_out[n] = (float)s;
      }
      n++;
    }
  free(score);
  return n;
}

/* --------------------------------------------------------------------------------------------------------------------
 * cv::undistortPoints (radtan) — cam/CamRadtan.h:102 ; cv::fisheye::undistortPoints — cam/CamEqui.h:121
 * fp64 inside, fp32 in/out.
 * ------------------------------------------------------------------------------------------------------------------*/
EXPORT void oracle_undistort_radtan(const float *uv, int n, const double *K4 /*fx fy cx cy*/, const double *D4, float *out) {
  double fx = K4[0], fy = K4[1], cx = K4[2], cy = K4[3];
  double k1 = D4[0], k2 = D4[1], p1 = D4[2], p2 = D4[3];
  double ifx = 1. / fx, ify = 1. / fy;
  for (int i = 0; i < n; i++) {
    double x = uv[2 * i], y = uv[2 * i + 1];
    x = (x - cx) * ifx;
    y = (y - cy) * ify;
    double x0 = x, y0 = y;
    for (int j = 0; j < 5; j++) {
      double r2 = x * x + y * y;
      double icdist = (1 + ((0 * r2 + 0) * r2 + 0) * r2) / (1 + ((0 * r2 + k2) * r2 + k1) * r2);
      double deltaX = 2 * p1 * x * y + p2 * (r2 + 2 * x * x);
      double deltaY = p1 * (r2 + 2 * y * y) + 2 * p2 * x * y;
      x = (x0 - deltaX) * icdist;
      y = (y0 - deltaY) * icdist;
    }
    out[2 * i] = (float)x;
    out[2 * i + 1] = (float)y;
  }
}

EXPORT void oracle_undistort_equi(const float *uv, int n, const double *K4, const double *D4, float *out) {
  double fx = K4[0], fy = K4[1], cx = K4[2], cy = K4[3];
  for (int i = 0; i < n; i++) {
    double pwx = ((double)uv[2 * i] - cx) / fx, pwy = ((double)uv[2 * i + 1] - cy) / fy;
    double theta_d = sqrt(pwx * pwx + pwy * pwy);
    if (theta_d > M_PI / 2) theta_d = M_PI / 2;
    if (theta_d < -M_PI / 2) theta_d = -M_PI / 2;
    double scale = 0.0;
    int converged = 0;
    if (!(theta_d > 1e-8)) {
      /* fabs(theta_d) <= EPS: the point is returned unscaled */
      out[2 * i] = (float)pwx;
      out[2 * i + 1] = (float)pwy;
      continue;
    }
    double theta = theta_d;
    const double EPS = 1e-8;
    for (int j = 0; j < 10; j++) {
      double t2 = theta * theta, t4 = t2 * t2, t6 = t4 * t2, t8 = t6 * t2;
      double k0t2 = D4[0] * t2, k1t4 = D4[1] * t4, k2t6 = D4[2] * t6, k3t8 = D4[3] * t8;
      double theta_fix = (theta * (1 + k0t2 + k1t4 + k2t6 + k3t8) - theta_d) / (1 + 3 * k0t2 + 5 * k1t4 + 7 * k2t6 + 9 * k3t8);
      theta = theta - theta_fix;
      if (fabs(theta_fix) < EPS) { converged = 1; break; }
    }
    scale = tan(theta) / theta_d;
    int flipped = (theta_d < 0 && theta > 0) || (theta_d > 0 && theta < 0);
    if (converged && !flipped) {
      out[2 * i] = (float)(pwx * scale);
      out[2 * i + 1] = (float)(pwy * scale);
    } else {
      out[2 * i] = -1000000.0f;
      out[2 * i + 1] = -1000000.0f;
    }
  }
}

/* --------------------------------------------------------------------------------------------------------------------
 * cv::getRectSubPix 8u -> 32f (used by cornerSubPix)
 * ------------------------------------------------------------------------------------------------------------------*/
static void get_rect_subpix_8u32f(const uint8_t *src, int sstep, int sw, int sh, float *dst, int dstep, int ww, int wh,
                                  float cx, float cy) {
  cx -= (float)(ww - 1) * 0.5f;
  cy -= (float)(wh - 1) * 0.5f;
  int ipx = (int)floorf(cx), ipy = (int)floorf(cy);
  if (0 <= ipx && ipx + ww < sw && 0 <= ipy && ipy + wh < sh) {
    float a = cx - (float)ipx, b = cy - (float)ipy;
    if (a < 0.0001f) a = 0.0001f;
    float a12 = a * (1.f - b), a22 = a * b, b1 = 1.f - b, b2 = b;
    double s = (1. - a) / a;
    const uint8_t *p = src + (size_t)ipy * sstep + ipx;
    for (int i = 0; i < wh; i++, p += sstep, dst += dstep) {
      float prev = (1 - a) * (b1 * p[0] + b2 * p[sstep]);
      for (int j = 0; j < ww; j++) {
        float t = a12 * p[j + 1] + a22 * p[j + 1 + sstep];
        dst[j] = prev + t;
        prev = (float)(t * s);
      }
    }
    return;
  }
  /* generic path (getRectSubPix_Cn_ with adjustRect): replicated border */
  float a = cx - (float)ipx, b = cy - (float)ipy;
  float a11 = (1.f - a) * (1.f - b), a12 = a * (1.f - b), a21 = (1.f - a) * b, a22 = a * b, b1 = 1.f - b, b2 = b;
  int rx, ry, rw, rh;
  const uint8_t *p = src;
  if (ipx >= 0) { p += ipx; rx = 0; } else { rx = -ipx; if (rx > ww) rx = ww; }
  if (ipx < sw - ww) rw = ww; else { rw = sw - ipx - 1; if (rw < 0) { p += rw; rw = 0; } }
  if (ipy >= 0) { p += (ptrdiff_t)ipy * sstep; ry = 0; } else ry = -ipy;
  if (ipy < sh - wh) rh = wh; else { rh = sh - ipy - 1; if (rh < 0) { p += (ptrdiff_t)rh * sstep; rh = 0; } }
  p -= rx;
  for (int i = 0; i < wh; i++, dst += dstep) {
    const uint8_t *p2 = p + sstep;
    if (i < ry || i >= rh) p2 -= sstep;
    float s0 = p[rx] * b1 + p2[rx] * b2;
    for (int j = 0; j < rx; j++) dst[j] = s0;
    s0 = p[rw] * b1 + p2[rw] * b2;
    for (int j = rw; j < ww; j++) dst[j] = s0;
    for (int j = rx; j < rw; j++) dst[j] = p[j] * a11 + p[j + 1] * a12 + p2[j] * a21 + p2[j + 1] * a22;
    if (i < rh) p = p2;
  }
}

/* --------------------------------------------------------------------------------------------------------------------
 * cv::cornerSubPix(img, pts, Size(5,5), Size(-1,-1), {COUNT+EPS, 20, 0.001}) — Grider_GRID.h:163-174
 * ------------------------------------------------------------------------------------------------------------------*/
EXPORT void oracle_corner_subpix(const uint8_t *img, int w, int h, int stride, float *pts, int n, int win, int max_iters,
                                 double eps) {
  int win_w = win * 2 + 1, win_h = win * 2 + 1;
  float *mask = (float *)malloc(sizeof(float) * win_w * win_h);
  float *buf = (float *)malloc(sizeof(float) * (win_w + 2) * (win_h + 2));
  eps *= eps;
  for (int i = 0; i < win_h; i++) {
    float y = (float)(i - win) / win;
    float vy = expf(-y * y);
    for (int j = 0; j < win_w; j++) {
      float x = (float)(j - win) / win;
      mask[i * win_w + j] = (float)(vy * expf(-x * x));
    }
  }
  for (int pi = 0; pi < n; pi++) {
    float cTx = pts[2 * pi], cTy = pts[2 * pi + 1], cIx = cTx, cIy = cTy;
    int iter = 0;
    double err = 0;
    do {
      double a = 0, b = 0, c = 0, bb1 = 0, bb2 = 0;
      get_rect_subpix_8u32f(img, stride, w, h, buf, win_w + 2, win_w + 2, win_h + 2, cIx, cIy);
      const float *subpix = buf + (win_w + 2) + 1;
      for (int i = 0, k = 0; i < win_h; i++, subpix += win_w + 2) {
        double py = i - win;
        for (int j = 0; j < win_w; j++, k++) {
          double m = mask[k];
          double tgx = subpix[j + 1] - subpix[j - 1];
          double tgy = subpix[j + win_w + 2] - subpix[j - win_w - 2];
          double gxx = tgx * tgx * m, gxy = tgx * tgy * m, gyy = tgy * tgy * m;
          double px = j - win;
          a += gxx; b += gxy; c += gyy;
          bb1 += gxx * px + gxy * py;
          bb2 += gxy * px + gyy * py;
        }
      }
      double det = a * c - b * b;
      if (fabs(det) <= DBL_EPSILON * DBL_EPSILON) break;
      double scale = 1.0 / det;
      float c2x = (float)(cIx + c * scale * bb1 - b * scale * bb2);
      float c2y = (float)(cIy - b * scale * bb1 + a * scale * bb2);
      err = (c2x - cIx) * (c2x - cIx) + (c2y - cIy) * (c2y - cIy);
      /* OpenCV 4.13: an update that leaves the image is discarded (the last in-image iterate is kept) */
      if (c2x < 0 || c2x >= w || c2y < 0 || c2y >= h) break;
      cIx = c2x; cIy = c2y;
    } while (++iter < max_iters && err > eps);
    if (fabsf(cIx - cTx) > win || fabsf(cIy - cTy) > win) { cIx = cTx; cIy = cTy; }
    pts[2 * pi] = cIx;
    pts[2 * pi + 1] = cIy;
  }
  free(mask);
  free(buf);
}

/* --------------------------------------------------------------------------------------------------------------------
 * cv::calcOpticalFlowPyrLK(prevPyr, nextPyr, prevPts, nextPts, status, err, Size(win,win), maxLevel,
 *                          {COUNT|EPS, max_count, eps}, OPTFLOW_USE_INITIAL_FLOW) — TrackKLT.cpp:671,858
 *
 * Pyramid levels are given unpadded; reads outside a level see what OpenCV's padded buffers hold: BORDER_REFLECT_101
 * for the intensity, zero for the Scharr derivative images.
 * accum_mode 0: A and b summed as exact integers, converted once.
 * accum_mode 1: fp32 accumulation in the order of OpenCV's SSE2 (baseline, no FMA) build: per row the first 8 window
 *               columns go through 4 float lanes (two groups of 4 for A; int32 pair sums (x, x+4) for b), columns 8..14
 *               through one scalar running float; lanes are reduced as (l0+l2)+(l1+l3) and added to the scalar.
 * ------------------------------------------------------------------------------------------------------------------*/
typedef struct { const uint8_t *p; int w, h, stride; } Lvl;

static inline int lvl_I(const Lvl *L, int x, int y) {
  return L->p[(size_t)reflect101(y, L->h) * L->stride + reflect101(x, L->w)];
}
static inline void lvl_deriv(const Lvl *L, int x, int y, int *dx, int *dy) {
  if (x < 0 || x >= L->w || y < 0 || y >= L->h) { *dx = 0; *dy = 0; return; }
  int xm = reflect101(x - 1, L->w), xp = reflect101(x + 1, L->w), ym = reflect101(y - 1, L->h), yp = reflect101(y + 1, L->h);
#define P(yy, xx) ((int)L->p[(size_t)(yy) * L->stride + (xx)])
  *dx = 3 * (P(ym, xp) - P(ym, xm)) + 10 * (P(y, xp) - P(y, xm)) + 3 * (P(yp, xp) - P(yp, xm));
  *dy = 3 * (P(yp, xm) - P(ym, xm)) + 10 * (P(yp, x) - P(ym, x)) + 3 * (P(yp, xp) - P(ym, xp));
#undef P
}

#define LK_MAXWIN 31
EXPORT void oracle_lk(const uint8_t *const *prev_lv, const uint8_t *const *next_lv, const int *lw, const int *lh,
                      const int *lstride, int max_level, const float *prev_pts, float *next_pts, uint8_t *status, int n,
                      int win, int max_count, double epsilon, int accum_mode) {
  const int W_BITS = 14;
  const float FLT_SCALE = 1.f / (1 << 20);
  const double min_eig_threshold = 1e-4;
  epsilon *= epsilon;
  float half = (win - 1) * 0.5f;
  static short Iw[LK_MAXWIN * LK_MAXWIN], dIw[LK_MAXWIN * LK_MAXWIN * 2];
  for (int i = 0; i < n; i++) status[i] = 1;
  for (int level = max_level; level >= 0; level--) {
    Lvl I = {prev_lv[level], lw[level], lh[level], lstride[level]};
    Lvl J = {next_lv[level], lw[level], lh[level], lstride[level]};
    for (int pi = 0; pi < n; pi++) {
      float ppx = prev_pts[2 * pi] * (float)(1. / (1 << level)), ppy = prev_pts[2 * pi + 1] * (float)(1. / (1 << level));
      float npx, npy;
      if (level == max_level) {
        npx = next_pts[2 * pi] * (float)(1. / (1 << level));
        npy = next_pts[2 * pi + 1] * (float)(1. / (1 << level));
      } else {
        npx = next_pts[2 * pi] * 2.f;
        npy = next_pts[2 * pi + 1] * 2.f;
      }
      next_pts[2 * pi] = npx;
      next_pts[2 * pi + 1] = npy;
      ppx -= half; ppy -= half;
      int ipx = (int)floorf(ppx), ipy = (int)floorf(ppy);
      if (ipx < -win || ipx >= I.w || ipy < -win || ipy >= I.h) {
        if (level == 0) status[pi] = 0;
        continue;
      }
      float a = ppx - ipx, b = ppy - ipy;
      int iw00 = cv_round_f((1.f - a) * (1.f - b) * (1 << W_BITS));
      int iw01 = cv_round_f(a * (1.f - b) * (1 << W_BITS));
      int iw10 = cv_round_f((1.f - a) * b * (1 << W_BITS));
      int iw11 = (1 << W_BITS) - iw00 - iw01 - iw10;
      long long eA11 = 0, eA12 = 0, eA22 = 0;
      float qA11[4] = {0, 0, 0, 0}, qA12[4] = {0, 0, 0, 0}, qA22[4] = {0, 0, 0, 0}, sA11 = 0, sA12 = 0, sA22 = 0;
      int nsimd = (win >= 8) ? (win / 8) * 8 : 0;
      for (int y = 0; y < win; y++)
        for (int x = 0; x < win; x++) {
          int X = ipx + x, Y = ipy + y;
          int ival = (lvl_I(&I, X, Y) * iw00 + lvl_I(&I, X + 1, Y) * iw01 + lvl_I(&I, X, Y + 1) * iw10 +
                      lvl_I(&I, X + 1, Y + 1) * iw11 + (1 << (W_BITS - 5 - 1))) >> (W_BITS - 5);
          int dx00, dy00, dx01, dy01, dx10, dy10, dx11, dy11;
          lvl_deriv(&I, X, Y, &dx00, &dy00);
          lvl_deriv(&I, X + 1, Y, &dx01, &dy01);
          lvl_deriv(&I, X, Y + 1, &dx10, &dy10);
          lvl_deriv(&I, X + 1, Y + 1, &dx11, &dy11);
          int ixval = (dx00 * iw00 + dx01 * iw01 + dx10 * iw10 + dx11 * iw11 + (1 << (W_BITS - 1))) >> W_BITS;
          int iyval = (dy00 * iw00 + dy01 * iw01 + dy10 * iw10 + dy11 * iw11 + (1 << (W_BITS - 1))) >> W_BITS;
          Iw[y * win + x] = (short)ival;
          dIw[(y * win + x) * 2] = (short)ixval;
          dIw[(y * win + x) * 2 + 1] = (short)iyval;
          eA11 += (long long)ixval * ixval; eA12 += (long long)ixval * iyval; eA22 += (long long)iyval * iyval;
          if (x < nsimd) {
            int l = x & 3;
            float fx = (float)ixval, fy = (float)iyval;
            qA22[l] = fy * fy + qA22[l]; qA12[l] = fx * fy + qA12[l]; qA11[l] = fx * fx + qA11[l];
          } else {
            sA11 += (float)(ixval * ixval); sA12 += (float)(ixval * iyval); sA22 += (float)(iyval * iyval);
          }
        }
      float A11, A12, A22;
      if (accum_mode == 0) {
        A11 = (float)eA11 * FLT_SCALE; A12 = (float)eA12 * FLT_SCALE; A22 = (float)eA22 * FLT_SCALE;
      } else {
        sA11 += (qA11[0] + qA11[2]) + (qA11[1] + qA11[3]);
        sA12 += (qA12[0] + qA12[2]) + (qA12[1] + qA12[3]);
        sA22 += (qA22[0] + qA22[2]) + (qA22[1] + qA22[3]);
        A11 = sA11 * FLT_SCALE; A12 = sA12 * FLT_SCALE; A22 = sA22 * FLT_SCALE;
      }
      float D = A11 * A22 - A12 * A12;
      float minEig = (A22 + A11 - sqrtf((A11 - A22) * (A11 - A22) + 4.f * A12 * A12)) / (2 * win * win);
      if (minEig < min_eig_threshold || D < FLT_EPSILON) {
        if (level == 0) status[pi] = 0;
        continue;
      }
      D = 1.f / D;
      npx -= half; npy -= half;
      float pdx = 0, pdy = 0;
      for (int j = 0; j < max_count; j++) {
        int inx = (int)floorf(npx), iny = (int)floorf(npy);
        if (inx < -win || inx >= J.w || iny < -win || iny >= J.h) {
          if (level == 0) status[pi] = 0;
          break;
        }
        a = npx - inx; b = npy - iny;
        iw00 = cv_round_f((1.f - a) * (1.f - b) * (1 << W_BITS));
        iw01 = cv_round_f(a * (1.f - b) * (1 << W_BITS));
        iw10 = cv_round_f((1.f - a) * b * (1 << W_BITS));
        iw11 = (1 << W_BITS) - iw00 - iw01 - iw10;
        long long eb1 = 0, eb2 = 0;
        float qb[4] = {0, 0, 0, 0}, qc[4] = {0, 0, 0, 0}, sb1 = 0, sb2 = 0;
        for (int y = 0; y < win; y++) {
          int diffrow[LK_MAXWIN];
          for (int x = 0; x < win; x++) {
            int X = inx + x, Y = iny + y;
            diffrow[x] = ((lvl_I(&J, X, Y) * iw00 + lvl_I(&J, X + 1, Y) * iw01 + lvl_I(&J, X, Y + 1) * iw10 +
                           lvl_I(&J, X + 1, Y + 1) * iw11 + (1 << (W_BITS - 5 - 1))) >> (W_BITS - 5)) - Iw[y * win + x];
            eb1 += (long long)diffrow[x] * dIw[(y * win + x) * 2];
            eb2 += (long long)diffrow[x] * dIw[(y * win + x) * 2 + 1];
          }
          for (int x0 = 0; x0 < nsimd; x0 += 8) {
            /* qb0 lanes: [bx(0,4) by(0,4) bx(1,5) by(1,5)], qb1 lanes: [bx(2,6) by(2,6) bx(3,7) by(3,7)] */
            for (int l = 0; l < 4; l++) {
              int xa = x0 + l, xb = x0 + l + 4;
              int sx = diffrow[xa] * dIw[(y * win + xa) * 2] + diffrow[xb] * dIw[(y * win + xb) * 2];
              int sy = diffrow[xa] * dIw[(y * win + xa) * 2 + 1] + diffrow[xb] * dIw[(y * win + xb) * 2 + 1];
              qb[l] += (float)sx;
              qc[l] += (float)sy;
            }
          }
          for (int x = nsimd; x < win; x++) {
            sb1 += (float)(diffrow[x] * dIw[(y * win + x) * 2]);
            sb2 += (float)(diffrow[x] * dIw[(y * win + x) * 2 + 1]);
          }
        }
        float b1, b2;
        if (accum_mode == 0) {
          b1 = (float)eb1 * FLT_SCALE; b2 = (float)eb2 * FLT_SCALE;
        } else {
          /* qb0+qb1 = [bx01 by01 bx23 by23] with bx01 = qb[0]+qb[2], bx23 = qb[1]+qb[3]; reduce = bx01 + bx23 */
          sb1 += ((qb[0] + qb[2]) + 0.f) + ((qb[1] + qb[3]) + 0.f);
          sb2 += ((qc[0] + qc[2]) + 0.f) + ((qc[1] + qc[3]) + 0.f);
          b1 = sb1 * FLT_SCALE; b2 = sb2 * FLT_SCALE;
        }
        float dx = (float)((A12 * b2 - A22 * b1) * D), dy = (float)((A12 * b1 - A11 * b2) * D);
        npx += dx; npy += dy;
        next_pts[2 * pi] = npx + half;
        next_pts[2 * pi + 1] = npy + half;
        if ((double)dx * dx + (double)dy * dy <= epsilon) break;
        if (j > 0 && fabs(dx + pdx) < 0.01 && fabs(dy + pdy) < 0.01) {
          next_pts[2 * pi] -= dx * 0.5f;
          next_pts[2 * pi + 1] -= dy * 0.5f;
          break;
        }
        pdx = dx; pdy = dy;
      }
      if (status[pi] && level == 0) {
        /* err is requested by the reference (TrackKLT.cpp:669-671), so OpenCV re-checks the final position */
        float fx = next_pts[2 * pi] - half, fy = next_pts[2 * pi + 1] - half;
        int ix = (int)floorf(fx), iy = (int)floorf(fy);
        if (ix < -win || ix >= J.w || iy < -win || iy >= J.h) status[pi] = 0;
      }
    }
  }
}
