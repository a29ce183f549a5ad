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
