// TEST SCAFFOLDING ONLY — never linked into the product library.
// Compiles the kernel bodies of uvio_b200/csrc/*_body.h with g++ and runs them with the one-lane GroupSeq, so the
// arithmetic of every CUDA kernel can be compared with OpenCV on a machine that has no GPU (pytest -m "not gpu").
// The product path (libuvio_klt.so) contains none of this and fails loudly without a CUDA device.
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../../uvio_b200/csrc/fast_body.h"
#include "../../uvio_b200/csrc/hostlogic.h"
#include "../../uvio_b200/csrc/image_body.h"
#include "../../uvio_b200/csrc/lk_body.h"
#include "../../uvio_b200/csrc/ransac_body.h"
#include "../../uvio_b200/csrc/subpix_body.h"
#include "../../uvio_b200/csrc/track_body.h"

using namespace uvio;

extern "C" {

void hs_equalize(const uint8_t *src, int w, int h, int pitch, uint8_t *dst, int dpitch) {
  GroupSeq g;
  unsigned hist[256] = {0};
  ImgView I{src, w, h, pitch};
  hist_rows(g, hist, I, 0, h);
  uint8_t lut[256];
  equalize_lut(hist, w * h, lut);
  lut_rows(g, lut, I, dst, dpitch, 0, h);
}

void hs_clahe(const uint8_t *src, int w, int h, int pitch, uint8_t *dst, int dpitch, double clip, int tx, int ty) {
  GroupSeq g;
  ImgView I{src, w, h, pitch};
  ClaheGeom q = clahe_geom(w, h, clip, tx, ty);
  std::vector<uint8_t> luts((size_t)tx * ty * 256);
  for (int j = 0; j < ty; j++)
    for (int i = 0; i < tx; i++) {
      unsigned hist[256] = {0};
      clahe_tile_hist(g, hist, I, q, i, j);
      clahe_tile_lut(hist, q, luts.data() + ((size_t)j * tx + i) * 256);
    }
  for (int y = 0; y < h; y++)
    for (int x = 0; x < w; x++) dst[(size_t)y * dpitch + x] = clahe_pixel(luts.data(), q, x, y, src[(size_t)y * pitch + x]);
}

void hs_pyrdown(const uint8_t *src, int w, int h, int pitch, uint8_t *dst, int dpitch) {
  GroupSeq g;
  ImgView I{src, w, h, pitch};
  int dw = (w + 1) / 2, dh = (h + 1) / 2;
  (void)g;
  for (int oy = 0; oy < dh; oy += PD_ROWS)
    for (int ox = 0; ox < dw; ox += PD_COLS) pyrdown_strip(I, dst, dw, dh, dpitch, ox, oy);
}

int hs_pyr_max_level(int w, int h, int req, int win) { return pyr_max_level(w, h, req, win); }

// FAST on the cell (x0,y0,cw,ch) of the image; returns count, list restored to row-major order
int hs_fast(const uint8_t *img, int w, int h, int pitch, int x0, int y0, int cw, int ch, int threshold, int32_t *xy,
            float *resp, int cap) {
  GroupSeq g;
  ImgView I{img, w, h, pitch};
  FastJob job{0, x0, y0, cw, ch, 0, ((cw + 1) / 2) * ((ch + 1) / 2)};
  std::vector<uint32_t> out(job.out_cap);
  unsigned count = 0;
  FastSmem *s = new FastSmem;
  for (int ty = 0; ty < ch; ty += FT_H)
    for (int tx = 0; tx < cw; tx += FT_W) fast_tile(g, *s, I, job, tx, ty, threshold, out.data(), &count);
  delete s;
  std::vector<uint32_t> v(out.begin(), out.begin() + count);
  std::sort(v.begin(), v.end(), [](uint32_t a, uint32_t b) {
    uint32_t ka = ((a >> 12) & 4095) << 12 | (a & 4095), kb = ((b >> 12) & 4095) << 12 | (b & 4095);
    return ka < kb;
  });
  for (unsigned i = 0; i < count && (int)i < cap; i++) {
    xy[2 * i] = v[i] & 4095;
    xy[2 * i + 1] = (v[i] >> 12) & 4095;
    resp[i] = (float)(v[i] >> 24);
  }
  return (int)count;
}

void hs_subpix(const uint8_t *img, int w, int h, int pitch, float *pts, int n, int max_iters, double eps) {
  GroupSeq g;
  ImgView I{img, w, h, pitch};
  float mask[SP_N];
  for (int i = 0; i < SP_W; i++) {
    float y = (float)(i - SP_WIN) / SP_WIN;
    float vy = expf(-y * y);
    for (int j = 0; j < SP_W; j++) {
      float x = (float)(j - SP_WIN) / SP_WIN;
      mask[i * SP_W + j] = (float)(vy * expf(-x * x));
    }
  }
  SubpixSmem *s = new SubpixSmem;
  for (int i = 0; i < n; i++) subpix_point(g, *s, I, mask, max_iters, eps * eps, pts[2 * i], pts[2 * i + 1]);
  delete s;
}

void hs_lk(const uint8_t *const *prev, const uint8_t *const *next, const int *lw, const int *lh, const int *lp,
           int max_level, const float *p0, float *p1, uint8_t *status, int n, int max_count, double eps) {
  GroupSeq g;
  ImgView P[16], N[16];
  for (int l = 0; l <= max_level; l++) {
    P[l] = ImgView{prev[l], lw[l], lh[l], lp[l]};
    N[l] = ImgView{next[l], lw[l], lh[l], lp[l]};
  }
  LkParams prm{max_level, max_count, eps * eps, 1e-4};
  LkSmem *s = new LkSmem;
  for (int i = 0; i < n; i++) lk_track_point(g, *s, P, N, prm, p0[2 * i], p0[2 * i + 1], p1[2 * i], p1[2 * i + 1], status[i]);
  delete s;
}

void hs_undistort(const float *uv, int n, int model, const double *K, const double *D, float *out) {
  CamModel c;
  c.model = model;
  for (int i = 0; i < 4; i++) { c.k[i] = K[i]; c.d[i] = D[i]; }
  for (int i = 0; i < n; i++) undistort_point(c, uv[2 * i], uv[2 * i + 1], out[2 * i], out[2 * i + 1]);
}

int hs_ransac(const float *p0, const float *p1, int n, double thr, double conf, int max_iters, uint8_t *mask) {
  GroupSeq g;
  RansacSmem *s = new RansacSmem;
  std::vector<float> buf((size_t)4 * n + 4);
  s->x0 = buf.data(); s->y0 = s->x0 + n; s->x1 = s->y0 + n; s->y1 = s->x1 + n;
  int ok = 0;
  if (n >= 7) ok = ransac_job(g, *s, p0, p1, n, thr, conf, max_iters, mask);
  else memset(mask, 0, n);
  delete s;
  return ok;
}

// per-cell top-K of the host logic: packed words (x | y<<12 | score<<24, any order) -> K keypoints (x, y, response)
int hs_cell_top_k(const uint32_t *words, int count, int cw, int ch, int K, float *out) {
  std::vector<uint32_t> tmp, tmp2;
  std::vector<KeyPointLite> top;
  cell_top_k(words, (unsigned)count, cw, ch, 0, 0, K, tmp, tmp2, top);
  for (size_t i = 0; i < top.size(); i++) {
    out[3 * i] = top[i].x;
    out[3 * i + 1] = top[i].y;
    out[3 * i + 2] = top[i].response;
  }
  return (int)top.size();
}

// the device-side pipeline's version of the same step (track_body.h): bitonic order by position, then the first K positions
// of libstdc++'s std::sort by response.  heap != 0 forces the heap-sort fallback of introsort over the whole range.
int hs_dv_top_k(const uint32_t *words, int count, int K, int heap, uint32_t *out) {
  int m = 1;
  while (m < count) m <<= 1;
  std::vector<uint32_t> buf((size_t)m, 0xFFFFFFFFu);
  memcpy(buf.data(), words, sizeof(uint32_t) * (size_t)count);
  dv_bitonic_by_position(GroupSeq(), buf.data(), m);
  if (heap) dv_heap_sort(buf.data(), buf.data() + count);
  else dv_sort_prefix(buf.data(), count, K);
  const int kn = count < K ? count : K;
  memcpy(out, buf.data(), sizeof(uint32_t) * (size_t)kn);
  return kn;
}
}  // extern "C"
