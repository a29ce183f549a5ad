// Image-wide integer passes: histogram equalisation, CLAHE, pyrDown.  
//
//   cv::equalizeHist                         ov_core/src/track/TrackKLT.cpp:59
//   cv::createCLAHE(10.0, Size(8,8))->apply  ov_core/src/track/TrackKLT.cpp:61-64
//   cv::buildOpticalFlowPyramid (pyrDown)    ov_core/src/track/TrackKLT.cpp:71
//
// All three are HBM-bound byte passes; bodies are written per thread block (GroupBlock on the device).
#pragma once
#include "simt.h"

namespace uvio {

#ifdef __CUDA_ARCH__
__device__ __forceinline__ void smem_inc(unsigned *p) { atomicAdd(p, 1u); }
__device__ __forceinline__ unsigned gmem_add(unsigned *p, unsigned v) { return atomicAdd(p, v); }
#else
inline void smem_inc(unsigned *p) { ++*p; }
inline unsigned gmem_add(unsigned *p, unsigned v) { unsigned o = *p; *p += v; return o; }
#endif

// ---- histogram of rows [row0,row1) of one image into a block-private histogram ---------------------------------------
template <class G>
UVIO_HD void hist_rows(G g, unsigned *sh /*256, zeroed*/, const ImgView &I, int row0, int row1) {
  const int w4 = I.w >> 2;   // pitch is a multiple of 16, base 16-byte aligned: rows can be read as 32-bit words
  for (int y = row0; y < row1; y++) {
    const uint32_t *r = (const uint32_t *)(I.p + (size_t)y * I.pitch);
    for (int x = g.lane(); x < w4; x += g.size()) {
      uint32_t v = r[x];
      smem_inc(&sh[v & 255]);
      smem_inc(&sh[(v >> 8) & 255]);
      smem_inc(&sh[(v >> 16) & 255]);
      smem_inc(&sh[v >> 24]);
    }
    for (int x = (w4 << 2) + g.lane(); x < I.w; x += g.size()) smem_inc(&sh[I.p[(size_t)y * I.pitch + x]]);
  }
}

// ---- equalizeHist LUT from a full-image histogram (one lane; 256 steps) ----------------------------------------------
UVIO_HD void equalize_lut(const unsigned *hist, int total, uint8_t *lut) {
  int i = 0;
  while (!hist[i]) ++i;
  if ((int)hist[i] == total) {
    for (int j = 0; j < 256; j++) lut[j] = (uint8_t)i;
    return;
  }
  float scale = 255.f / (float)(total - (int)hist[i]);
  for (int j = 0; j <= i; j++) lut[j] = 0;
  int sum = 0;
  for (++i; i < 256; ++i) {
    sum += (int)hist[i];
    lut[i] = sat_u8(round_f2i(fmul((float)sum, scale)));
  }
}

// ---- apply a 256-entry LUT to rows [row0,row1) ------------------------------------------------------------------------
template <class G>
UVIO_HD void lut_rows(G g, const uint8_t *lut /*shared*/, const ImgView &I, uint8_t *dst, int dpitch, int row0, int row1) {
  const int w4 = I.w >> 2;
  for (int y = row0; y < row1; y++) {
    const uint32_t *r = (const uint32_t *)(I.p + (size_t)y * I.pitch);
    uint32_t *d = (uint32_t *)(dst + (size_t)y * dpitch);
    for (int x = g.lane(); x < w4; x += g.size()) {
      uint32_t v = r[x];
      d[x] = (uint32_t)lut[v & 255] | ((uint32_t)lut[(v >> 8) & 255] << 8) | ((uint32_t)lut[(v >> 16) & 255] << 16) |
             ((uint32_t)lut[v >> 24] << 24);
    }
    for (int x = (w4 << 2) + g.lane(); x < I.w; x += g.size()) dst[(size_t)y * dpitch + x] = lut[I.p[(size_t)y * I.pitch + x]];
  }
}

// ---- CLAHE ------------------------------------------------------------------------------------------------------------
struct ClaheGeom {
  int tiles_x, tiles_y, tw, th, clip;
  float lut_scale, inv_tw, inv_th;
};

UVIO_HD ClaheGeom clahe_geom(int w, int h, double clip_limit, int tiles_x, int tiles_y) {
  ClaheGeom q;
  int wp = w, hp = h;
  if (w % tiles_x != 0 || h % tiles_y != 0) {   // OpenCV pads both directions as soon as either is not divisible
    wp = w + (tiles_x - w % tiles_x);
    hp = h + (tiles_y - h % tiles_y);
  }
  q.tiles_x = tiles_x;
  q.tiles_y = tiles_y;
  q.tw = wp / tiles_x;
  q.th = hp / tiles_y;
  int area = q.tw * q.th;
  q.clip = 0;
  if (clip_limit > 0.0) {
    q.clip = (int)(clip_limit * area / 256);
    if (q.clip < 1) q.clip = 1;
  }
  q.lut_scale = 255.f / (float)area;
  q.inv_tw = 1.0f / (float)q.tw;
  q.inv_th = 1.0f / (float)q.th;
  return q;
}

// histogram of tile (tx,ty) (REFLECT_101 beyond the image when it was padded) into sh[256] (zeroed)
template <class G>
UVIO_HD void clahe_tile_hist(G g, unsigned *sh, const ImgView &I, const ClaheGeom &q, int tx, int ty) {
  int n = q.tw * q.th;
  for (int i = g.lane(); i < n; i += g.size()) {
    int yy = i / q.tw, xx = i - yy * q.tw;
    int y = ty * q.th + yy, x = tx * q.tw + xx;
    if (y >= I.h) y = reflect101(y, I.h);
    if (x >= I.w) x = reflect101(x, I.w);
    smem_inc(&sh[I.p[(size_t)y * I.pitch + x]]);
  }
}

// clip + redistribute + LUT for one tile (one lane)
UVIO_HD void clahe_tile_lut(unsigned *hist, const ClaheGeom &q, uint8_t *lut) {
  if (q.clip > 0) {
    int clipped = 0;
    for (int i = 0; i < 256; i++)
      if ((int)hist[i] > q.clip) {
        clipped += (int)hist[i] - q.clip;
        hist[i] = (unsigned)q.clip;
      }
    int redist = clipped / 256, residual = clipped - redist * 256;
    for (int i = 0; i < 256; i++) hist[i] += (unsigned)redist;
    if (residual != 0) {
      int step = 256 / residual;
      if (step < 1) step = 1;
      for (int i = 0; i < 256 && residual > 0; i += step, residual--) hist[i]++;
    }
  }
  int sum = 0;
  for (int i = 0; i < 256; i++) {
    sum += (int)hist[i];
    lut[i] = sat_u8(round_f2i(fmul((float)sum, q.lut_scale)));
  }
}

// bilinear blend of the four neighbouring tile LUTs; luts = [tiles_y][tiles_x][256]
UVIO_HD uint8_t clahe_pixel(const uint8_t *luts, const ClaheGeom &q, int x, int y, int v) {
  float tyf = fsub(fmul((float)y, q.inv_th), 0.5f);
  int ty1 = floor_f2i(tyf), ty2 = ty1 + 1;
  float ya = fsub(tyf, (float)ty1), ya1 = fsub(1.0f, ya);
  ty1 = clampi(ty1, 0, q.tiles_y - 1);
  ty2 = clampi(ty2, 0, q.tiles_y - 1);
  float txf = fsub(fmul((float)x, q.inv_tw), 0.5f);
  int tx1 = floor_f2i(txf), tx2 = tx1 + 1;
  float xa = fsub(txf, (float)tx1), xa1 = fsub(1.0f, xa);
  tx1 = clampi(tx1, 0, q.tiles_x - 1);
  tx2 = clampi(tx2, 0, q.tiles_x - 1);
  float l11 = (float)luts[(ty1 * q.tiles_x + tx1) * 256 + v], l12 = (float)luts[(ty1 * q.tiles_x + tx2) * 256 + v];
  float l21 = (float)luts[(ty2 * q.tiles_x + tx1) * 256 + v], l22 = (float)luts[(ty2 * q.tiles_x + tx2) * 256 + v];
  float res = fadd(fmul(fadd(fmul(l11, xa1), fmul(l12, xa)), ya1), fmul(fadd(fmul(l21, xa1), fmul(l22, xa)), ya));
  return sat_u8(round_f2i(res));
}

// ---- pyrDown: one block computes a PD_TW x PD_TH tile of the output level --------------------------------------------
constexpr int PD_TW = 64, PD_TH = 8;
constexpr int PD_SW = 2 * PD_TW + 3, PD_SH = 2 * PD_TH + 3;   // staged source footprint 131 x 19
constexpr int PD_SP = PD_SW + 1;                              // staging pitch

struct PyrSmem {
  uint8_t src[PD_SH * PD_SP];
  unsigned short hsum[PD_SH * PD_TW];   // horizontal [1 4 6 4 1] sums (<= 4080)
};

template <class G>
UVIO_HD void pyrdown_tile(G g, PyrSmem &s, const ImgView &I, uint8_t *dst, int dw, int dh, int dpitch, int ox0, int oy0) {
  int sx0 = 2 * ox0 - 2, sy0 = 2 * oy0 - 2;
  for (int i = g.lane(); i < PD_SH * PD_SW; i += g.size()) {
    int yy = i / PD_SW, xx = i - yy * PD_SW;
    int X = reflect101(sx0 + xx, I.w), Y = reflect101(sy0 + yy, I.h);
    s.src[yy * PD_SP + xx] = I.p[(size_t)Y * I.pitch + X];
  }
  g.sync();
  for (int i = g.lane(); i < PD_SH * PD_TW; i += g.size()) {
    int yy = i / PD_TW, ox = i - yy * PD_TW;
    const uint8_t *r = s.src + yy * PD_SP + 2 * ox;
    s.hsum[i] = (unsigned short)(r[0] + 4 * r[1] + 6 * r[2] + 4 * r[3] + r[4]);
  }
  g.sync();
  for (int i = g.lane(); i < PD_TH * PD_TW; i += g.size()) {
    int oy = i / PD_TW, ox = i - oy * PD_TW;
    if (ox0 + ox < dw && oy0 + oy < dh) {
      const unsigned short *c = s.hsum + (2 * oy) * PD_TW + ox;
      int v = c[0] + 4 * c[PD_TW] + 6 * c[2 * PD_TW] + 4 * c[3 * PD_TW] + c[4 * PD_TW];
      dst[(size_t)(oy0 + oy) * dpitch + ox0 + ox] = (uint8_t)((v + 128) >> 8);
    }
  }
  g.sync();
}

// Last pyramid level cv::buildOpticalFlowPyramid creates: it stops before a level that is <= the LK window.
UVIO_HD int pyr_max_level(int w, int h, int requested, int win) {
  int level = 0;
  while (level < requested) {
    int nw = (w + 1) / 2, nh = (h + 1) / 2;
    if (nw <= win || nh <= win) break;
    w = nw;
    h = nh;
    level++;
  }
  return level;
}

}  // namespace uvio
