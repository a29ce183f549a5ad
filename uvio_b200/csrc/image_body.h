// Image-wide integer passes: histogram equalisation, CLAHE, pyrDown.  
//
//   cv::equalizeHist                         ov_core/src/track/TrackKLT.cpp:59
//   cv::createCLAHE(10.0, Size(8,8))->apply  ov_core/src/track/TrackKLT.cpp:61-64
//   cv::buildOpticalFlowPyramid (pyrDown)    ov_core/src/track/TrackKLT.cpp:71
//
// All three are HBM-bound byte passes; bodies are written per thread block (GroupBlock on the device).
#pragma once
#include <string.h>

#include "simt.h"

namespace uvio {

#ifdef __CUDA_ARCH__
__device__ __forceinline__ void smem_inc(unsigned *p) { atomicAdd(p, 1u); }
__device__ __forceinline__ unsigned gmem_add(unsigned *p, unsigned v) { return atomicAdd(p, v); }
__device__ __forceinline__ unsigned smem_add(unsigned *p, unsigned v) { return atomicAdd(p, v); }
#else
inline unsigned smem_add(unsigned *p, unsigned v) { unsigned o = *p; *p += v; return o; }
inline void smem_inc(unsigned *p) { ++*p; }
inline unsigned gmem_add(unsigned *p, unsigned v) { unsigned o = *p; *p += v; return o; }
#endif

// ---- histogram of rows [row0,row1) of one image into a block-private histogram ---------------------------------------
// sh holds `nsub` (a power of two) histograms of 256 bins; warp k of the group counts into copy k % nsub, which keeps the
// shared-memory atomics of different warps apart.  Rows are read as 16-byte vectors when the layout allows it.
struct Vec16 { uint32_t x, y, z, w; };

UVIO_HD void hist_word(unsigned *sh, uint32_t v) {
  smem_inc(&sh[v & 255]);
  smem_inc(&sh[(v >> 8) & 255]);
  smem_inc(&sh[(v >> 16) & 255]);
  smem_inc(&sh[v >> 24]);
}

#ifdef __CUDA_ARCH__
__device__ __forceinline__ Vec16 load_vec16(const uint8_t *p) {
  uint4 v = *(const uint4 *)p;
  return Vec16{v.x, v.y, v.z, v.w};
}
__device__ __forceinline__ void store_vec16(uint8_t *p, const Vec16 &v) { *(uint4 *)p = make_uint4(v.x, v.y, v.z, v.w); }
#else
inline Vec16 load_vec16(const uint8_t *p) { Vec16 v; memcpy(&v, p, 16); return v; }
inline void store_vec16(uint8_t *p, const Vec16 &v) { memcpy(p, &v, 16); }
#endif

template <class G>
UVIO_HD void hist_rows(G g, unsigned *sh /*nsub*256, zeroed*/, const ImgView &I, int row0, int row1, int nsub = 1) {
  unsigned *mine = sh + ((g.lane() >> 5) & (nsub - 1)) * 256;
  const bool vec = ((I.pitch & 15) == 0) && ((((size_t)I.p) & 15) == 0);
  const int w16 = vec ? (I.w >> 4) : 0, rows = row1 - row0;
  for (int i = g.lane(); i < rows * w16; i += g.size()) {
    int y = i / w16, x = i - y * w16;
    Vec16 v = load_vec16(I.p + (size_t)(row0 + y) * I.pitch + 16 * x);
    hist_word(mine, v.x);
    hist_word(mine, v.y);
    hist_word(mine, v.z);
    hist_word(mine, v.w);
  }
  const int x0 = w16 << 4, rem = I.w - x0;   // columns the vectors did not cover
  for (int i = g.lane(); i < rows * rem; i += g.size()) {
    int y = i / rem, x = i - y * rem;
    smem_inc(&mine[I.p[(size_t)(row0 + y) * I.pitch + x0 + x]]);
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
UVIO_HD uint32_t lut_word(const uint8_t *lut, uint32_t v) {
  return (uint32_t)lut[v & 255] | ((uint32_t)lut[(v >> 8) & 255] << 8) | ((uint32_t)lut[(v >> 16) & 255] << 16) |
         ((uint32_t)lut[v >> 24] << 24);
}

template <class G>
UVIO_HD void lut_rows(G g, const uint8_t *lut /*shared*/, const ImgView &I, uint8_t *dst, int dpitch, int row0, int row1) {
  const bool vec = ((I.pitch & 15) == 0) && ((((size_t)I.p) & 15) == 0) && ((dpitch & 15) == 0) && ((((size_t)dst) & 15) == 0);
  const int w16 = vec ? (I.w >> 4) : 0, rows = row1 - row0;
  for (int i = g.lane(); i < rows * w16; i += g.size()) {
    int y = i / w16, x = i - y * w16;
    Vec16 v = load_vec16(I.p + (size_t)(row0 + y) * I.pitch + 16 * x);
    v.x = lut_word(lut, v.x);
    v.y = lut_word(lut, v.y);
    v.z = lut_word(lut, v.z);
    v.w = lut_word(lut, v.w);
    store_vec16(dst + (size_t)(row0 + y) * dpitch + 16 * x, v);
  }
  const int x0 = w16 << 4, rem = I.w - x0;
  for (int i = g.lane(); i < rows * rem; i += g.size()) {
    int y = i / rem, x = i - y * rem;
    dst[(size_t)(row0 + y) * dpitch + x0 + x] = lut[I.p[(size_t)(row0 + y) * I.pitch + x0 + x]];
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

// ---- pyrDown: one lane computes a PD_COLS x PD_ROWS strip of the output level ----------------------------------------
constexpr int PD_COLS = 4;    // output columns per lane (one 32-bit store)
constexpr int PD_ROWS = 4;    // output rows per lane: 2*PD_ROWS+3 source rows, all fetched before the arithmetic starts

// one aligned source word, REFLECT_101 at the frame border
UVIO_HD uint32_t pyr_src_word(const uint8_t *row, int X0, int w, bool aligned) {
  if (aligned && X0 >= 0 && X0 + 3 < w) return *(const uint32_t *)(row + X0);
  return (uint32_t)row[reflect101(X0, w)] | ((uint32_t)row[reflect101(X0 + 1, w)] << 8) |
         ((uint32_t)row[reflect101(X0 + 2, w)] << 16) | ((uint32_t)row[reflect101(X0 + 3, w)] << 24);
}

// horizontal [1 4 6 4 1] sums for output columns ox0..ox0+3 from the four words around them, packed as 16-bit pairs
// (each <= 4080).  b(k) = source pixel 2*ox0 + k, k = -2 .. 8.
UVIO_HD void pyr_hsum(uint32_t wm, uint32_t w0, uint32_t w1, uint32_t wp, uint32_t &h01, uint32_t &h23) {
  uint32_t bm2 = (wm >> 16) & 255, bm1 = wm >> 24;
  uint32_t b0 = w0 & 255, b1 = (w0 >> 8) & 255, b2 = (w0 >> 16) & 255, b3 = w0 >> 24;
  uint32_t b4 = w1 & 255, b5 = (w1 >> 8) & 255, b6 = (w1 >> 16) & 255, b7 = w1 >> 24;
  uint32_t b8 = wp & 255;
  uint32_t h0 = bm2 + b2 + 4 * (bm1 + b1) + 6 * b0;
  uint32_t h1 = b0 + b4 + 4 * (b1 + b3) + 6 * b2;
  uint32_t h2 = b2 + b6 + 4 * (b3 + b5) + 6 * b4;
  uint32_t h3 = b4 + b8 + 4 * (b5 + b7) + 6 * b6;
  h01 = h0 | (h1 << 16);
  h23 = h2 | (h3 << 16);
}

// vertical [1 4 6 4 1] on packed pairs: each 16-bit lane <= 16*4080 = 65280, no carry between the lanes
UVIO_HD uint32_t pyr_vsum(uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t a4) {
  return a0 + a4 + 4 * (a1 + a3) + 6 * a2;
}

// One lane's share of cv::pyrDown (App. A.3): output columns ox0..ox0+3 (ox0 a multiple of 4), rows oy0..oy0+PD_ROWS-1.
// Every source byte is fetched as part of an aligned word; there is no shared-memory staging and no barrier.  INTERIOR:
// the lane's source columns 2*ox0 .. 2*ox0+11 lie inside the frame and rows are word-aligned, so the loads carry no
// branches and the compiler can issue all of them up front (the two reflected pixels left of column 0 are taken from
// the first word in registers).  Only the lanes at the right border of the frame take the general path.
template <bool INTERIOR, bool ONCE>
UVIO_HD void pyrdown_strip_t(const ImgView &I, uint8_t *UVIO_RESTRICT dst, int dw, int dh, int dpitch, int ox0, int oy0) {
  constexpr int NR = 2 * PD_ROWS + 3;
  const int sx = 2 * ox0, sy = 2 * oy0 - 2;
  uint32_t a[NR], c[NR];   // columns (0,1) in a, (2,3) in c
#pragma unroll
  for (int r = 0; r < NR; r++) {
    const uint8_t *UVIO_RESTRICT row = I.p + (size_t)((INTERIOR || ONCE) ? reflect101_once(sy + r, I.h) : reflect101(sy + r, I.h)) * I.pitch;
    uint32_t wm, w0, w1, wp;
    if (INTERIOR) {
      const uint32_t *q = (const uint32_t *)(row + sx);
      w0 = q[0]; w1 = q[1]; wp = q[2];
      wm = sx ? q[-1] : ((w0 & 0x00ff0000u) | ((w0 & 0x0000ff00u) << 16));   // pixels -2,-1 mirror pixels 2,1
    } else if (ONCE) {
      // right border (or an unaligned image): the 11 source pixels one by one, reflected without branches
      uint32_t px[11];
#pragma unroll
      for (int k = 0; k < 11; k++) px[k] = row[reflect101_once(sx - 2 + k, I.w)];
      wm = (px[0] << 16) | (px[1] << 24);
      w0 = px[2] | (px[3] << 8) | (px[4] << 16) | (px[5] << 24);
      w1 = px[6] | (px[7] << 8) | (px[8] << 16) | (px[9] << 24);
      wp = px[10];
    } else {
      const bool al = ((I.pitch & 3) == 0) && ((((size_t)I.p) & 3) == 0);
      wm = pyr_src_word(row, sx - 4, I.w, al); w0 = pyr_src_word(row, sx, I.w, al);
      w1 = pyr_src_word(row, sx + 4, I.w, al); wp = pyr_src_word(row, sx + 8, I.w, al);
    }
    pyr_hsum(wm, w0, w1, wp, a[r], c[r]);
  }
  const bool dst_word = ((dpitch & 3) == 0) && ((((size_t)dst) & 3) == 0) && ox0 + 3 < dw;
#pragma unroll
  for (int o = 0; o < PD_ROWS; o++) {
    if (oy0 + o >= dh) break;
    uint32_t v01 = pyr_vsum(a[2 * o], a[2 * o + 1], a[2 * o + 2], a[2 * o + 3], a[2 * o + 4]) + 0x00800080u;   // +128 per lane
    uint32_t v23 = pyr_vsum(c[2 * o], c[2 * o + 1], c[2 * o + 2], c[2 * o + 3], c[2 * o + 4]) + 0x00800080u;
    uint32_t out = ((v01 >> 8) & 255) | ((v01 >> 24) << 8) | (((v23 >> 8) & 255) << 16) | ((v23 >> 24) << 24);
    uint8_t *d = dst + (size_t)(oy0 + o) * dpitch + ox0;
    if (dst_word) *(uint32_t *)d = out;
    else
      for (int k = 0; k < 4 && ox0 + k < dw; k++) d[k] = (uint8_t)(out >> (8 * k));
  }
}

UVIO_HD void pyrdown_strip(const ImgView &I, uint8_t *dst, int dw, int dh, int dpitch, int ox0, int oy0) {
  if (ox0 >= dw || oy0 >= dh) return;
  const bool aligned = ((I.pitch & 3) == 0) && ((((size_t)I.p) & 3) == 0);
  // a strip reads source rows 2*oy0-2 .. 2*oy0+2*PD_ROWS and columns 2*ox0-2 .. 2*ox0+8 whether or not all of its outputs exist:
  // one reflection covers them when the image is at least 2*PD_ROWS+8 = 16 high / wide (every pyramid source level is)
  constexpr int MIN_ONCE = 2 * PD_ROWS + 8;
  if (aligned && 2 * ox0 + 12 <= I.w && I.h >= MIN_ONCE) pyrdown_strip_t<true, false>(I, dst, dw, dh, dpitch, ox0, oy0);
  else if (I.w >= MIN_ONCE && I.h >= MIN_ONCE) pyrdown_strip_t<false, true>(I, dst, dw, dh, dpitch, ox0, oy0);
  else pyrdown_strip_t<false, false>(I, dst, dw, dh, dpitch, ox0, oy0);
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
