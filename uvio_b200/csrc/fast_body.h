// FAST-9/16 with 3x3 non-max suppression on a sub-rectangle (grid cell) of an image.
//
// Replaces cv::FAST(img(cv::Rect(x, y, size_x, size_y)), pts, threshold, true) at
// ov_core/src/track/Grider_GRID.h:125.  ROI semantics are kept: positions within 3 px of the
// *cell* edge never fire and count as score 0 for their neighbours' suppression test.
// One thread block handles a FT_W x FT_H tile of the cell; corners are appended (unordered) to the job's list as
// packed words  x | y << 12 | score << 24  (cell-relative x, y).  The host restores cv::FAST's row-major emission
// order by sorting on (y, x) before it applies the reference's std::sort-by-response (Grider_GRID.h:128).
#pragma once
#include "image_body.h"
#include "simt.h"
#ifndef __CUDACC__
#include <algorithm>
using std::max;
using std::min;
#endif

namespace uvio {

constexpr int FT_W = 64, FT_H = 16;
constexpr int FT_PW = FT_W + 8, FT_PH = FT_H + 8;     // pixel footprint (tile + 1 nms halo + 3 ring halo each side)
constexpr int FT_SW = FT_W + 2, FT_SH = FT_H + 2;     // score footprint (tile + 1 nms halo each side)

struct FastSmem {
  uint8_t pix[FT_PH * FT_PW];
  uint8_t score[FT_SH * FT_SW];
  uint16_t cand[FT_SH * FT_SW];   // score positions that pass the four-point pre-test, in no particular order
  unsigned ncand;
};

struct FastJob {          // one grid cell of one image
  int img;                // index into the image table of the launch
  int x0, y0, w, h;       // cell rectangle (Grider_GRID.h:113-114, integer-division cell size)
  int out_off;            // first word of this job's list in the output buffer
  int out_cap;            // capacity in words (sized to the NMS worst case, ceil(w/2)*ceil(h/2): cannot overflow)
  int threshold;          // cv::FAST threshold of this cell's tracker
};

UVIO_HD int fast_ring_off(int k, int pitch) {
  // (dx,dy), k = 0..15: (0,3)(1,3)(2,2)(3,1)(3,0)(3,-1)(2,-2)(1,-3)(0,-3)(-1,-3)(-2,-2)(-3,-1)(-3,0)(-3,1)(-2,2)(-1,3)
  const int dx = k < 4 ? k : (k < 6 ? 3 : (k < 12 ? 8 - k : (k < 14 ? -3 : k - 16)));
  const int dy = k < 2 ? 3 : (k < 7 ? 4 - k : (k < 10 ? -3 : (k < 15 ? k - 12 : 3)));
  return dy * pitch + dx;
}

// Necessary condition for a 9-of-16 arc: any nine contiguous ring positions hold at least two of the four compass points
// (k = 0, 4, 8, 12), so a corner has two of them darker than centre - t or two brighter than centre + t.  Five loads decide
// for most pixels of a natural image that the sixteen-point test need not run.
UVIO_HD bool fast_maybe(const uint8_t *p, int pitch, int threshold) {
  const int v = p[0];
  const int d0 = v - (int)p[fast_ring_off(0, pitch)], d4 = v - (int)p[fast_ring_off(4, pitch)], d8 = v - (int)p[fast_ring_off(8, pitch)],
            d12 = v - (int)p[fast_ring_off(12, pitch)];
  const int up = (d0 > threshold) + (d4 > threshold) + (d8 > threshold) + (d12 > threshold);
  const int dn = (d0 < -threshold) + (d4 < -threshold) + (d8 < -threshold) + (d12 < -threshold);
  return up >= 2 || dn >= 2;
}

// largest t' such that the pixel is still a corner at threshold t' (0 if it is not a corner at `threshold`)
UVIO_HD int fast_score(const uint8_t *p, int pitch, int threshold) {
  int d[16];
  int v = p[0];
  unsigned brighter = 0, darker = 0;   // ring darker than centre by > t  /  ring brighter than centre by > t
#pragma unroll
  for (int k = 0; k < 16; k++) {
    d[k] = v - (int)p[fast_ring_off(k, pitch)];
    brighter |= (unsigned)(d[k] > threshold) << k;
    darker |= (unsigned)(d[k] < -threshold) << k;
  }
  unsigned m1 = brighter | (brighter << 16), m2 = darker | (darker << 16);
  unsigned r1 = m1 & (m1 >> 1), r2 = m2 & (m2 >> 1);
  r1 &= r1 >> 2; r2 &= r2 >> 2;
  r1 &= r1 >> 4; r2 &= r2 >> 4;
  r1 &= m1 >> 8; r2 &= m2 >> 8;
  if (!((r1 | r2) & 0xFFFFu)) return 0;
  // Arc minima by doubling (m2 over d[k..k+1], m4 over 4, m8 over 8, then the ninth element), once for d = centre - ring
  // and once for e = ring - centre. Written without negating a min/max result: ptxas 12.9 for sm_100a drops the negation
  // when it fuses max(x, -max(..)) into VIMNMX3 (seen on B200: scores came out as max(d) instead of the arc minimum).
  int e[16];
#pragma unroll
  for (int k = 0; k < 16; k++) e[k] = (int)p[fast_ring_off(k, pitch)] - v;
  int best = -1000;
  {
    int m2[16], m4[16];
#pragma unroll
    for (int k = 0; k < 16; k++) m2[k] = min(d[k], d[(k + 1) & 15]);
#pragma unroll
    for (int k = 0; k < 16; k++) m4[k] = min(m2[k], m2[(k + 2) & 15]);
#pragma unroll
    for (int k = 0; k < 16; k++) best = max(best, min(min(m4[k], m4[(k + 4) & 15]), d[(k + 8) & 15]));
  }
  {
    int m2[16], m4[16];
#pragma unroll
    for (int k = 0; k < 16; k++) m2[k] = min(e[k], e[(k + 1) & 15]);
#pragma unroll
    for (int k = 0; k < 16; k++) m4[k] = min(m2[k], m2[(k + 2) & 15]);
#pragma unroll
    for (int k = 0; k < 16; k++) best = max(best, min(min(m4[k], m4[(k + 4) & 15]), e[(k + 8) & 15]));
  }
  return best > threshold ? best - 1 : 0;
}

template <class G>
UVIO_HD void fast_tile(G g, FastSmem &s, const ImgView &I, const FastJob &job, int tx0, int ty0, int threshold,
                       uint32_t *out, unsigned *count) {
  // stage pixels: cell-relative (tx0-4 .. tx0+FT_W+3, ty0-4 .. ty0+FT_H+3), clamped into the cell (clamped values are
  // only ever read for positions whose score is forced to 0)
  for (int i = g.lane(); i < FT_PH * FT_PW; i += g.size()) {
    int yy = i / FT_PW, xx = i - yy * FT_PW;
    int x = clampi(tx0 - 4 + xx, 0, job.w - 1), y = clampi(ty0 - 4 + yy, 0, job.h - 1);
    s.pix[i] = I.p[(size_t)(job.y0 + y) * I.pitch + job.x0 + x];
  }
  if (g.lane() == 0) s.ncand = 0;
  g.sync();
  // pass 1: the cheap pre-test on every score position; survivors are listed so that pass 2 runs on full warps (a warp in
  // which one lane of 32 needs the sixteen-point test would otherwise pay for it in all of them)
  for (int i = g.lane(); i < FT_SH * FT_SW; i += g.size()) {
    int yy = i / FT_SW, xx = i - yy * FT_SW;
    int x = tx0 - 1 + xx, y = ty0 - 1 + yy;
    s.score[i] = 0;
    if (x >= 3 && x < job.w - 3 && y >= 3 && y < job.h - 3 && fast_maybe(s.pix + (yy + 3) * FT_PW + (xx + 3), FT_PW, threshold))
      s.cand[smem_add(&s.ncand, 1u)] = (uint16_t)i;
  }
  g.sync();
  const int ncand = (int)s.ncand;
  for (int c = g.lane(); c < ncand; c += g.size()) {
    const int i = s.cand[c];
    int yy = i / FT_SW, xx = i - yy * FT_SW;
    s.score[i] = (uint8_t)fast_score(s.pix + (yy + 3) * FT_PW + (xx + 3), FT_PW, threshold);
  }
  g.sync();
  for (int i = g.lane(); i < FT_H * FT_W; i += g.size()) {
    int yy = i / FT_W, xx = i - yy * FT_W;
    int x = tx0 + xx, y = ty0 + yy;
    if (x >= job.w || y >= job.h) continue;
    const uint8_t *c = s.score + (yy + 1) * FT_SW + (xx + 1);
    int sc = c[0];
    if (!sc) continue;
    if (sc > c[-1] && sc > c[1] && sc > c[-FT_SW - 1] && sc > c[-FT_SW] && sc > c[-FT_SW + 1] && sc > c[FT_SW - 1] &&
        sc > c[FT_SW] && sc > c[FT_SW + 1]) {
      unsigned idx = gmem_add(count, 1u);
      if ((int)idx < job.out_cap) out[job.out_off + idx] = (uint32_t)x | ((uint32_t)y << 12) | ((uint32_t)sc << 24);
    }
  }
  g.sync();
}

}  // namespace uvio
