// Pyramidal Lucas-Kanade, one execution group (a warp on the device) per feature.
//
// Replaces cv::calcOpticalFlowPyrLK(prevPyr, nextPyr, prevPts, nextPts, status, err, Size(15,15), maxLevel,
// {COUNT|EPS, 30, 0.01}, OPTFLOW_USE_INITIAL_FLOW) at ov_core/src/track/TrackKLT.cpp:671 and :858.
//
// Arithmetic: 14-bit integer bilinear weights, template I = intensity*32 (s16), Scharr derivative
// patch (s16) blended from the level's derivative image (zero outside the frame), 2x2 Gram matrix and mismatch vector
// accumulated in fp32.  No derivative image is materialised: the Scharr values are computed from an 18x18 intensity
// footprint staged in shared memory.  The fp32 sums are accumulated in the order OpenCV's baseline (SSE2) build uses
// (per window row: columns 0..7 in four lanes, columns 8..14 in one running scalar), so positions come out identical
// to the CPU tracker rather than merely within tolerance: 15 (Gram) / 10 (mismatch) independent ordered chains are
// spread over the lanes of the group; everything else is lane-parallel integer work.
#pragma once
#include <float.h>

#include "simt.h"

namespace uvio {

constexpr int LK_WIN = 15;                 // TrackKLT.h:144 (win_size 15x15, not configurable in the reference)
constexpr int LK_NPIX = LK_WIN * LK_WIN;   // 225
constexpr int LK_FP = 18;                  // template footprint rows/cols (16 + 1-px Scharr halo each side)
constexpr int LK_FPP = 20;                 // footprint pitch
constexpr int LK_JP = 16;                  // J footprint rows/cols

struct LkSmem {
  uint8_t foot[LK_FP * LK_FPP];            // I(ip-1 .. ip+16), REFLECT_101
  uint8_t jfoot[LK_JP * LK_JP];            // J(in .. in+15), REFLECT_101
  short dI[16 * 16 * 2];                   // Scharr (dx,dy) at ip .. ip+15, zero outside the frame
  short Iw[LK_NPIX];                       // template, intensity * 32
  short dIw[LK_NPIX * 2];                  // blended derivative patch
  short diff[LK_NPIX];                     // J patch - I patch of the current iteration
  float red[16];                           // results of the ordered accumulation chains
};

struct LkParams {
  int max_level;      // effective top level (min(requested, levels built))
  int max_count;      // 30
  double eps_sq;      // 0.01^2
  double min_eig;     // 1e-4
};

template <class G>
UVIO_HD void lk_stage_template(G g, LkSmem &s, const ImgView &I, int ipx, int ipy, int iw00, int iw01, int iw10, int iw11) {
  for (int i = g.lane(); i < LK_FP * LK_FP; i += g.size()) {
    int yy = i / LK_FP, xx = i - yy * LK_FP;
    int X = reflect101(ipx - 1 + xx, I.w), Y = reflect101(ipy - 1 + yy, I.h);
    s.foot[yy * LK_FPP + xx] = I.p[(size_t)Y * I.pitch + X];
  }
  g.sync();
  for (int i = g.lane(); i < 256; i += g.size()) {
    int y = i >> 4, x = i & 15;
    int X = ipx + x, Y = ipy + y;
    int dx = 0, dy = 0;
    if (X >= 0 && X < I.w && Y >= 0 && Y < I.h) {
      const uint8_t *c = s.foot + (y + 1) * LK_FPP + (x + 1);
      int tl = c[-LK_FPP - 1], tc = c[-LK_FPP], tr = c[-LK_FPP + 1], ml = c[-1], mr = c[1];
      int bl = c[LK_FPP - 1], bc = c[LK_FPP], br = c[LK_FPP + 1];
      dx = 3 * (tr - tl) + 10 * (mr - ml) + 3 * (br - bl);
      dy = 3 * (bl - tl) + 10 * (bc - tc) + 3 * (br - tr);
    }
    s.dI[2 * i] = (short)dx;
    s.dI[2 * i + 1] = (short)dy;
  }
  g.sync();
  for (int i = g.lane(); i < LK_NPIX; i += g.size()) {
    int y = i / LK_WIN, x = i - y * LK_WIN;
    const uint8_t *c = s.foot + (y + 1) * LK_FPP + (x + 1);
    int ival = (c[0] * iw00 + c[1] * iw01 + c[LK_FPP] * iw10 + c[LK_FPP + 1] * iw11 + (1 << 8)) >> 9;
    const short *d = s.dI + 2 * (y * 16 + x);
    int ix = (d[0] * iw00 + d[2] * iw01 + d[32] * iw10 + d[34] * iw11 + (1 << 13)) >> 14;
    int iy = (d[1] * iw00 + d[3] * iw01 + d[33] * iw10 + d[35] * iw11 + (1 << 13)) >> 14;
    s.Iw[i] = (short)ival;
    s.dIw[2 * i] = (short)ix;
    s.dIw[2 * i + 1] = (short)iy;
  }
  g.sync();
}

// 15 ordered chains -> red[0..14]; chain = acc*5 + k, acc in {A11, A12, A22}, k = 0..3 SIMD lane, 4 = scalar tail.
template <class G>
UVIO_HD void lk_accumulate_gram(G g, LkSmem &s) {
  for (int c = g.lane(); c < 15; c += g.size()) {
    int acc = c / 5, k = c - acc * 5;
    int oa = acc == 2 ? 1 : 0, ob = acc == 0 ? 0 : 1;     // A11: ix*ix, A12: ix*iy, A22: iy*iy
    float q = 0.f;
    if (k < 4) {
      for (int y = 0; y < LK_WIN; y++) {
        const short *r = s.dIw + 2 * (y * LK_WIN);
        q = fadd(fmul((float)r[2 * k + oa], (float)r[2 * k + ob]), q);
        q = fadd(fmul((float)r[2 * (k + 4) + oa], (float)r[2 * (k + 4) + ob]), q);
      }
    } else {
      for (int y = 0; y < LK_WIN; y++) {
        const short *r = s.dIw + 2 * (y * LK_WIN);
        for (int x = 8; x < LK_WIN; x++) q = fadd(q, (float)((int)r[2 * x + oa] * (int)r[2 * x + ob]));
      }
    }
    s.red[c] = q;
  }
  g.sync();
}

// 10 ordered chains -> red[0..9]; chain = acc*5 + k, acc in {b1, b2}.
template <class G>
UVIO_HD void lk_accumulate_mismatch(G g, LkSmem &s) {
  for (int c = g.lane(); c < 10; c += g.size()) {
    int acc = c / 5, k = c - acc * 5;
    float q = 0.f;
    if (k < 4) {
      for (int y = 0; y < LK_WIN; y++) {
        const short *d = s.diff + y * LK_WIN;
        const short *r = s.dIw + 2 * (y * LK_WIN) + acc;
        int pair = (int)d[k] * (int)r[2 * k] + (int)d[k + 4] * (int)r[2 * (k + 4)];
        q = fadd(q, (float)pair);
      }
    } else {
      for (int y = 0; y < LK_WIN; y++) {
        const short *d = s.diff + y * LK_WIN;
        const short *r = s.dIw + 2 * (y * LK_WIN) + acc;
        for (int x = 8; x < LK_WIN; x++) q = fadd(q, (float)((int)d[x] * (int)r[2 * x]));
      }
    }
    s.red[c] = q;
  }
  g.sync();
}

UVIO_HD float lk_fold(const float *r) {   // scalar tail += (l0+l2)+(l1+l3)
  return fadd(r[4], fadd(fadd(r[0], r[2]), fadd(r[1], r[3])));
}

template <class G>
UVIO_HD void lk_stage_diff(G g, LkSmem &s, const ImgView &J, int inx, int iny, int iw00, int iw01, int iw10, int iw11) {
  for (int i = g.lane(); i < LK_JP * LK_JP; i += g.size()) {
    int yy = i >> 4, xx = i & 15;
    int X = reflect101(inx + xx, J.w), Y = reflect101(iny + yy, J.h);
    s.jfoot[i] = J.p[(size_t)Y * J.pitch + X];
  }
  g.sync();
  for (int i = g.lane(); i < LK_NPIX; i += g.size()) {
    int y = i / LK_WIN, x = i - y * LK_WIN;
    const uint8_t *c = s.jfoot + y * LK_JP + x;
    int jv = (c[0] * iw00 + c[1] * iw01 + c[LK_JP] * iw10 + c[LK_JP + 1] * iw11 + (1 << 8)) >> 9;
    s.diff[i] = (short)(jv - s.Iw[i]);
  }
  g.sync();
}

UVIO_HD void lk_weights(float a, float b, int &iw00, int &iw01, int &iw10, int &iw11) {
  float a1 = fsub(1.f, a), b1 = fsub(1.f, b);
  iw00 = round_f2i(fmul(fmul(a1, b1), 16384.f));
  iw01 = round_f2i(fmul(fmul(a, b1), 16384.f));
  iw10 = round_f2i(fmul(fmul(a1, b), 16384.f));
  iw11 = 16384 - iw00 - iw01 - iw10;
}

// Tracks one point through all levels. Every lane holds the same scalars (they are recomputed redundantly).
// prev/next: arrays of max_level+1 level views. (nx, ny) holds the initial guess on entry and the result on return.
template <class G>
UVIO_HD void lk_track_point(G g, LkSmem &s, const ImgView *prev, const ImgView *next, const LkParams &P, float px,
                            float py, float &nx, float &ny, uint8_t &status, int *work = nullptr) {
  const float FLT_SCALE = 1.f / (1 << 20);
  const float half = (LK_WIN - 1) * 0.5f;
  status = 1;
  float outx = nx, outy = ny;
  for (int level = P.max_level; level >= 0; level--) {
    const ImgView I = prev[level], J = next[level];
    float sc = (float)(1. / (1 << level));
    float ppx = fmul(px, sc), ppy = fmul(py, sc);
    float npx, npy;
    if (level == P.max_level) {
      npx = fmul(outx, sc);
      npy = fmul(outy, sc);
    } else {
      npx = fmul(outx, 2.f);
      npy = fmul(outy, 2.f);
    }
    outx = npx;
    outy = npy;
    ppx = fsub(ppx, half);
    ppy = fsub(ppy, half);
    int ipx = floor_f2i(ppx), ipy = floor_f2i(ppy);
    if (ipx < -LK_WIN || ipx >= I.w || ipy < -LK_WIN || ipy >= I.h) {
      if (level == 0) status = 0;
      continue;
    }
    int iw00, iw01, iw10, iw11;
    lk_weights(fsub(ppx, (float)ipx), fsub(ppy, (float)ipy), iw00, iw01, iw10, iw11);
    lk_stage_template(g, s, I, ipx, ipy, iw00, iw01, iw10, iw11);
    if (work) work[0]++;   // levels whose template was staged
    lk_accumulate_gram(g, s);
    float A11 = fmul(lk_fold(s.red + 0), FLT_SCALE);
    float A12 = fmul(lk_fold(s.red + 5), FLT_SCALE);
    float A22 = fmul(lk_fold(s.red + 10), FLT_SCALE);
    g.sync();   // red[] is reused by the mismatch chains
    float D = fsub(fmul(A11, A22), fmul(A12, A12));
    float dA = fsub(A11, A22);
    float minEig = fsub(fadd(A22, A11), sqrtf(fadd(fmul(dA, dA), fmul(fmul(4.f, A12), A12)))) / (float)(2 * LK_WIN * LK_WIN);
    if ((double)minEig < P.min_eig || D < FLT_EPSILON) {
      if (level == 0) status = 0;
      continue;
    }
    D = 1.f / D;
    npx = fsub(npx, half);
    npy = fsub(npy, half);
    float pdx = 0.f, pdy = 0.f;
    for (int j = 0; j < P.max_count; j++) {
      int inx = floor_f2i(npx), iny = floor_f2i(npy);
      if (inx < -LK_WIN || inx >= J.w || iny < -LK_WIN || iny >= J.h) {
        if (level == 0) status = 0;
        break;
      }
      lk_weights(fsub(npx, (float)inx), fsub(npy, (float)iny), iw00, iw01, iw10, iw11);
      lk_stage_diff(g, s, J, inx, iny, iw00, iw01, iw10, iw11);
      if (work) work[1]++;   // iterations (J footprints staged)
      lk_accumulate_mismatch(g, s);
      float b1 = fmul(lk_fold(s.red + 0), FLT_SCALE);
      float b2 = fmul(lk_fold(s.red + 5), FLT_SCALE);
      g.sync();
      float dx = fmul(fsub(fmul(A12, b2), fmul(A22, b1)), D);
      float dy = fmul(fsub(fmul(A12, b1), fmul(A11, b2)), D);
      npx = fadd(npx, dx);
      npy = fadd(npy, dy);
      outx = fadd(npx, half);
      outy = fadd(npy, half);
      if (dadd(dmul((double)dx, (double)dx), dmul((double)dy, (double)dy)) <= P.eps_sq) break;
      if (j > 0 && fabs((double)fadd(dx, pdx)) < 0.01 && fabs((double)fadd(dy, pdy)) < 0.01) {
        outx = fsub(outx, fmul(dx, 0.5f));
        outy = fsub(outy, fmul(dy, 0.5f));
        break;
      }
      pdx = dx;
      pdy = dy;
    }
    if (status && level == 0) {
      // the reference passes an err vector (TrackKLT.cpp:669-671), so OpenCV re-checks the final position
      int ix = floor_f2i(fsub(outx, half)), iy = floor_f2i(fsub(outy, half));
      if (ix < -LK_WIN || ix >= J.w || iy < -LK_WIN || iy >= J.h) status = 0;
    }
  }
  nx = outx;
  ny = outy;
}

}  // namespace uvio
