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
constexpr int LK_FPW = 6;                  // 32-bit words staged per footprint row (24 bytes cover 18 + alignment slack)
constexpr int LK_BW = 17;                  // blended-intensity grid of an interior template: 15 + a 1-px Scharr ring
constexpr int LK_JW = 24;                  // rows/cols of the J window staged per level (search range +-4 px around the guess; re-centred beyond)
constexpr int LK_JWW = 7;                  // words per staged J row (28 bytes cover 24 + alignment slack)

constexpr int LK_TAIL = 7 * LK_WIN;         // 105 window pixels in columns 8..14: the scalar tail of every row
constexpr int LK_TAILP = 108;               // padded to whole 16-byte vectors

struct Short2 { short x, y; };
struct Float4 { float x, y, z, w; };

// Shared memory of one feature.  Arrays of the ordered sums are laid out the way the chains walk them: columns 0..7 of
// every row ("simd", 8 or 4 entries per row) and columns 8..14 ("tail", 7 per row, contiguous so that the one lane that
// owns the scalar chain reads four values per load).
struct alignas(16) LkSmem {
  uint32_t footw[LK_FP * LK_FPW];          // I rows (ip-1 .. ip+16), REFLECT_101, as aligned words
  union {
    uint32_t jwin[LK_JW * LK_JWW];         // J window rows, REFLECT_101, as aligned words (after the template is built)
    Short2 dI[16 * 16];                    // Scharr (dx,dy) at ip .. ip+15, zero outside the frame (while it is built)
    int blend[LK_BW * LK_BW];              // blended intensities (weights not yet shifted out) at ip-1 .. ip+15 (interior windows)
  };
  alignas(16) Short2 dIw[LK_NPIX + 3];     // blended derivative patch (aligned: the sizes above are not multiples of 16)
  short Iw[LK_NPIX + 7];                   // template, intensity * 32
  union alignas(16) {                      // read with 16-byte loads (lk_tail_chain)
    struct {
      float gsimd[3][LK_WIN * 8];          // fp32(Ix*Ix), fp32(Ix*Iy), fp32(Iy*Iy), columns 0..7
      float gtail[3][LK_TAILP];            // the same, columns 8..14
    };
    struct {
      float psimd[2][LK_WIN * 4];          // per iteration: fp32(diff*Ix (col k) + diff*Ix (col k+4)), k = 0..3; same for Iy
      float ptail[2][LK_TAILP];            // fp32(diff*Ix), fp32(diff*Iy), columns 8..14
    };
  };
  float red[16];                           // results of the ordered accumulation chains
};

struct LkParams {
  int max_level;      // effective top level (min(requested, levels built))
  int max_count;      // 30
  double eps_sq;      // 0.01^2
  double min_eig;     // 1e-4
};

#ifdef __CUDA_ARCH__
__device__ __forceinline__ Float4 load_float4(const float *p) {
  float4 v = *(const float4 *)p;
  return Float4{v.x, v.y, v.z, v.w};
}
#else
inline Float4 load_float4(const float *p) { return Float4{p[0], p[1], p[2], p[3]}; }
#endif

// Stage ROWS image rows starting at (x_aligned, y0) as WORDS aligned 32-bit words each.  A window that lies inside the
// frame is fetched with plain aligned loads; otherwise words that touch the border are assembled bytewise (REFLECT_101).
template <int ROWS, int WORDS, class G>
UVIO_HD void lk_stage_rows(G g, uint32_t *dst, const ImgView &I, int x_aligned, int y0) {
  const bool aligned = ((I.pitch & 3) == 0) && ((((size_t)I.p) & 3) == 0);
  constexpr int N = ROWS * WORDS;
  if (aligned && x_aligned >= 0 && x_aligned + 4 * WORDS <= I.w && y0 >= 0 && y0 + ROWS <= I.h) {
    const uint8_t *base = I.p + (size_t)y0 * I.pitch + x_aligned;
    for (int i = g.lane(); i < N; i += g.size()) {
      int r = i / WORDS, k = i - r * WORDS;
      dst[i] = *(const uint32_t *)(base + (size_t)r * I.pitch + 4 * k);
    }
    return;
  }
  // The window sticks out of the frame.  When one reflection is enough for every coordinate it can take (always, except
  // on images smaller than the window's reach) the border handling is branch-free index arithmetic.
  const bool once = x_aligned > -I.w && x_aligned + 4 * WORDS <= 2 * I.w - 2 && y0 > -I.h && y0 + ROWS <= 2 * I.h - 2 &&
                    I.w >= 2 && I.h >= 2;
  if (once) {
    for (int i = g.lane(); i < N; i += g.size()) {
      int r = i / WORDS, k = i - r * WORDS;
      int X0 = x_aligned + 4 * k;
      const uint8_t *row = I.p + (size_t)reflect101_once(y0 + r, I.h) * I.pitch;
      uint32_t v;
      if (aligned && X0 >= 0 && X0 + 3 < I.w) {
        v = *(const uint32_t *)(row + X0);
      } else {
        v = (uint32_t)row[reflect101_once(X0, I.w)] | ((uint32_t)row[reflect101_once(X0 + 1, I.w)] << 8) |
            ((uint32_t)row[reflect101_once(X0 + 2, I.w)] << 16) | ((uint32_t)row[reflect101_once(X0 + 3, I.w)] << 24);
      }
      dst[i] = v;
    }
    return;
  }
  for (int i = g.lane(); i < N; i += g.size()) {
    int r = i / WORDS, k = i - r * WORDS;
    int X0 = x_aligned + 4 * k;
    const uint8_t *row = I.p + (size_t)reflect101(y0 + r, I.h) * I.pitch;
    uint32_t v = (uint32_t)row[reflect101(X0, I.w)] | ((uint32_t)row[reflect101(X0 + 1, I.w)] << 8) |
                 ((uint32_t)row[reflect101(X0 + 2, I.w)] << 16) | ((uint32_t)row[reflect101(X0 + 3, I.w)] << 24);
    dst[i] = v;
  }
}

template <class G>
UVIO_HD void lk_stage_template(G g, LkSmem &s, const ImgView &I, int ipx, int ipy, int iw00, int iw01, int iw10, int iw11) {
  const int xa = (ipx - 1) & ~3;
  lk_stage_rows<LK_FP, LK_FPW>(g, s.footw, I, xa, ipy - 1);
  g.sync();
  const uint8_t *foot = (const uint8_t *)s.footw + ((ipx - 1) - xa);   // foot[yy*24 + xx] = I(ip-1+xx, ip-1+yy)
  constexpr int FPP = LK_FPW * 4;
  const bool inside = ipx >= 0 && ipx + 16 <= I.w && ipy >= 0 && ipy + 16 <= I.h;   // no derivative sample falls outside
  if (inside) {
    // Blend first, differentiate second.  The bilinear blend (integer weights, no shift yet) and the Scharr stencil are
    // both exact integer linear maps, so the blend of the Scharr image equals the Scharr stencil applied to the blended
    // intensities B(x,y) = sum_w w * I(ip-1+x.., ip-1+y..), 17x17 of them (22 bits each, 27 after the stencil); the
    // rounding shifts come last, exactly where the two-step form applies them.  The template itself is the centre tap:
    // Iw = (B + 2^8) >> 9.  One pass of four byte loads per B replaces the 16x16 Scharr pass and the three blends per pixel.
    for (int i = g.lane(); i < LK_BW * LK_BW; i += g.size()) {
      int y = i / LK_BW, x = i - y * LK_BW;
      const uint8_t *c = foot + y * FPP + x;
      s.blend[i] = c[0] * iw00 + c[1] * iw01 + c[FPP] * iw10 + c[FPP + 1] * iw11;
    }
    g.sync();
    for (int i = g.lane(); i < LK_NPIX; i += g.size()) {
      int y = i / LK_WIN, x = i - y * LK_WIN;
      const int *b = s.blend + (y + 1) * LK_BW + (x + 1);
      int tl = b[-LK_BW - 1], tc = b[-LK_BW], tr = b[-LK_BW + 1], ml = b[-1], mr = b[1];
      int bl = b[LK_BW - 1], bc = b[LK_BW], br = b[LK_BW + 1];
      int ival = (b[0] + (1 << 8)) >> 9;
      int ix = (3 * (tr - tl) + 10 * (mr - ml) + 3 * (br - bl) + (1 << 13)) >> 14;
      int iy = (3 * (bl - tl) + 10 * (bc - tc) + 3 * (br - tr) + (1 << 13)) >> 14;
      s.Iw[i] = (short)ival;
      s.dIw[i] = Short2{(short)ix, (short)iy};
      float xx = (float)(ix * ix), xy = (float)(ix * iy), yy = (float)(iy * iy);
      if (x < 8) {
        s.gsimd[0][y * 8 + x] = xx;
        s.gsimd[1][y * 8 + x] = xy;
        s.gsimd[2][y * 8 + x] = yy;
      } else {
        s.gtail[0][y * 7 + x - 8] = xx;
        s.gtail[1][y * 7 + x - 8] = xy;
        s.gtail[2][y * 7 + x - 8] = yy;
      }
    }
    g.sync();
    return;
  }
  // Windows that touch the frame border: derivative samples outside the frame are zero (calcScharrDeriv + BORDER_CONSTANT), which
  // the blend must see as zeros, so the derivative image is formed first.
  for (int i = g.lane(); i < 256; i += g.size()) {
    int y = i >> 4, x = i & 15;
    int dx = 0, dy = 0;
    if (ipx + x >= 0 && ipx + x < I.w && ipy + y >= 0 && ipy + y < I.h) {
      const uint8_t *c = foot + (y + 1) * FPP + (x + 1);
      int tl = c[-FPP - 1], tc = c[-FPP], tr = c[-FPP + 1], ml = c[-1], mr = c[1];
      int bl = c[FPP - 1], bc = c[FPP], br = c[FPP + 1];
      dx = 3 * (tr - tl) + 10 * (mr - ml) + 3 * (br - bl);
      dy = 3 * (bl - tl) + 10 * (bc - tc) + 3 * (br - tr);
    }
    s.dI[i] = Short2{(short)dx, (short)dy};
  }
  g.sync();
  for (int i = g.lane(); i < LK_NPIX; i += g.size()) {
    int y = i / LK_WIN, x = i - y * LK_WIN;
    const uint8_t *c = foot + (y + 1) * FPP + (x + 1);
    int ival = (c[0] * iw00 + c[1] * iw01 + c[FPP] * iw10 + c[FPP + 1] * iw11 + (1 << 8)) >> 9;
    const Short2 *d = s.dI + (y * 16 + x);
    Short2 d00 = d[0], d01 = d[1], d10 = d[16], d11 = d[17];
    int ix = (d00.x * iw00 + d01.x * iw01 + d10.x * iw10 + d11.x * iw11 + (1 << 13)) >> 14;
    int iy = (d00.y * iw00 + d01.y * iw01 + d10.y * iw10 + d11.y * iw11 + (1 << 13)) >> 14;
    s.Iw[i] = (short)ival;
    s.dIw[i] = Short2{(short)ix, (short)iy};
    // |ix|,|iy| < 2^15: the int products are exact and (float) rounds them once, as the CPU's float multiply does
    float xx = (float)(ix * ix), xy = (float)(ix * iy), yy = (float)(iy * iy);
    if (x < 8) {
      s.gsimd[0][y * 8 + x] = xx;
      s.gsimd[1][y * 8 + x] = xy;
      s.gsimd[2][y * 8 + x] = yy;
    } else {
      s.gtail[0][y * 7 + x - 8] = xx;
      s.gtail[1][y * 7 + x - 8] = xy;
      s.gtail[2][y * 7 + x - 8] = yy;
    }
  }
  g.sync();
}

// the scalar tail of an ordered sum: 105 values added one after the other, fetched four at a time
UVIO_HD float lk_tail_chain(const float *v) {
  float q = 0.f;
#pragma unroll
  for (int j = 0; j < LK_TAIL / 4; j++) {
    Float4 f = load_float4(v + 4 * j);
    q = fadd(q, f.x);
    q = fadd(q, f.y);
    q = fadd(q, f.z);
    q = fadd(q, f.w);
  }
  return fadd(q, v[LK_TAIL - 1]);
}

// 15 ordered chains -> red[0..14]; chain = acc*5 + k, acc in {A11, A12, A22}, k = 0..3 SIMD lane, 4 = scalar tail.
template <class G>
UVIO_HD void lk_accumulate_gram(G g, LkSmem &s) {
  for (int c = g.lane(); c < 15; c += g.size()) {
    int acc = c / 5, k = c - acc * 5;
    float q = 0.f;
    if (k < 4) {
      const float *v = s.gsimd[acc] + k;
#pragma unroll
      for (int y = 0; y < LK_WIN; y++) {
        q = fadd(v[y * 8], q);
        q = fadd(v[y * 8 + 4], q);
      }
    } else {
      q = lk_tail_chain(s.gtail[acc]);
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
      const float *v = s.psimd[acc] + k;
#pragma unroll
      for (int y = 0; y < LK_WIN; y++) q = fadd(q, v[y * 4]);
    } else {
      q = lk_tail_chain(s.ptail[acc]);
    }
    s.red[c] = q;
  }
  g.sync();
}

UVIO_HD float lk_fold(const float *r) {   // scalar tail += (l0+l2)+(l1+l3)
  return fadd(r[4], fadd(fadd(r[0], r[2]), fadd(r[1], r[3])));
}

// J window of the level: rows jy0 .. jy0+31, columns from the aligned x <= jx0
template <class G>
UVIO_HD void lk_stage_window(G g, LkSmem &s, const ImgView &J, int jx0, int jy0) {
  lk_stage_rows<LK_JW, LK_JWW>(g, s.jwin, J, jx0 & ~3, jy0);
  g.sync();
}

UVIO_HD int lk_bilinear(const uint8_t *c, int pitch, int iw00, int iw01, int iw10, int iw11) {
  return (c[0] * iw00 + c[1] * iw01 + c[pitch] * iw10 + c[pitch + 1] * iw11 + (1 << 8)) >> 9;
}

// diff = J patch - I patch at integer offset (ox, oy) of the staged window, and the mismatch products.  One lane handles
// either the two pixels (y,k), (y,k+4) that meet in SIMD lane k of the CPU code, or one pixel of the scalar tail.
//
// The products are integers, so the CPU's ordered fp32 sums are EXACT (and therefore order-independent) whenever no
// partial sum can reach 2^24: every partial sum of a subset of the terms lies in [-NEG, POS] (NEG / POS = the sums of the
// negative / positive terms).  With max(POS, NEG) < 2^24 the result is fp32(integer sum), which a warp reduction
// delivers; this holds for ~90 % of the iterations.  Otherwise the products go to shared memory and the ordered chains of
// lk_accumulate_mismatch run.  Returns true when (b1, b2) (unscaled) were produced here.
template <class G>
UVIO_HD bool lk_mismatch_products(G g, LkSmem &s, int jx0, int ox, int oy, int iw00, int iw01, int iw10, int iw11, float &b1,
                                  float &b2) {
  constexpr int JP = LK_JWW * 4;
  constexpr int NPAIR = LK_WIN * 4;
  constexpr int RP = (NPAIR + G::SIZE - 1) / G::SIZE, RT = (LK_TAIL + G::SIZE - 1) / G::SIZE;
  typedef typename G::acc_t acc_t;
  const uint8_t *win = (const uint8_t *)s.jwin + (jx0 - (jx0 & ~3)) + oy * JP + ox;
  int pa[RP], pb[RP], ta[RT], tb[RT];
  acc_t sum1 = 0, pos1 = 0, sum2 = 0, pos2 = 0;
#pragma unroll
  for (int r = 0; r < RP; r++) {
    int i = g.lane() + r * G::SIZE;
    int a = 0, b = 0;
    if (i < NPAIR) {
      int y = i >> 2, k = i & 3, e = y * LK_WIN + k;
      const uint8_t *c = win + y * JP + k;
      int d0 = lk_bilinear(c, JP, iw00, iw01, iw10, iw11) - s.Iw[e];
      int d1 = lk_bilinear(c + 4, JP, iw00, iw01, iw10, iw11) - s.Iw[e + 4];
      Short2 g0 = s.dIw[e], g1 = s.dIw[e + 4];
      a = d0 * g0.x + d1 * g1.x;
      b = d0 * g0.y + d1 * g1.y;
    }
    pa[r] = a;
    pb[r] = b;
    sum1 += a;
    pos1 += a > 0 ? a : 0;
    sum2 += b;
    pos2 += b > 0 ? b : 0;
  }
#pragma unroll
  for (int r = 0; r < RT; r++) {
    int t = g.lane() + r * G::SIZE;
    int a = 0, b = 0;
    if (t < LK_TAIL) {
      int y = t / 7, x = 8 + t - 7 * y, e = y * LK_WIN + x;
      int d = lk_bilinear(win + y * JP + x, JP, iw00, iw01, iw10, iw11) - s.Iw[e];
      Short2 gd = s.dIw[e];
      a = d * gd.x;
      b = d * gd.y;
    }
    ta[r] = a;
    tb[r] = b;
    sum1 += a;
    pos1 += a > 0 ? a : 0;
    sum2 += b;
    pos2 += b > 0 ? b : 0;
  }
  // |term| <= 2 * 8160 * 4080 and a lane of a warp owns at most 6 terms: the int accumulators cannot overflow
  const acc_t LIM = (acc_t)1 << 24;
  acc_t neg1 = pos1 - sum1, neg2 = pos2 - sum2;
  unsigned P1 = g.reduce_add((unsigned)(pos1 < LIM ? pos1 : LIM)), N1 = g.reduce_add((unsigned)(neg1 < LIM ? neg1 : LIM));
  unsigned P2 = g.reduce_add((unsigned)(pos2 < LIM ? pos2 : LIM)), N2 = g.reduce_add((unsigned)(neg2 < LIM ? neg2 : LIM));
  unsigned worst = P1 > N1 ? P1 : N1, w2 = P2 > N2 ? P2 : N2;
  worst = worst > w2 ? worst : w2;
  if (worst < (unsigned)LIM) {
    b1 = (float)((int)P1 - (int)N1);
    b2 = (float)((int)P2 - (int)N2);
    return true;
  }
#pragma unroll
  for (int r = 0; r < RP; r++) {
    int i = g.lane() + r * G::SIZE;
    if (i < NPAIR) {
      s.psimd[0][i] = (float)pa[r];
      s.psimd[1][i] = (float)pb[r];
    }
  }
#pragma unroll
  for (int r = 0; r < RT; r++) {
    int t = g.lane() + r * G::SIZE;
    if (t < LK_TAIL) {
      s.ptail[0][t] = (float)ta[r];
      s.ptail[1][t] = (float)tb[r];
    }
  }
  g.sync();
  return false;
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
    int jx0 = 0, jy0 = 0;
    bool have_window = false;
    for (int j = 0; j < P.max_count; j++) {
      int inx = floor_f2i(npx), iny = floor_f2i(npy);
      if (inx < -LK_WIN || inx >= J.w || iny < -LK_WIN || iny >= J.h) {
        if (level == 0) status = 0;
        break;
      }
      lk_weights(fsub(npx, (float)inx), fsub(npy, (float)iny), iw00, iw01, iw10, iw11);
      if (!have_window || inx < jx0 || inx > jx0 + (LK_JW - 16) || iny < jy0 || iny > jy0 + (LK_JW - 16)) {
        jx0 = inx - (LK_JW - 16) / 2;   // (re)centre the staged window on the current estimate
        jy0 = iny - (LK_JW - 16) / 2;
        lk_stage_window(g, s, J, jx0, jy0);
        have_window = true;
      }
      float b1, b2;
      if (work) work[1]++;   // iterations
      if (!lk_mismatch_products(g, s, jx0, inx - jx0, iny - jy0, iw00, iw01, iw10, iw11, b1, b2)) {
        if (work) work[2]++;   // iterations that needed the ordered chains
        lk_accumulate_mismatch(g, s);
        b1 = lk_fold(s.red + 0);
        b2 = lk_fold(s.red + 5);
        g.sync();
      }
      b1 = fmul(b1, FLT_SCALE);
      b2 = fmul(b2, FLT_SCALE);
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
