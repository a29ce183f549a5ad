// The reference's descriptor tracker (ov_core/src/track/TrackDescriptor.cpp, Grider_FAST.h) as device code: per-cell top-K in
// std::sort order and the mask test, ORB descriptors of the refined corners (cv::ORB::compute on provided key points: the 7x7
// sigma-2 blur of a sub-matrix, the 256 learned point pairs), brute-force Hamming 2-nearest-neighbour matching in both directions
// with the ratio and symmetry tests, fundamental-matrix outlier rejection, the stereo / mono id assignment and the temporal
// merge.  FAST, cornerSubPix, undistortion and RANSAC are the bodies the KLT path already has (fast_body.h, subpix_body.h,
// ransac_body.h); equalisation is k_preprocess with a one-level pyramid.
//
// As in track_body.h the reference's sequential loops are restated order-free: "push_back the survivors" = ordered
// compaction by block-wide exclusive scans, "skip if the occupancy cell is already taken" = the first point per cell wins
// (atomicMin of the list index), "the match whose trainIdx is i" = an inverse map (symmetric matches are one to one).
// The same text is compiled for the host by tests/hostsim (GroupSeq, one lane).
#pragma once
#include "fast_body.h"
#include "featdb_body.h"   // fdb_min / fdb_add_ret / fdb_ldi
#include "kernels.h"       // SubpixItem, PyrDesc
#include "ransac_body.h"
#include "simt.h"
#include "orb_tables.h"    // ORB_BLUR_TAPS (the point pairs are passed in)
#include "track_body.h"    // dv_sort_prefix, dv_bitonic_by_position, DvObs

namespace uvio {

constexpr int DESC_WORDS = 8;          // 256-bit descriptor
constexpr int DESC_EDGE = 31;          // cv::ORB edgeThreshold (default): key points nearer to the border are dropped by compute()
constexpr int DESC_MAX_P = 4096;       // most key points per camera a tracker may be sized for

struct DescSide {                      // one camera of a unit
  ImgView img;                         // equalised frame (device memory)
  const uint8_t *mask;                 // w x h, pitch w; all zeros when the message carries none
  unsigned long long cam_id;
  int cam;                             // index into the CamModel table of the launch
  int pyr;                             // index into the PyrDesc table of the launch (one level: img)
  int w, h;
  int job0, ncells;                    // FAST jobs of this side, [job0, job0 + ncells) in Grider_FAST's r order (Grider_FAST.h:105-108)
  int slot0;                           // first of this side's P slots in the launch-wide sub-pixel result array
  uint32_t *cell_top;                  // [ncells * K] per cell: the top-K corner words that passed the mask test (image coordinates)
  int *cell_n;                         // [ncells]
  int *cnt;                            // [0] extracted (pts_ext), [1] after ORB's border filter, [2] new list, [3] last list
  float *kp_xy;                        // [2P] key points that have descriptors (pts_ext after compute())
  uint32_t *kp_desc;                   // [8P]
  float *nw_xy;                        // [2P] this frame's list (pts0 / pts1 of perform_detection_*)
  uint32_t *nw_desc;
  unsigned long long *nw_ids;
  float *la_xy;                        // [2P] pts_last / desc_last / ids_last (TrackBase.h:186-189, TrackDescriptor.h:170)
  uint32_t *la_desc;
  unsigned long long *la_ids;
};

struct DescUnit {                      // a stereo pair or one mono camera of one tracker
  DescSide s[2];
  int stereo;
  int P, K, mpd;
  int gw[2], gh[2];                    // occupancy grid sizes (TrackDescriptor.cpp:374, :436-439)
  double knn_ratio;
  double thr[3];                       // 1 / max focal length of the two cameras of match job 0 (L->R), 1 (L->L), 2 (R->R)  (:530-533)
  unsigned long long *currid;          // TrackBase::currid after the feed (device copy, written by the tracker's last unit)
  unsigned long long id_base;          // TrackBase::currid before this feed (the host mirrors it: every feed ends with a synchronisation)
  const int *sib_cnt[3];               // a non-stereo tracker fed several cameras at once: the counters (DescSide::cnt) of its units BEFORE this
                                       // one, in message order (null-terminated): their new points take their ids first
  int last_of_rig;                     // this unit publishes the tracker's id counter
  int *grid;                           // mono occupancy grid (gw[0] * gh[0] words, DV_INF between feeds)
  int *m_q, *m_t;                      // [3P] the matches of the three jobs (queryIdx, trainIdx), in robust_match's order
  int *m_n;                            // [3]
  int *fwd, *bwd;                      // [3P] scratch: surviving best match per query / per train descriptor (-1: none)
  float *p0n, *p1n;                    // [3 * 2P] scratch: normalised points handed to RANSAC
  uint8_t *rs_mask;                    // [3P]
  DvObs *obs;                          // [2P] this feed's update_feature calls, in order (mapped host memory)
  int *head;                           // mapped host memory: [0] observations, [1] points in the last list, [2..3] currid after the feed
};

struct DescGlobals {                   // launch-wide tables and scratch
  const FastJob *fast_jobs;
  uint32_t *fast_out;
  unsigned *fast_cnt;
  SubpixItem *sp_items;
  int *n_sp;
  float *sp_out;                       // [2 * slots]
  const CamModel *cams;
};

struct OrbSmem {
  uint8_t pix[37 * 40];
  float hor[37 * 32];
  uint8_t blur[31 * 32];
};

UVIO_HD int desc_popc(uint32_t v) {
#ifdef __CUDA_ARCH__
  return __popc(v);
#else
  return __builtin_popcount(v);
#endif
}

// ---- Grider_FAST: top-K of one cell in std::sort order, then the mask test (Grider_FAST.h:120-145) -------------------------------
template <class G>
UVIO_HD void desc_topk(G g, const DescUnit &U, const DescGlobals &X, int side, int r, uint32_t *smem, int smem_cap) {
  const DescSide &S = U.s[side];
  const int slot = S.job0 + r;
  const FastJob job = X.fast_jobs[slot];
  if (job.w == 0) {                                         // the cell would leave the image (:112)
    if (g.lane() == 0) S.cell_n[r] = 0;
    return;
  }
  unsigned n = X.fast_cnt[slot];
  if (n > (unsigned)job.out_cap) n = (unsigned)job.out_cap;
  int m = 1;
  while (m < (int)n) m <<= 1;
  uint32_t *src = X.fast_out + job.out_off;
  uint32_t *buf = m <= smem_cap ? smem : src;               // the job's own region holds the next power of two
  if (buf != src)
    for (int i = g.lane(); i < (int)n; i += g.size()) buf[i] = src[i];
  for (int i = (int)n + g.lane(); i < m; i += g.size()) buf[i] = 0xFFFFFFFFu;
  g.sync();
  dv_bitonic_by_position(g, buf, m);                        // cv::FAST emits row-major: the order std::sort starts from
  if (g.lane() == 0) {
    dv_sort_prefix(buf, (long)n, (long)U.K);                // :123
    const int kn = (int)n < U.K ? (int)n : U.K;
    int kept = 0;
    for (int k = 0; k < kn; k++) {                          // :128-145
      const uint32_t w = buf[k];
      const int x = job.x0 + (int)(w & 0xFFFu), y = job.y0 + (int)((w >> 12) & 0xFFFu);
      if (x < 0 || x > S.w || y < 0 || y > S.h) continue;   // :136 (cannot happen)
      if (S.mask[(size_t)y * S.w + x] > 127) continue;      // :141
      S.cell_top[r * U.K + kept++] = (uint32_t)x | ((uint32_t)y << 12) | (w & 0xFF000000u);
    }
    S.cell_n[r] = kept;
    X.fast_cnt[slot] = 0;                                   // re-armed for the next feed
  }
  g.sync();
}

// ---- the cells' lists joined in r order (:150-152) -> sub-pixel items (:158-172); one group per side ------------------------------
template <class G>
UVIO_HD void desc_gather(G g, const DescUnit &U, const DescGlobals &X, int side, int *s_base) {
  const DescSide &S = U.s[side];
  int total = 0;
  for (int base = 0; base < S.ncells; base += g.size()) {   // pass 1: how many
    const int r = base + g.lane();
    int t;
    g.exclusive_sum(r < S.ncells ? S.cell_n[r] : 0, t);
    total += t;
  }
  if (g.lane() == 0) {
    *s_base = total ? fdb_add_ret(X.n_sp, total) : 0;
    S.cnt[0] = total;
  }
  g.sync();
  const int item0 = *s_base;
  int off = 0;
  for (int base = 0; base < S.ncells; base += g.size()) {   // pass 2: where
    const int r = base + g.lane();
    const int n = r < S.ncells ? S.cell_n[r] : 0;
    int t;
    const int pos = off + g.exclusive_sum(n, t);
    for (int k = 0; k < n; k++) {
      const uint32_t w = S.cell_top[r * U.K + k];
      SubpixItem it;
      it.x = (float)(w & 0xFFFu);
      it.y = (float)((w >> 12) & 0xFFFu);
      it.pyr = S.pyr;
      it.pad = S.slot0 + pos + k;
      X.sp_items[item0 + pos + k] = it;
    }
    off += t;
  }
  g.sync();
}

// ---- cv::ORB::compute on provided key points -----------------------------------------------------------------------------------------
// KeyPointsFilter::runByImageBorder(keypoints, size, 31): the key point is converted to an integer point (cvRound) and must
// lie inside Rect(31, 31, w - 62, h - 62).
UVIO_HD bool orb_keeps(float x, float y, int w, int h) {
  if (w <= 2 * DESC_EDGE || h <= 2 * DESC_EDGE) return false;
  const int xr = round_f2i(x), yr = round_f2i(y);
  return xr >= DESC_EDGE && xr < w - DESC_EDGE && yr >= DESC_EDGE && yr < h - DESC_EDGE;
}

// 256-bit descriptor of the key point at integer position (cx, cy) (= cvRound of its sub-pixel position, level 0): the 31x31
// patch of the blurred image around it and bit i = blurred(a_i) < blurred(b_i) for the 256 point pairs of `pairs`
// (tools/make_orb_pattern.py; FAST key points carry angle -1 deg, which rotates no pair off its pixel).  The blur is
// cv::GaussianBlur(Size(7,7), 2, 2) as ORB gets it: its pyramid level is a sub-matrix of a bordered buffer, which OpenCV
// filters with sepFilter2D and the fp32 taps of getGaussianKernel(7, 2, CV_32F) -- rows tap by tap, columns in the symmetric
// form, one cvRound at the end -- not with the fixed-point kernel whole 8-bit images get.  out: 8 words, byte j = bits 8j..8j+7.
template <class G>
UVIO_HD void orb_describe(G g, OrbSmem &s, const ImgView &I, int cx, int cy, const int8_t *pairs, uint32_t *out) {
  const float k[7] = ORB_BLUR_TAPS;
  for (int i = g.lane(); i < 37 * 37; i += g.size()) {
    const int r = i / 37, c = i - r * 37;
    s.pix[r * 40 + c] = I.p[(size_t)(cy - 18 + r) * I.pitch + (cx - 18 + c)];
  }
  g.sync();
  for (int i = g.lane(); i < 37 * 31; i += g.size()) {      // RowFilter<uchar, float>: s = k0*S0; s += k1*S1; ...
    const int r = i / 31, c = i - r * 31;
    const uint8_t *p = s.pix + r * 40 + c;
    float v = fmul(k[0], (float)p[0]);
    for (int t = 1; t < 7; t++) v = fadd(v, fmul(k[t], (float)p[t]));
    s.hor[r * 32 + c] = v;
  }
  g.sync();
  for (int i = g.lane(); i < 31 * 31; i += g.size()) {      // SymmColumnFilter<Cast<float, uchar>>: centre, then mirrored pairs
    const int r = i / 31, c = i - r * 31;
    const float *h = s.hor + r * 32 + c;
    float v = fmul(k[3], h[3 * 32]);
    v = fadd(v, fmul(k[4], fadd(h[4 * 32], h[2 * 32])));
    v = fadd(v, fmul(k[5], fadd(h[5 * 32], h[1 * 32])));
    v = fadd(v, fmul(k[6], fadd(h[6 * 32], h[0])));
    s.blur[r * 32 + c] = sat_u8(round_f2i(v));
  }
  g.sync();
  for (int b = g.lane(); b < 32; b += g.size()) {
    unsigned byte = 0;
    for (int k = 0; k < 8; k++) {
      const int8_t *q = pairs + 4 * (8 * b + k);
      const unsigned va = s.blur[(q[1] + 15) * 32 + q[0] + 15], vb = s.blur[(q[3] + 15) * 32 + q[2] + 15];
      byte |= (va < vb ? 1u : 0u) << k;
    }
    s.pix[b] = (uint8_t)byte;
  }
  g.sync();
  for (int w = g.lane(); w < DESC_WORDS; w += g.size())
    out[w] = (uint32_t)s.pix[4 * w] | ((uint32_t)s.pix[4 * w + 1] << 8) | ((uint32_t)s.pix[4 * w + 2] << 16) | ((uint32_t)s.pix[4 * w + 3] << 24);
  g.sync();
}

// rank of extracted point i among the points compute() keeps, for every i of one side (rank[i] = -1: dropped); also writes the
// kept positions and the count when `publish`.  rank: at least U.P shorts private to the group (P <= DESC_MAX_P).
template <class G>
UVIO_HD int desc_orb_ranks(G g, const DescUnit &U, const DescGlobals &X, int side, short *rank, bool publish) {
  const DescSide &S = U.s[side];
  const int n = fdb_ldi(&S.cnt[0]);
  int off = 0;
  for (int base = 0; base < n; base += g.size()) {
    const int i = base + g.lane();
    float x = 0.f, y = 0.f;
    int keep = 0;
    if (i < n) {
      x = X.sp_out[2 * (S.slot0 + i)];
      y = X.sp_out[2 * (S.slot0 + i) + 1];
      keep = orb_keeps(x, y, S.w, S.h) ? 1 : 0;
    }
    int t;
    const int pos = off + g.exclusive_sum(keep, t);
    if (i < n) rank[i] = (short)(keep ? pos : -1);
    if (keep && publish) {
      S.kp_xy[2 * pos] = x;
      S.kp_xy[2 * pos + 1] = y;
    }
    off += t;
  }
  if (publish && g.lane() == 0) S.cnt[1] = off;
  g.sync();
  return off;
}

// ---- BruteForce-Hamming knnMatch(k = 2) + ratio test for one query (TrackDescriptor.cpp:494-499, :542-556) ---------------------------
// knnMatch returns the two nearest in ascending distance, the lower index first among equals; fewer than two train descriptors
// leave fewer than two neighbours and the ratio test clears the query.  Returns the surviving best train index or -1.
UVIO_HD int desc_best(const uint32_t *dq, const uint32_t *dt, int nt, double knn_ratio) {
  if (nt < 2) return -1;
  uint32_t q[DESC_WORDS];
  for (int w = 0; w < DESC_WORDS; w++) q[w] = dq[w];
  int i0 = -1, d0 = 0x7fffffff, d1 = 0x7fffffff;
  for (int j = 0; j < nt; j++) {
    const uint32_t *t = dt + (size_t)DESC_WORDS * j;
    int d = 0;
    for (int w = 0; w < DESC_WORDS; w++) d += desc_popc(q[w] ^ t[w]);
    if (d < d0) { d1 = d0; d0 = d; i0 = j; }
    else if (d < d1) d1 = d;
  }
  const float ratio = (float)d0 / (float)d1;               // DMatch::distance is a float; 0 / 0 = NaN compares false: kept
  return ((double)ratio > knn_ratio) ? -1 : i0;
}

// ---- robust_match (TrackDescriptor.cpp:487-540) of one job ----------------------------------------------------------------------------
// job 0: this frame's left key points -> right key points; 1: last left list -> new left list; 2: last right -> new right.
// Leaves the matches in U.m_q / U.m_t [job * P ...] and their count in U.m_n[job].
template <class G>
UVIO_HD void desc_match(G g, RansacSmem &rs, const DescUnit &U, const DescGlobals &X, int job) {
  const DescSide &A = U.s[job == 2 ? 1 : 0], &B = U.s[job == 1 ? 0 : 1];
  const float *xa = job == 0 ? A.kp_xy : A.la_xy, *xb = job == 0 ? B.kp_xy : B.nw_xy;
  const uint32_t *da = job == 0 ? A.kp_desc : A.la_desc, *db = job == 0 ? B.kp_desc : B.nw_desc;
  const int na = fdb_ldi(&A.cnt[job == 0 ? 1 : 3]), nb = fdb_ldi(&B.cnt[job == 0 ? 1 : 2]);
  int *fwd = U.fwd + job * U.P, *bwd = U.bwd + job * U.P;
  int *mq = U.m_q + job * U.P, *mt = U.m_t + job * U.P;
  for (int i = g.lane(); i < na; i += g.size()) fwd[i] = desc_best(da + (size_t)DESC_WORDS * i, db, nb, U.knn_ratio);
  for (int j = g.lane(); j < nb; j += g.size()) bwd[j] = desc_best(db + (size_t)DESC_WORDS * j, da, na, U.knn_ratio);
  g.sync();
  // symmetry test (:558-575): matches1 in query order; the partner is the entry of matches2 whose query index is our train index
  int ngood = 0;
  for (int base = 0; base < na; base += g.size()) {
    const int i = base + g.lane();
    int keep = 0, t = -1;
    if (i < na) {
      t = fwd[i];
      keep = (t >= 0 && bwd[t] == i) ? 1 : 0;
    }
    int tot;
    const int pos = ngood + g.exclusive_sum(keep, tot);
    if (keep) { mq[pos] = i; mt[pos] = t; }
    ngood += tot;
  }
  g.sync();
  if (ngood < 10) {                                         // :517-518
    if (g.lane() == 0) U.m_n[job] = 0;
    g.sync();
    return;
  }
  const CamModel ca = X.cams[A.cam], cb = X.cams[B.cam];
  float *p0 = U.p0n + (size_t)job * 2 * U.P, *p1 = U.p1n + (size_t)job * 2 * U.P;
  uint8_t *mask = U.rs_mask + (size_t)job * U.P;
  for (int k = g.lane(); k < ngood; k += g.size()) {        // :522-526
    undistort_point(ca, xa[2 * mq[k]], xa[2 * mq[k] + 1], p0[2 * k], p0[2 * k + 1]);
    undistort_point(cb, xb[2 * mt[k]], xb[2 * mt[k] + 1], p1[2 * k], p1[2 * k + 1]);
  }
  g.sync();
  const int found = ransac_job(g, rs, p0, p1, ngood, U.thr[job], 0.999, 1000, mask);   // :533
  g.sync();
  // keep the inliers, in order (:536-539).  No model: the reference reads an empty mask (undefined); nothing survives here.
  int nin = 0;
  for (int base = 0; base < ngood; base += g.size()) {
    const int k = base + g.lane();
    int keep = 0, q = 0, t = 0;
    if (k < ngood) {
      q = mq[k];
      t = mt[k];
      keep = (found && mask[k] == 1) ? 1 : 0;
    }
    int tot;
    const int pos = nin + g.exclusive_sum(keep, tot);      // (the scan synchronises: every read above precedes the writes below)
    if (keep) { mq[pos] = q; mt[pos] = t; }
    nin += tot;
  }
  if (g.lane() == 0) U.m_n[job] = nin;
  g.sync();
}

UVIO_HD void desc_copy_desc(uint32_t *dst, const uint32_t *src) {
  for (int w = 0; w < DESC_WORDS; w++) dst[w] = src[w];
}

// ---- perform_detection_stereo, after the L->R match (TrackDescriptor.cpp:433-484) --------------------------------------------------------
template <class G>
UVIO_HD void desc_assign_stereo(G g, const DescUnit &U) {
  const DescSide &L = U.s[0], &R = U.s[1];
  const int n = fdb_ldi(&U.m_n[0]);
  const float mpd = (float)U.mpd;
  const unsigned long long id0 = U.id_base;
  int off = 0;
  for (int base = 0; base < n; base += g.size()) {
    const int k = base + g.lane();
    int keep = 0, q = 0, t = 0;
    if (k < n) {
      q = U.m_q[k];
      t = U.m_t[k];
      const float ax = L.kp_xy[2 * q], ay = L.kp_xy[2 * q + 1], bx = R.kp_xy[2 * t], by = R.kp_xy[2 * t + 1];
      const int x0 = (int)ax, y0 = (int)ay, x0g = (int)(ax / mpd), y0g = (int)(ay / mpd);
      const int x1 = (int)bx, y1 = (int)by, x1g = (int)(bx / mpd), y1g = (int)(by / mpd);
      const bool out0 = x0g < 0 || x0g >= U.gw[0] || y0g < 0 || y0g >= U.gh[0] || x0 < 0 || x0 >= L.w || y0 < 0 || y0 >= L.h;   // :452-455
      const bool out1 = x1g < 0 || x1g >= U.gw[1] || y1g < 0 || y1g >= U.gh[1] || x1 < 0 || x1 >= L.w || y1 < 0 || y1 >= L.h;   // :460-463 (img0's size, as written)
      // :468-469 tests the two occupancy grids, which the reference never marks in the stereo path: no point is rejected by them
      keep = (!out0 && !out1) ? 1 : 0;
    }
    int tot;
    const int pos = off + g.exclusive_sum(keep, tot);
    if (keep) {                                             // :472-483
      L.nw_xy[2 * pos] = L.kp_xy[2 * q];
      L.nw_xy[2 * pos + 1] = L.kp_xy[2 * q + 1];
      R.nw_xy[2 * pos] = R.kp_xy[2 * t];
      R.nw_xy[2 * pos + 1] = R.kp_xy[2 * t + 1];
      desc_copy_desc(L.nw_desc + (size_t)DESC_WORDS * pos, L.kp_desc + (size_t)DESC_WORDS * q);
      desc_copy_desc(R.nw_desc + (size_t)DESC_WORDS * pos, R.kp_desc + (size_t)DESC_WORDS * t);
      L.nw_ids[pos] = R.nw_ids[pos] = id0 + 1 + (unsigned long long)pos;   // size_t temp = ++currid
    }
    off += tot;
  }
  g.sync();
  if (g.lane() == 0) L.cnt[2] = R.cnt[2] = off;
  g.sync();
}

// ---- perform_detection_monocular, after compute() (TrackDescriptor.cpp:371-404) ------------------------------------------------------------
template <class G>
UVIO_HD void desc_assign_mono(G g, const DescUnit &U) {
  const DescSide &S = U.s[0];
  const int n = fdb_ldi(&S.cnt[1]);
  const float mpd = (float)U.mpd;
  // the first point (in list order) that is in bounds takes its occupancy cell; later ones in the same cell are skipped (:394-403)
  for (int i = g.lane(); i < n; i += g.size()) {
    const float px = S.kp_xy[2 * i], py = S.kp_xy[2 * i + 1];
    const int x = (int)px, y = (int)py, xg = (int)(px / mpd), yg = (int)(py / mpd);
    if (xg < 0 || xg >= U.gw[0] || yg < 0 || yg >= U.gh[0] || x < 0 || x >= S.w || y < 0 || y >= S.h) continue;   // :389-391
    fdb_min(&U.grid[yg * U.gw[0] + xg], i);
  }
  g.sync();
  int off = 0;
  for (int base = 0; base < n; base += g.size()) {
    const int i = base + g.lane();
    int keep = 0, cell = -1;
    if (i < n) {
      const float px = S.kp_xy[2 * i], py = S.kp_xy[2 * i + 1];
      const int x = (int)px, y = (int)py, xg = (int)(px / mpd), yg = (int)(py / mpd);
      if (!(xg < 0 || xg >= U.gw[0] || yg < 0 || yg >= U.gh[0] || x < 0 || x >= S.w || y < 0 || y >= S.h)) {
        cell = yg * U.gw[0] + xg;
        keep = fdb_ldi(&U.grid[cell]) == i ? 1 : 0;
      }
    }
    int tot;
    const int pos = off + g.exclusive_sum(keep, tot);
    if (keep) {
      S.nw_xy[2 * pos] = S.kp_xy[2 * i];
      S.nw_xy[2 * pos + 1] = S.kp_xy[2 * i + 1];
      desc_copy_desc(S.nw_desc + (size_t)DESC_WORDS * pos, S.kp_desc + (size_t)DESC_WORDS * i);
      S.nw_ids[pos] = (unsigned long long)pos;              // rank; desc_merge adds the id counter
    }
    off += tot;
  }
  g.sync();
  // ids: ranks here; desc_merge adds the tracker's counter and the new points of the cameras before this one (the reference
  // feeds the cameras of a non-stereo tracker in parallel, each id drawn with ++currid (:50-55): camera by camera in message
  // order is one of the orders that can happen, and the one that does not change from run to run)
  if (g.lane() == 0) {
    U.m_n[0] = 0;
    S.cnt[2] = off;
  }
  g.sync();
  for (int i = g.lane(); i < n; i += g.size()) {            // re-arm the grid
    const float px = S.kp_xy[2 * i], py = S.kp_xy[2 * i + 1];
    const int xg = (int)(px / mpd), yg = (int)(py / mpd);
    if (xg >= 0 && xg < U.gw[0] && yg >= 0 && yg < U.gh[0]) U.grid[yg * U.gw[0] + xg] = DV_INF;
  }
  g.sync();
}

// ---- temporal merge, database update and "move forward in time" (feed_stereo :252-345, feed_monocular :115-168) ----------------------------
// inv: 2P ints private to the group (the match whose trainIdx is i, per side)
template <class G>
UVIO_HD void desc_merge(G g, const DescUnit &U, const DescGlobals &X, int *inv) {
  const int sides = U.stereo ? 2 : 1;
  const int n = fdb_ldi(&U.s[0].cnt[2]);
  const bool first = fdb_ldi(&U.s[0].cnt[3]) == 0 || (U.stereo && fdb_ldi(&U.s[1].cnt[3]) == 0);   // :208 / :86
  int n_obs = 0;
  unsigned long long before = 0;                            // new points of this tracker's cameras that come before this one
  for (int k = 0; k < 3 && U.sib_cnt[k]; k++) before += (unsigned long long)fdb_ldi(&U.sib_cnt[k][2]);
  if (!U.stereo) {                                          // size_t temp = ++currid (:399), camera by camera
    for (int i = g.lane(); i < n; i += g.size()) U.s[0].nw_ids[i] += U.id_base + before + 1;
    g.sync();
  }
  if (!first) {
    for (int i = g.lane(); i < sides * U.P; i += g.size()) inv[i] = -1;
    g.sync();
    for (int sd = 0; sd < sides; sd++) {
      const int job = 1 + sd, nm = fdb_ldi(&U.m_n[job]);
      for (int k = g.lane(); k < nm; k += g.size()) inv[sd * U.P + U.m_t[job * U.P + k]] = U.m_q[job * U.P + k];   // one to one
    }
    g.sync();
    for (int i = g.lane(); i < n; i += g.size()) {
      const int idll = inv[i];
      unsigned long long id = U.s[0].nw_ids[i];             // ids_left_new for both cameras (:304-305)
      if (U.stereo) {
        const int idrr = inv[U.P + i];
        if (idll != -1 && idrr != -1 && U.s[0].la_ids[idll] == U.s[1].la_ids[idrr]) id = U.s[0].la_ids[idll];   // :289
      } else if (idll != -1) {
        id = U.s[0].la_ids[idll];                           // :141-142
      }
      for (int sd = 0; sd < sides; sd++) {                  // :311-326 / :152-155
        const DescSide &S = U.s[sd];
        DvObs o;
        o.id = id;
        o.cam_id = S.cam_id;
        o.u = S.nw_xy[2 * i];
        o.v = S.nw_xy[2 * i + 1];
        undistort_point(X.cams[S.cam], o.u, o.v, o.un, o.vn);
        U.obs[sides * i + sd] = o;
        S.nw_ids[i] = id;
      }
    }
    n_obs = sides * n;
    g.sync();
  }
  for (int sd = 0; sd < sides; sd++) {                      // :333-345 / :161-168
    const DescSide &S = U.s[sd];
    for (int i = g.lane(); i < n; i += g.size()) {
      S.la_xy[2 * i] = S.nw_xy[2 * i];
      S.la_xy[2 * i + 1] = S.nw_xy[2 * i + 1];
      S.la_ids[i] = S.nw_ids[i];
    }
    for (int i = g.lane(); i < n * DESC_WORDS; i += g.size()) S.la_desc[i] = S.nw_desc[i];
    if (g.lane() == 0) S.cnt[3] = n;
  }
  if (g.lane() == 0) {
    const unsigned long long c = U.id_base + before + (unsigned long long)n;
    if (U.last_of_rig) *U.currid = c;
    U.head[0] = n_obs;
    U.head[1] = n;
    U.head[2] = (int)(unsigned)(c & 0xFFFFFFFFull);
    U.head[3] = (int)(unsigned)(c >> 32);
  }
  g.sync();
}

}  // namespace uvio
