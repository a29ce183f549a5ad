// The reference's per-frame bookkeeping as DEVICE code: occupancy passes, cell selection, per-cell top-K in std::sort order,
// min-distance reject, id assignment, the stereo merge and the observation list -- everything engine.cpp's host path does
// between its three stream synchronisations (hostlogic.h, phase_a .. phase_d) -- so that a feed is one chain of kernel
// launches with a single synchronisation at its end and per-GPU throughput no longer depends on host cores.
//
// The reference's loops are sequential and order-dependent (TrackKLT.cpp:409-464, :496-527, :675-708, :306-354).  They are
// kept exact, not by running them on one thread, but by stating what each loop computes in order-free terms:
//   * "erase if the close-grid cell is already occupied" keeps, per cell, the FIRST point (in list order) that passes the
//     order-independent tests  ->  an atomicMin of the list index per cell;
//   * "push_back the survivors" / "id = ++currid" are ordered compactions  ->  block-wide exclusive scans over keep flags;
//   * membership tests (std::find over id vectors) are brute-force comparisons over <= a few hundred ids per camera.
// One thread block handles one unit (a stereo pair or a mono camera) of one tracker.  The per-cell top-K keeps the
// reference's std::sort tie order with a replica of libstdc++'s introsort run by one lane per cell (dv_sort_prefix).
//
// The same text is compiled for the host by tests/hostsim (GroupSeq, one lane), where it runs against the golden sequences.
#pragma once
#include "fast_body.h"
#include "image_body.h"   // Vec16
#include "featdb_body.h"   // fdb_scan, fdb_min / fdb_add, fdb_ldi
#include "kernels.h"
#include "simt.h"

namespace uvio {

constexpr int DV_INF = 0x7fffffff;

struct DvObs {                    // == Obs / UvioObs
  unsigned long long id, cam_id;
  float u, v, un, vn;
};

struct DvSide {                   // one camera of a unit during one feed
  int w, h;
  int pyr_slot[3];                // the camera's three pyramids (table indices); which is which in this feed comes with the feed bits
  const uint8_t *mask_slot[3];    // device copies of the masks that go with them
  int det_pyr, last_pyr, cur_pyr; // (derived per feed by dv_apply_feed) detection image, temporal LK from / to
  int cam_model;
  unsigned long long cam_id;
  double thr;                     // 2.0 / max focal length (TrackKLT.cpp:870-873)
  const uint8_t *det_mask;        // (derived) mask of the detection image, pitch w (null = all zeros)
  const uint8_t *cur_mask;        // (derived) mask of the message being fed (mono merge, :166)
  // persistent state (TrackBase.h:186-189 pts_last / ids_last), device memory
  float *pts_last;                // [2 * P]
  unsigned long long *ids_last;   // [P]
  int *n_last;
  // per-feed work, device memory
  float *pts;                     // [2 * P] kept old points, then the new ones: pts_old of perform_matching
  unsigned long long *ids;        // [P]
  uint8_t *close;                 // [cw * ch]  grid_2d_close
  int *first;                     // [cw * ch]  first-come scratch
  int *grid;                      // [gx * gy]  grid_2d_grid (read saturated at 255)
  int *cells;                     // [C] cells FAST runs on, (x << 16) | y, x-outer / y-inner
  uint32_t *top;                  // [C * K] per cell: top-K corner words (cell-relative) in std::sort order
  int *top_n;                     // [C]
  int *ctl;                       // [0] points in pts/ids, [1] cells, [2] detect, [3] lk_run, [4] kept old points
  // host mirror of the state after the feed (mapped host memory)
  float *h_pts_last;
  unsigned long long *h_ids_last;
  int *h_n_last;
};

struct DvUnit {
  int stereo, first;
  int publish_last;                                // copy pts_last / ids_last to the host mirror at the end of the feed (someone reads them)
  int num_features, gx, gy, mpd, fast_threshold;   // TrackKLT ctor arguments (occupancy grid = gx x gy)
  int size_x, size_y, per_cell;                    // Grider_GRID geometry (Grider_GRID.h:88-100)
  int P, C, K;                                     // capacities: points per camera, cells, corners per cell
  int job_base;                                    // FAST job slots [job_base, job_base + 2C): left cells, right cells
  int fast_base, cell_cap;                         // FAST output words: job slot s owns [fast_base + s' * cell_cap, ...)
  int sp_base;                                     // sub-pixel + stereo-LK slots of the left corners [sp_base, sp_base + CK), compact
  int spr_base;                                    // sub-pixel slots of the right corners [spr_base, spr_base + CK), by (cell, k); all left slots of a launch come first
  int lk_base;                                     // temporal-LK slots [lk_base, lk_base + 2P): left, right
  int lkn_base;                                    // temporal-LK slots of this feed's new points [lkn_base, lkn_base + 2CK): left, right
  int rs_base;                                     // RANSAC jobs rs_base (left), rs_base + 1 (right)
  unsigned long long *currid;                      // TrackBase::currid of the tracker (device)
  unsigned long long *h_currid;                    // ... and its host mirror (mapped)
  DvSide s[2];
  int *scr_a, *scr_b, *scr_c, *scr_d;              // [max(P, C * K) + 2] flags / scans
  int *part;                                       // [1024] partial sums of the block scans
  float *tmp_pts;                                  // [2 * P] new right points / right top-up candidates
  unsigned long long *tmp_ids;                     // [P]
  int *uctl;                                       // [0] selected left corners, [1] error flag
  DvObs *obs;                                      // mapped host memory: the update_feature calls of this feed, in order
  DvObs *obs_dev;                                  // ... built here (device memory) and copied out in 16-byte words at the end
  int *n_obs;
  int obs_cap;
};

struct DvHead { int first, count; };

// What changes from feed to feed in a unit's descriptor, packed in one word so that the descriptors themselves stay in device
// memory untouched while the batch keeps its shape: bit 0 first frame, bit 1 publish pts_last / ids_last, then per side (8 bits
// each, left from bit 2): detection slot (2), current slot (2), detection mask present (1), current mask present (1).
constexpr int DV_MAX_UNITS = 256;                 // units per launch (a bigger batch is cut into more groups)
struct DvFeedBits { unsigned v[DV_MAX_UNITS]; };
UVIO_HD unsigned dv_side_bits(int det_slot, int cur_slot, bool det_mask, bool cur_mask) {
  return (unsigned)det_slot | ((unsigned)cur_slot << 2) | ((unsigned)det_mask << 4) | ((unsigned)cur_mask << 5);
}
UVIO_HD void dv_apply_feed(DvUnit &U, unsigned bits) {
  U.first = (int)(bits & 1u);
  U.publish_last = (int)((bits >> 1) & 1u);
  for (int side = 0; side < 2; side++) {
    DvSide &S = U.s[side];
    const unsigned b = bits >> (2 + 8 * side);
    const int det = (int)(b & 3u), cur = (int)((b >> 2) & 3u);
    S.det_pyr = S.pyr_slot[det];
    S.cur_pyr = S.pyr_slot[cur];
    S.last_pyr = S.pyr_slot[(cur + 2) % 3];
    S.det_mask = ((b >> 4) & 1u) ? S.mask_slot[det] : nullptr;
    S.cur_mask = ((b >> 5) & 1u) ? S.mask_slot[cur] : nullptr;
  }
}   // the units of one tracker in this feed, handled by one block in order (they share currid)

struct DvGlobals {
  FastJob *fast_jobs;
  uint32_t *fast_out;
  unsigned *fast_cnt;
  int *fast_active, *fast_n_active;  // job slots that hold a cell to scan, compacted (any order), and their number
  unsigned *feed_bits;            // [units] this feed's DvFeedBits, written by the occupancy kernel for the kernels after it
  int *n_items;                   // [0] sub-pixel items, [1] stereo-LK items, [2] new-point temporal-LK items appended so far (zeroed by the occupancy kernel)
  SubpixItem *sp_items;           // compact lists in any order: every item names the slot its result goes to
  const float *sp_out;            // [2 * slot]
  LkItem *lk1_items;              // stereo LK of new left corners; result slot = the corner's sub-pixel slot
  int *lk1_out;
  const LkOut *lk1_res;
  const uint8_t *lk1_status;
  LkItem *lk2_items;              // temporal LK of the points kept by the occupancy pass (issued right after it)
  LkItem *lk2n_items;             // temporal LK of the points detected in this feed, compact: unit slots [lkn_base, lkn_base + 2CK)
  int *lk2n_out;                  // ... and where each result goes in lk2_out / lk2_status
  const LkOut *lk2_out;
  RansacJob *rs_jobs;
  const float *p1n;
  const uint8_t *rs_mask;         // klt status && RANSAC inlier
};

// Fast scratch private to the block (shared memory on the device): the id / point lists the brute-force membership and box
// tests walk.  Each holds at least P entries.
struct DvSmem {
  unsigned long long *ids0, *ids1;
  float *pts;                   // [2 * P]
};
template <class G>
UVIO_HD void dv_stage_ids(G g, unsigned long long *dst, const unsigned long long *src, int n) {
  for (int i = g.lane(); i < n; i += g.size()) dst[i] = src[i];
  g.sync();
}
template <class G>
UVIO_HD void dv_stage_pts(G g, float *dst, const float *src, int n) {
  for (int i = g.lane(); i < 2 * n; i += g.size()) dst[i] = src[i];
  g.sync();
}

UVIO_HD int dv_resize_nn_src(int d, int dst_len, int src_len) {   // cv::resize INTER_NEAREST source index (TrackKLT.cpp:480)
  double f = (double)dst_len / (double)src_len;
  double inv = 1. / f;
  int s = (int)floor(d * inv);
  return s < src_len - 1 ? s : src_len - 1;
}

// ---- libstdc++ std::sort, first K positions only (see StdSortPrefix in hostlogic.h for why that is exact) -------------------
UVIO_HD bool dv_by_response(uint32_t a, uint32_t b) { return (a >> 24) > (b >> 24); }   // Grider_FAST::compare_response

UVIO_HD void dv_push_heap(uint32_t *first, long hole, long top, uint32_t value) {       // std::__push_heap
  long parent = (hole - 1) / 2;
  while (hole > top && dv_by_response(first[parent], value)) {
    first[hole] = first[parent];
    hole = parent;
    parent = (hole - 1) / 2;
  }
  first[hole] = value;
}
UVIO_HD void dv_adjust_heap(uint32_t *first, long hole, long len, uint32_t value) {     // std::__adjust_heap
  const long top = hole;
  long child = hole;
  while (child < (len - 1) / 2) {
    child = 2 * (child + 1);
    if (dv_by_response(first[child], first[child - 1])) child--;
    first[hole] = first[child];
    hole = child;
  }
  if ((len & 1) == 0 && child == (len - 2) / 2) {
    child = 2 * (child + 1);
    first[hole] = first[child - 1];
    hole = child - 1;
  }
  dv_push_heap(first, hole, top, value);
}
UVIO_HD void dv_heap_sort(uint32_t *first, uint32_t *last) {                            // std::__partial_sort(first, last, last)
  const long len = last - first;
  if (len >= 2) {                                                                       // std::__make_heap
    long parent = (len - 2) / 2;
    while (true) {
      uint32_t value = first[parent];
      dv_adjust_heap(first, parent, len, value);
      if (parent == 0) break;
      parent--;
    }
  }
  while (last - first > 1) {                                                            // std::__sort_heap / __pop_heap
    --last;
    uint32_t value = *last;
    *last = *first;
    dv_adjust_heap(first, 0, last - first, value);
  }
}
UVIO_HD void dv_swap(uint32_t *a, uint32_t *b) { uint32_t t = *a; *a = *b; *b = t; }
UVIO_HD void dv_median_to_first(uint32_t *result, uint32_t *a, uint32_t *b, uint32_t *c) {   // std::__move_median_to_first
  if (dv_by_response(*a, *b)) {
    if (dv_by_response(*b, *c)) dv_swap(result, b);
    else if (dv_by_response(*a, *c)) dv_swap(result, c);
    else dv_swap(result, a);
  } else if (dv_by_response(*a, *c)) dv_swap(result, a);
  else if (dv_by_response(*b, *c)) dv_swap(result, c);
  else dv_swap(result, b);
}
UVIO_HD uint32_t *dv_partition(uint32_t *first, uint32_t *last, uint32_t *pivot) {      // std::__unguarded_partition
  while (true) {
    while (dv_by_response(*first, *pivot)) ++first;
    --last;
    while (dv_by_response(*pivot, *last)) --last;
    if (!(first < last)) return first;
    dv_swap(first, last);
    ++first;
  }
}
UVIO_HD void dv_linear_insert(uint32_t *last) {                                         // std::__unguarded_linear_insert
  uint32_t val = *last;
  uint32_t *next = last - 1;
  while (dv_by_response(val, *next)) {
    *last = *next;
    last = next;
    --next;
  }
  *last = val;
}
// positions [0, k) of a[0..n) end up exactly as std::sort(a, a + n, compare_response) leaves them
UVIO_HD void dv_sort_prefix(uint32_t *a, long n, long k) {
  if (n <= 0) return;
  const long keep = k < n ? k : n;
  long bound = n;
  long lg = 0;
  for (long m = n; m > 1; m >>= 1) lg++;
  // std::__introsort_loop with its recursion (into the right part) unrolled onto an explicit stack
  struct Frame { uint32_t *first, *last; long depth; };
  Frame stack[72];
  int sp = 0;
  stack[sp++] = Frame{a, a + n, lg * 2};
  while (sp > 0) {
    Frame f = stack[--sp];
    uint32_t *first = f.first, *last = f.last;
    long depth = f.depth;
    // the reference recursion handles the right part BEFORE it continues with the left one; parts are disjoint, so the order
    // in which they are processed does not change what each ends up as
    while (last - first > 16) {
      if (depth == 0) {
        dv_heap_sort(first, last);
        break;
      }
      --depth;
      uint32_t *mid = first + (last - first) / 2;
      dv_median_to_first(first, first + 1, mid, last - 1);
      uint32_t *cut = dv_partition(first + 1, last, first);
      if (cut - a < keep) {
        if (sp < 72) stack[sp++] = Frame{cut, last, depth};
      } else if (cut - a < bound) {
        bound = cut - a;
      }
      last = cut;
    }
  }
  // std::__final_insertion_sort: guarded insertion for the first 16 positions, unguarded after them
  const long guarded = n > 16 ? 16 : n;
  for (long i = 1; i < guarded && i < bound; i++) {
    if (dv_by_response(a[i], a[0])) {
      uint32_t val = a[i];
      for (long j = i; j > 0; j--) a[j] = a[j - 1];
      a[0] = val;
    } else {
      dv_linear_insert(a + i);
    }
  }
  for (long i = 16; i < bound; i++) dv_linear_insert(a + i);
}

// bitonic sort of buf[0..m) (m a power of two) by the low 24 bits (y << 12 | x: cv::FAST's row-major emission order)
template <class G>
UVIO_HD void dv_bitonic_by_position(G g, uint32_t *buf, int m) {
  for (int k = 2; k <= m; k <<= 1)
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int i = g.lane(); i < m; i += g.size()) {
        const int l = i ^ j;
        if (l > i) {
          const uint32_t a = buf[i], b = buf[l];
          const bool up = (i & k) == 0;
          if (((a & 0xFFFFFFu) > (b & 0xFFFFFFu)) == up) { buf[i] = b; buf[l] = a; }
        }
      }
      g.sync();
    }
}

// ---- per-cell top-K (Grider_GRID.h:125-141): one block per FAST job slot -------------------------------------------------------
// work: a buffer of at least `work_cap` words private to the block (shared memory when the cell's corners fit, else global)
template <class G>
UVIO_HD void dv_topk(G g, const DvUnit &U, const DvGlobals &X, int side, int r, uint32_t *smem, int smem_cap) {
  const DvSide &S = U.s[side];
  const int slot = U.job_base + side * U.C + r;
  const FastJob job = X.fast_jobs[slot];
  if (job.w == 0 || r >= S.ctl[1] || !S.ctl[2]) {
    if (g.lane() == 0) S.top_n[r] = 0;
    return;
  }
  unsigned n = X.fast_cnt[slot];
  if (n > (unsigned)job.out_cap) n = (unsigned)job.out_cap;
  int m = 1;
  while (m < (int)n) m <<= 1;
  uint32_t *src = X.fast_out + job.out_off;
  uint32_t *buf = m <= smem_cap ? smem : src;            // the job's own region holds the next power of two (cell_cap is one)
  if (buf != src)
    for (int i = g.lane(); i < (int)n; i += g.size()) buf[i] = src[i];
  for (int i = (int)n + g.lane(); i < m; i += g.size()) buf[i] = 0xFFFFFFFFu;
  g.sync();
  dv_bitonic_by_position(g, buf, m);
  if (g.lane() == 0) {
    dv_sort_prefix(buf, (long)n, (long)U.K);
    const int kn = (int)n < U.K ? (int)n : U.K;
    for (int k = 0; k < kn; k++) S.top[r * U.K + k] = buf[k];
    S.top_n[r] = kn;
    X.fast_cnt[slot] = 0;                                // re-armed for the next feed
  }
  g.sync();
}

// ---- occupancy pass of one camera (TrackKLT.cpp:409-464 / :545-600 / :722-780), cell selection and FAST jobs -----------------
// is a pixel inside one of the 255-boxes painted around the first n points of S.pts (mask_updated minus the mask, :457-461)?
UVIO_HD bool dv_boxed(const DvSide &S, const float *pts, int n, int mpd, long at) {
  const int y = (int)(at / S.w), x = (int)(at - (long)y * S.w);
  for (int i = 0; i < n; i++) {
    const int bx = (int)pts[2 * i], by = (int)pts[2 * i + 1];
    if (bx - mpd < 0 || bx + mpd >= S.w || by - mpd < 0 || by + mpd >= S.h) continue;   // the box was not painted (:457)
    const int dx = x - bx, dy = y - by;
    if (dx >= -mpd && dx <= mpd && dy >= -mpd && dy <= mpd) return true;
  }
  return false;
}

// order-independent tests of the occupancy loop for one point; returns close cell + 1 (0 = erased) and the coarse cell
UVIO_HD int dv_occ_base(const DvUnit &U, const DvSide &S, float px, float py, int &gcell) {
  const int cols = S.w, rows = S.h, mpd = U.mpd;
  const int cw = (int)((float)cols / (float)mpd), ch = (int)((float)rows / (float)mpd);
  const float size_x = (float)cols / (float)U.gx, size_y = (float)rows / (float)U.gy;
  const int x = (int)px, y = (int)py;
  const int edge = 10;
  if (x < edge || x >= cols - edge || y < edge || y >= rows - edge) return 0;
  const int x_close = (int)(px / (float)mpd), y_close = (int)(py / (float)mpd);
  if (x_close < 0 || x_close >= cw || y_close < 0 || y_close >= ch) return 0;
  const int x_grid = (int)floorf(px / size_x), y_grid = (int)floorf(py / size_y);
  if (x_grid < 0 || x_grid >= U.gx || y_grid < 0 || y_grid >= U.gy) return 0;
  if (S.det_mask && S.det_mask[(size_t)y * cols + x] > 127) return 0;
  gcell = y_grid * U.gx + x_grid;
  return y_close * cw + x_close + 1;
}

// the cells to extract from (x-outer / y-inner, :483-492) and their FAST jobs; the whole group (lane i <-> cell x = i / gy, y = i % gy)
template <class G>
UVIO_HD void dv_cells_and_jobs(G g, const DvUnit &U, const DvGlobals &X, int side, bool detect) {
  const DvSide &S = U.s[side];
  const int nfg = (int)((double)U.num_features / (double)(U.gx * U.gy)) + 1;
  int req = (int)(0.5 * nfg);
  if (req < 1) req = 1;
  int total = 0;
  for (int base = 0; base < U.C; base += g.size()) {        // C <= group size in practice: one round
    const int i = base + g.lane();
    int valid = 0;
    if (detect && i < U.C) {
      const int x = i / U.gy, y = i - x * U.gy;
      int m = 0;
      if (S.det_mask) m = S.det_mask[(size_t)dv_resize_nn_src(y, U.gy, S.h) * S.w + dv_resize_nn_src(x, U.gx, S.w)];
      int gcount = fdb_ldi(&S.grid[y * U.gx + x]);
      if (gcount > 255) gcount = 255;
      valid = gcount < req && m != 255;
    }
    int round_total;
    const int pos = g.exclusive_sum(valid, round_total);
    if (valid) S.cells[total + pos] = ((i / U.gy) << 16) | (i % U.gy);
    total += round_total;
  }
  g.sync();
  if (g.lane() == 0) {
    S.ctl[1] = total;
    S.ctl[2] = detect;
  }
  for (int r = g.lane(); r < U.C; r += g.size()) {
    FastJob j;
    j.img = S.det_pyr;
    j.x0 = j.y0 = j.w = j.h = 0;
    j.out_off = U.fast_base + (side * U.C + r) * U.cell_cap;
    j.out_cap = U.cell_cap;
    j.threshold = U.fast_threshold;
    if (r < total) {
      const int x = (S.cells[r] >> 16) * U.size_x, y = (S.cells[r] & 0xFFFF) * U.size_y;
      if (!(x + U.size_x > S.w || y + U.size_y > S.h)) {   // Grider_GRID.h:117
        j.x0 = x; j.y0 = y; j.w = U.size_x; j.h = U.size_y;
        const int cap = ((j.w + 1) / 2) * ((j.h + 1) / 2);
        if (cap < j.out_cap) j.out_cap = cap;
        X.fast_active[fdb_add_ret(X.fast_n_active, 1)] = U.job_base + side * U.C + r;   // compact list of the slots FAST must visit (any order)
      }
    }
    X.fast_jobs[U.job_base + side * U.C + r] = j;
  }
  g.sync();
}

template <class G>
UVIO_HD void dv_occupancy_side(G g, const DvUnit &U, const DvGlobals &X, const DvSmem &sm, int side) {
  const DvSide &S = U.s[side];
  const DvSide &L = U.s[0];
  const int mpd = U.mpd;
  const int cw = (int)((float)S.w / (float)mpd), ch = (int)((float)S.h / (float)mpd);
  for (int i = g.lane(); i < cw * ch; i += g.size()) { S.close[i] = 0; S.first[i] = DV_INF; }
  for (int i = g.lane(); i < U.gx * U.gy; i += g.size()) S.grid[i] = 0;
  g.sync();
  const int n = U.first ? 0 : *S.n_last;
  const int nl = side ? L.ctl[0] : 0;                       // right pass: ids of the kept left points are exempt (:754-759)
  if (side) dv_stage_ids(g, sm.ids0, L.ids, nl);
  for (int i = g.lane(); i < n; i += g.size()) {
    int gcell = 0;
    const int c = dv_occ_base(U, S, S.pts_last[2 * i], S.pts_last[2 * i + 1], gcell);
    U.scr_a[i] = c;
    U.scr_b[i] = gcell;
    if (c) fdb_min(&S.first[c - 1], i);
  }
  g.sync();
  for (int i = g.lane(); i < n; i += g.size()) {
    const int c = U.scr_a[i];
    bool keep = false;
    if (c) {
      keep = fdb_ldi(&S.first[c - 1]) == i;                // the first point of its close cell
      if (!keep && side) {                                  // a later one survives only as a stereo id
        const unsigned long long id = S.ids_last[i];
        for (int k = 0; k < nl && !keep; k++) keep = sm.ids0[k] == id;
      }
    }
    U.scr_c[i] = keep;
  }
  g.sync();
  fdb_scan(g, U.scr_c, U.scr_d, n, U.part);
  const int kept = U.scr_d[n];
  for (int i = g.lane(); i < n; i += g.size()) {
    if (!U.scr_c[i]) continue;
    const int d = U.scr_d[i];
    S.pts[2 * d] = S.pts_last[2 * i];
    S.pts[2 * d + 1] = S.pts_last[2 * i + 1];
    S.ids[d] = S.ids_last[i];
    S.close[U.scr_a[i] - 1] = 255;
    fdb_add(&S.grid[U.scr_b[i]], 1);
  }
  g.sync();
  if (g.lane() == 0) {
    S.ctl[0] = kept;
    S.ctl[4] = kept;
  }
  // The temporal LK of the kept points does not depend on the detection: its items are ready now, so the (issue-bound) launch
  // can run beside the (latency-bound) detection chain.  Whether perform_matching runs at all (>= 10 points, :848) is only
  // known after the detection; if it does not, these tracks are simply not used.
  for (int i = g.lane(); i < U.P; i += g.size()) {
    LkItem it{0, 0, 0, 0, -1, -1};
    if (!U.first && i < kept) it = LkItem{S.pts[2 * i], S.pts[2 * i + 1], S.pts[2 * i], S.pts[2 * i + 1], S.last_pyr, S.cur_pyr};
    X.lk2_items[U.lk_base + side * U.P + i] = it;
  }
  const int thr = 20 < (int)(0.5 * U.num_features) ? 20 : (int)(0.5 * U.num_features);
  const int need = U.num_features - kept;
  // mono returns when need < thr (:470); stereo detects when need > thr (:609, :785)
  const bool detect = (U.stereo || side) ? (need > thr) : !(need < thr);
  dv_cells_and_jobs(g, U, X, side, detect);
}

template <class G>
UVIO_HD void dv_occupancy(G g, const DvUnit &U, const DvGlobals &X, const DvSmem &sm) {
  if (g.lane() == 0) { U.uctl[0] = 0; U.uctl[1] = 0; }
  dv_occupancy_side(g, U, X, sm, 0);
  if (U.stereo) {
    dv_occupancy_side(g, U, X, sm, 1);
  } else {                                                  // the right job slots of a mono unit exist but are never used
    for (int r = g.lane(); r < U.C; r += g.size()) {
      FastJob j;
      j.img = U.s[0].det_pyr;
      j.x0 = j.y0 = j.w = j.h = 0;
      j.out_off = U.fast_base;
      j.out_cap = 0;
      j.threshold = U.fast_threshold;
      X.fast_jobs[U.job_base + U.C + r] = j;
    }
    for (int i = g.lane(); i < U.P; i += g.size()) X.lk2_items[U.lk_base + U.P + i] = LkItem{0, 0, 0, 0, -1, -1};
    g.sync();
  }
}

// ---- selection: left mask test against mask_updated (Grider_GRID.h:142-148), sub-pixel / stereo-LK items ---------------------
UVIO_HD void dv_corner_xy(const DvUnit &U, const DvSide &S, int r, uint32_t w, float &x, float &y) {
  x = (float)(w & 4095u) + (float)((S.cells[r] >> 16) * U.size_x);
  y = (float)((w >> 12) & 4095u) + (float)((S.cells[r] & 0xFFFF) * U.size_y);
}

template <class G>
UVIO_HD void dv_select(G g, const DvUnit &U, const DvGlobals &X, const DvSmem &sm) {
  const DvSide &L = U.s[0], &R = U.s[1];
  const int CK = U.C * U.K;
  const int n_old = L.ctl[4];
  dv_stage_pts(g, sm.pts, L.pts, L.ctl[2] ? n_old : 0);
  for (int idx = g.lane(); idx < CK; idx += g.size()) {
    const int r = idx / U.K, k = idx - r * U.K;
    int ok = 0;
    if (L.ctl[2] && r < L.ctl[1] && k < L.top_n[r]) {
      float fx, fy;
      dv_corner_xy(U, L, r, L.top[idx], fx, fy);
      const int x = (int)fx, y = (int)fy;
      if (!(x < 0 || x > L.w || y < 0 || y > L.h)) {
        const long at = (long)y * L.w + x;
        ok = !((L.det_mask && L.det_mask[at] > 127) || dv_boxed(L, sm.pts, n_old, U.mpd, at));
      }
    }
    U.scr_a[idx] = ok;
  }
  g.sync();
  fdb_scan(g, U.scr_a, U.scr_b, CK, U.part);
  const int nsel = U.scr_b[CK];
  for (int idx = g.lane(); idx < CK; idx += g.size()) {
    const int r = idx / U.K, k = idx - r * U.K;
    if (U.scr_a[idx]) {
      const int j = U.scr_b[idx];
      float fx, fy;
      dv_corner_xy(U, L, r, L.top[idx], fx, fy);
      X.sp_items[fdb_add_ret(&X.n_items[0], 1)] = SubpixItem{fx, fy, L.det_pyr, U.sp_base + j};
      if (U.stereo) {   // left -> right LK of the refined corner (TrackKLT.cpp:661-672): starts from the sub-pixel result of the same slot
        const int q = fdb_add_ret(&X.n_items[1], 1);
        X.lk1_items[q] = LkItem{0, 0, 0, 0, L.det_pyr, R.det_pyr};
        X.lk1_out[q] = U.sp_base + j;
      }
    }
    // right: every top-K corner of every cell FAST ran on is refined (a superset of what the right top-up will take)
    if (U.stereo && R.ctl[2] && r < R.ctl[1] && k < R.top_n[r]) {
      SubpixItem it{0, 0, R.det_pyr, U.spr_base + idx};
      dv_corner_xy(U, R, r, R.top[idx], it.x, it.y);
      X.sp_items[fdb_add_ret(&X.n_items[0], 1)] = it;
    }
  }
  if (g.lane() == 0) U.uctl[0] = nsel;
  g.sync();
}

// ---- min-distance reject of refined corners (TrackKLT.cpp:497-512 / :635-650 / :811-825) ---------------------------------------
// pts[0..n): candidates in order; flags out[i] = kept; marks S.close.  Uses scr_a (cells) and S.first.
template <class G>
UVIO_HD void dv_close_reject(G g, const DvUnit &U, const DvSide &S, const float *pts, int n, int *keep) {
  const int mpd = U.mpd;
  const int cw = (int)((float)S.w / (float)mpd), ch = (int)((float)S.h / (float)mpd);
  for (int i = g.lane(); i < cw * ch; i += g.size()) S.first[i] = DV_INF;
  g.sync();
  for (int i = g.lane(); i < n; i += g.size()) {
    const int xg = (int)(pts[2 * i] / (float)mpd), yg = (int)(pts[2 * i + 1] / (float)mpd);
    int c = 0;
    if (!(xg < 0 || xg >= cw || yg < 0 || yg >= ch) && !(S.close[yg * cw + xg] > 127)) {
      c = yg * cw + xg + 1;
      fdb_min(&S.first[c - 1], i);
    }
    U.scr_a[i] = c;
  }
  g.sync();
  for (int i = g.lane(); i < n; i += g.size()) {
    const int c = U.scr_a[i];
    keep[i] = c && fdb_ldi(&S.first[c - 1]) == i;
  }
  g.sync();
  for (int i = g.lane(); i < n; i += g.size())
    if (keep[i]) S.close[U.scr_a[i] - 1] = 255;
  g.sync();
}

// ---- finish the detection (phase C of the host path) and emit the temporal LK / RANSAC work -------------------------------------
template <class G>
UVIO_HD void dv_assign(G g, const DvUnit &U, const DvGlobals &X, const DvSmem &sm) {
  const DvSide &L = U.s[0], &R = U.s[1];
  const int CK = U.C * U.K;
  const int nsel = U.uctl[0];
  const float *refined = X.sp_out + 2 * (size_t)U.sp_base;
  const int thr = 20 < (int)(0.5 * U.num_features) ? 20 : (int)(0.5 * U.num_features);
  unsigned long long currid = *U.currid;
  int nL = L.ctl[0];
  g.sync();
  dv_close_reject(g, U, L, refined, nsel, U.scr_b);                         // scr_b = keep
  if (!U.stereo) {
    fdb_scan(g, U.scr_b, U.scr_c, nsel, U.part);                            // :519-527
    const int total = U.scr_c[nsel];
    for (int j = g.lane(); j < nsel; j += g.size())
      if (U.scr_b[j]) {
        const int d = nL + U.scr_c[j];
        L.pts[2 * d] = refined[2 * j];
        L.pts[2 * d + 1] = refined[2 * j + 1];
        L.ids[d] = currid + 1 + (unsigned long long)U.scr_c[j];
      }
    nL += total;
    currid += (unsigned long long)total;
    g.sync();
  } else {
    int nR = R.ctl[0];
    // :675-708: a kept corner becomes a stereo point (one id, both cameras), a left-only point, or nothing
    for (int j = g.lane(); j < nsel; j += g.size()) {
      int cls = 0;
      if (U.scr_b[j]) {
        const float x0 = refined[2 * j], y0 = refined[2 * j + 1];
        const LkOut o = X.lk1_res[U.sp_base + j];
        const bool oob_left = (int)x0 < 0 || (int)x0 >= L.w || (int)y0 < 0 || (int)y0 >= L.h;
        const bool oob_right = (int)o.x < 0 || (int)o.x >= R.w || (int)o.y < 0 || (int)o.y >= R.h;
        if (!oob_left && !oob_right && X.lk1_status[U.sp_base + j] == 1) cls = 2;
        else if (!oob_left) cls = 1;
      }
      U.scr_a[j] = cls > 0;
      U.scr_d[j] = cls == 2;
    }
    g.sync();
    fdb_scan(g, U.scr_a, U.scr_c, nsel, U.part);                            // order among the points that get an id
    const int totalL = U.scr_c[nsel];
    for (int j = g.lane(); j < nsel; j += g.size())
      if (U.scr_a[j]) {
        const int d = nL + U.scr_c[j];
        L.pts[2 * d] = refined[2 * j];
        L.pts[2 * d + 1] = refined[2 * j + 1];
        L.ids[d] = currid + 1 + (unsigned long long)U.scr_c[j];
      }
    g.sync();
    fdb_scan(g, U.scr_d, U.scr_b, nsel, U.part);                            // order among the stereo points
    const int totalR = U.scr_b[nsel];
    for (int j = g.lane(); j < nsel; j += g.size())
      if (U.scr_d[j]) {
        const LkOut o = X.lk1_res[U.sp_base + j];
        const int d = U.scr_b[j];
        U.tmp_pts[2 * d] = o.x;
        U.tmp_pts[2 * d + 1] = o.y;
        U.tmp_ids[d] = currid + 1 + (unsigned long long)U.scr_c[j];
      }
    nL += totalL;
    currid += (unsigned long long)totalL;
    g.sync();
    // right: the occupancy loop continues over the appended stereo points (:722-780); their ids are all in ids0, so the
    // closeness erase never applies to them
    for (int q = g.lane(); q < totalR; q += g.size()) {
      int gcell = 0;
      const int c = dv_occ_base(U, R, U.tmp_pts[2 * q], U.tmp_pts[2 * q + 1], gcell);
      U.scr_a[q] = c;
      U.scr_d[q] = gcell;
      U.scr_c[q] = c != 0;
    }
    g.sync();
    fdb_scan(g, U.scr_c, U.scr_b, totalR, U.part);
    const int keptR = U.scr_b[totalR];
    for (int q = g.lane(); q < totalR; q += g.size())
      if (U.scr_c[q]) {
        const int d = nR + U.scr_b[q];
        R.pts[2 * d] = U.tmp_pts[2 * q];
        R.pts[2 * d + 1] = U.tmp_pts[2 * q + 1];
        R.ids[d] = U.tmp_ids[q];
        R.close[U.scr_a[q] - 1] = 255;
        fdb_add(&R.grid[U.scr_d[q]], 1);
      }
    nR += keptR;
    g.sync();
    // right top-up (:784-826) from the cells that are valid NOW, a subset of the cells FAST ran on
    const int need = U.num_features - nR;
    if (need > thr) {
      // scr_d[0..nv): row in R.cells of every valid cell, in x-outer / y-inner order (one lane: <= C cells)
      if (g.lane() == 0) {
        const int nfg = (int)((double)U.num_features / (double)(U.gx * U.gy)) + 1;
        int req = (int)(0.5 * nfg);
        if (req < 1) req = 1;
        int nv = 0;
        for (int x = 0; x < U.gx; x++)
          for (int y = 0; y < U.gy; y++) {
            int m = 0;
            if (R.det_mask) m = R.det_mask[(size_t)dv_resize_nn_src(y, U.gy, R.h) * R.w + dv_resize_nn_src(x, U.gx, R.w)];
            int gcount = fdb_ldi(&R.grid[y * U.gx + x]);
            if (gcount > 255) gcount = 255;
            if (gcount < req && m != 255) {
              int r = 0;
              while (r < R.ctl[1] && R.cells[r] != ((x << 16) | y)) r++;
              if (r == R.ctl[1] || !R.ctl[2]) U.uctl[1] = 1;   // cannot happen: R.cells is a superset
              U.scr_d[U.C * U.K + 1 + nv++] = r;
            }
          }
        U.scr_d[U.C * U.K] = nv;
      }
      g.sync();
      const int nv = U.scr_d[CK];
      dv_stage_pts(g, sm.pts, R.pts, nR);
      const uint8_t *mask0 = (L.w == R.w && L.h == R.h) ? L.det_mask : nullptr;   // :721 clones mask0 (sic)
      for (int idx = g.lane(); idx < nv * U.K; idx += g.size()) {
        const int vi = idx / U.K, k = idx - vi * U.K;
        const int r = U.scr_d[CK + 1 + vi];
        int ok = 0;
        if (r < R.ctl[1] && k < R.top_n[r]) {
          float fx, fy;
          dv_corner_xy(U, R, r, R.top[r * U.K + k], fx, fy);
          const int x = (int)fx, y = (int)fy;
          if (!(x < 0 || x > R.w || y < 0 || y > R.h)) {
            const long at = (long)y * R.w + x;
            ok = !((mask0 && mask0[at] > 127) || dv_boxed(R, sm.pts, nR, U.mpd, at));
          }
        }
        U.scr_c[idx] = ok;
      }
      g.sync();
      fdb_scan(g, U.scr_c, U.scr_b, nv * U.K, U.part);
      const int next = U.scr_b[nv * U.K];
      for (int idx = g.lane(); idx < nv * U.K; idx += g.size())
        if (U.scr_c[idx]) {
          const int vi = idx / U.K, k = idx - vi * U.K;
          const int r = U.scr_d[CK + 1 + vi];
          const int sp = U.spr_base + r * U.K + k;
          const int d = U.scr_b[idx];
          U.tmp_pts[2 * d] = X.sp_out[2 * (size_t)sp];
          U.tmp_pts[2 * d + 1] = X.sp_out[2 * (size_t)sp + 1];
        }
      g.sync();
      dv_close_reject(g, U, R, U.tmp_pts, next, U.scr_b);                   // :811-825
      fdb_scan(g, U.scr_b, U.scr_c, next, U.part);
      const int total = U.scr_c[next];
      for (int e = g.lane(); e < next; e += g.size())
        if (U.scr_b[e]) {
          const int d = nR + U.scr_c[e];
          R.pts[2 * d] = U.tmp_pts[2 * e];
          R.pts[2 * d + 1] = U.tmp_pts[2 * e + 1];
          R.ids[d] = currid + 1 + (unsigned long long)U.scr_c[e];
        }
      nR += total;
      currid += (unsigned long long)total;
      g.sync();
    }
    if (g.lane() == 0) R.ctl[0] = nR;
  }
  if (g.lane() == 0) {
    L.ctl[0] = nL;
    *U.currid = currid;
    *U.h_currid = currid;
  }
  g.sync();
  // first frame of a unit: store, no tracking, no database update (TrackKLT.cpp:110-123 / :221-240); else the temporal LK
  for (int side = 0; side < (U.stereo ? 2 : 1); side++) {
    const DvSide &S = U.s[side];
    const int n = S.ctl[0];
    const bool run = !U.first && n >= 10;                                   // :848-852
    const int kept_old = S.ctl[4];
    if (run)
      for (int i = kept_old + g.lane(); i < n; i += g.size()) {             // the points detected in this feed
        const int q = fdb_add_ret(&X.n_items[2], 1);
        X.lk2n_items[q] = LkItem{S.pts[2 * i], S.pts[2 * i + 1], S.pts[2 * i], S.pts[2 * i + 1], S.last_pyr, S.cur_pyr};
        X.lk2n_out[q] = U.lk_base + side * U.P + i;
      }
    if (run)                                                                // RANSAC undistorts the old positions from the item list
      for (int i = kept_old + g.lane(); i < n; i += g.size())
        X.lk2_items[U.lk_base + side * U.P + i] = LkItem{S.pts[2 * i], S.pts[2 * i + 1], S.pts[2 * i], S.pts[2 * i + 1], -1, -1};
    if (g.lane() == 0) {
      S.ctl[3] = run;
      RansacJob j;
      j.off = U.lk_base + side * U.P;
      j.n = run ? n : 0;
      j.cam0 = j.cam1 = S.cam_model;
      j.thr = S.thr;
      j.run = run;
      X.rs_jobs[U.rs_base + side] = j;
    }
    if (U.first) {
      for (int i = g.lane(); i < n; i += g.size()) {
        S.pts_last[2 * i] = S.pts[2 * i];
        S.pts_last[2 * i + 1] = S.pts[2 * i + 1];
        S.ids_last[i] = S.ids[i];
      }
      if (g.lane() == 0) *S.n_last = n;
      g.sync();
      dv_publish_side(g, U, S, n);
    }
  }
  if (!U.stereo && g.lane() == 0) {                                         // the unused right slots of a mono unit
    RansacJob j;
    j.off = U.lk_base + U.P;
    j.n = 0;
    j.cam0 = j.cam1 = L.cam_model;
    j.thr = L.thr;
    j.run = 0;
    X.rs_jobs[U.rs_base + 1] = j;
  }
  if (U.first && g.lane() == 0) *U.n_obs = 0;
  g.sync();
}

// Results leave for mapped host memory as one linear copy in 16-byte words (full PCIe write bursts), never as scattered stores.
template <class G>
UVIO_HD void dv_copy_out(G g, void *dst, const void *src, size_t bytes) {
  const size_t n16 = (bytes + 15) / 16;                     // the arrays are 256-byte aligned and padded
  Vec16 *d = (Vec16 *)dst;
  const Vec16 *q = (const Vec16 *)src;
  for (size_t i = g.lane(); i < n16; i += g.size()) d[i] = q[i];
}
template <class G>
UVIO_HD void dv_publish_side(G g, const DvUnit &U, const DvSide &S, int n) {
  if (U.publish_last) {
    dv_copy_out(g, S.h_pts_last, S.pts_last, sizeof(float) * 2 * (size_t)n);
    dv_copy_out(g, S.h_ids_last, S.ids_last, sizeof(unsigned long long) * (size_t)n);
  }
  if (g.lane() == 0) *S.h_n_last = n;
}

// ---- merge + database observations + pts_last / ids_last (TrackKLT.cpp:143-189, :285-381) -------------------------------------
template <class G>
UVIO_HD void dv_store_last(G g, const DvSide &S, const float *pts, const unsigned long long *ids, const int *keep, const int *pos, int n, int base) {
  for (int i = g.lane(); i < n; i += g.size())
    if (keep[i]) {
      const int d = base + pos[i];
      S.pts_last[2 * d] = pts[2 * i];
      S.pts_last[2 * d + 1] = pts[2 * i + 1];
      S.ids_last[d] = ids[i];
    }
}

template <class G>
UVIO_HD void dv_merge(G g, const DvUnit &U, const DvGlobals &X, const DvSmem &sm) {
  if (U.first) return;                                                      // stored by dv_assign
  const DvSide &L = U.s[0], &R = U.s[1];
  const int nL = L.ctl[0], offL = U.lk_base, offR = U.lk_base + U.P;
  const bool runL = L.ctl[3] != 0;
  g.sync();
  if (!U.stereo) {
    // :159-173
    for (int i = g.lane(); i < nL; i += g.size()) {
      int keep = 0;
      if (runL) {
        const LkOut p = X.lk2_out[offL + i];
        if (!(p.x < 0 || p.y < 0 || (int)p.x >= L.w || (int)p.y >= L.h) &&
            !(L.cur_mask && L.cur_mask[(size_t)(int)p.y * L.w + (int)p.x] > 127) && X.rs_mask[offL + i])
          keep = 1;
      }
      U.scr_a[i] = keep;
    }
    g.sync();
    fdb_scan(g, U.scr_a, U.scr_b, nL, U.part);
    const int total = U.scr_b[nL];
    for (int i = g.lane(); i < nL; i += g.size())
      if (U.scr_a[i]) {
        const int d = U.scr_b[i];
        const LkOut p = X.lk2_out[offL + i];
        L.pts_last[2 * d] = p.x;
        L.pts_last[2 * d + 1] = p.y;
        L.ids_last[d] = L.ids[i];
        if (d < U.obs_cap) U.obs_dev[d] = DvObs{L.ids[i], L.cam_id, p.x, p.y, X.p1n[2 * (size_t)(offL + i)], X.p1n[2 * (size_t)(offL + i) + 1]};
      }
    if (g.lane() == 0) *L.n_last = total;
    g.sync();
    dv_publish_side(g, U, L, total);
    dv_copy_out(g, U.obs, U.obs_dev, sizeof(DvObs) * (size_t)(total < U.obs_cap ? total : U.obs_cap));
    if (g.lane() == 0) *U.n_obs = total;
    g.sync();
    return;
  }
  const int nR = R.ctl[0];
  const bool runR = R.ctl[3] != 0;
  dv_stage_ids(g, sm.ids0, R.ids, nR);
  // left loop (:306-338).  scr_a = left kept, scr_c = index of the paired right point or -1
  for (int i = g.lane(); i < nL; i += g.size()) {
    int keep = 0, pair = -1;
    if (runL) {
      const LkOut p = X.lk2_out[offL + i];
      if (!(p.x < 0 || p.y < 0 || (int)p.x > L.w || (int)p.y > L.h)) {       // '>' on the left (:308-309)
        const unsigned long long id = L.ids[i];
        int ir = -1;                                                         // first index in ids_right_old (:314-320)
        for (int k = 0; k < nR; k++)
          if (sm.ids0[k] == id) { ir = k; break; }
        const bool okL = X.rs_mask[offL + i] != 0;
        if (okL && ir >= 0 && runR && X.rs_mask[offR + ir]) {
          const LkOut q = X.lk2_out[offR + ir];
          if (!(q.x < 0 || q.y < 0 || (int)q.x >= R.w || (int)q.y >= R.h)) { keep = 1; pair = ir; }
        } else if (okL) {
          keep = 1;
        }
      }
    }
    U.scr_a[i] = keep;
    U.scr_c[i] = pair;
  }
  g.sync();
  fdb_scan(g, U.scr_a, U.scr_b, nL, U.part);
  const int totL = U.scr_b[nL];
  for (int i = g.lane(); i < nL; i += g.size())
    if (U.scr_a[i]) {
      const int d = U.scr_b[i];
      const LkOut p = X.lk2_out[offL + i];
      L.pts_last[2 * d] = p.x;
      L.pts_last[2 * d + 1] = p.y;
      L.ids_last[d] = L.ids[i];
      if (d < U.obs_cap) U.obs_dev[d] = DvObs{L.ids[i], L.cam_id, p.x, p.y, X.p1n[2 * (size_t)(offL + i)], X.p1n[2 * (size_t)(offL + i) + 1]};
    }
  g.sync();
  // right points paired with a left one come first, in the left loop's order
  for (int i = g.lane(); i < nL; i += g.size()) U.scr_d[i] = U.scr_c[i] >= 0;
  g.sync();
  fdb_scan(g, U.scr_d, U.scr_b, nL, U.part);
  const int nPair = U.scr_b[nL];
  for (int i = g.lane(); i < nL; i += g.size())
    if (U.scr_d[i]) {
      const int d = U.scr_b[i], ir = U.scr_c[i];
      const LkOut q = X.lk2_out[offR + ir];
      R.pts_last[2 * d] = q.x;
      R.pts_last[2 * d + 1] = q.y;
      R.ids_last[d] = R.ids[ir];
      if (totL + d < U.obs_cap)
        U.obs_dev[totL + d] = DvObs{R.ids[ir], R.cam_id, q.x, q.y, X.p1n[2 * (size_t)(offR + ir)], X.p1n[2 * (size_t)(offR + ir) + 1]};
    }
  g.sync();
  // right loop (:341-354): tracked right points whose id is not in good_ids_right yet.  scr_a = kept
  for (int l = g.lane(); l < nL; l += g.size()) sm.ids1[l] = U.scr_c[l] >= 0 ? sm.ids0[U.scr_c[l]] : ~0ull;   // ids paired by the left loop
  g.sync();
  for (int i = g.lane(); i < nR; i += g.size()) {
    int keep = 0;
    if (runR) {
      const LkOut q = X.lk2_out[offR + i];
      if (!(q.x < 0 || q.y < 0 || (int)q.x >= R.w || (int)q.y >= R.h) && X.rs_mask[offR + i]) {
        const unsigned long long id = sm.ids0[i];
        bool added = false;
        for (int l = 0; l < nL && !added; l++) added = sm.ids1[l] == id;
        for (int k = 0; k < i && !added; k++) added = sm.ids0[k] == id && X.rs_mask[offR + k];   // a duplicate id earlier in this loop
        keep = !added;
      }
    }
    U.scr_a[i] = keep;
  }
  g.sync();
  fdb_scan(g, U.scr_a, U.scr_b, nR, U.part);
  const int nSolo = U.scr_b[nR];
  for (int i = g.lane(); i < nR; i += g.size())
    if (U.scr_a[i]) {
      const int d = nPair + U.scr_b[i];
      const LkOut q = X.lk2_out[offR + i];
      R.pts_last[2 * d] = q.x;
      R.pts_last[2 * d + 1] = q.y;
      R.ids_last[d] = R.ids[i];
      if (totL + d < U.obs_cap)
        U.obs_dev[totL + d] = DvObs{R.ids[i], R.cam_id, q.x, q.y, X.p1n[2 * (size_t)(offR + i)], X.p1n[2 * (size_t)(offR + i) + 1]};
    }
  if (g.lane() == 0) {
    *L.n_last = totL;
    *R.n_last = nPair + nSolo;
  }
  g.sync();
  const int n_obs = totL + nPair + nSolo;
  dv_publish_side(g, U, L, totL);
  dv_publish_side(g, U, R, nPair + nSolo);
  dv_copy_out(g, U.obs, U.obs_dev, sizeof(DvObs) * (size_t)(n_obs < U.obs_cap ? n_obs : U.obs_cap));
  if (g.lane() == 0) *U.n_obs = n_obs;
  g.sync();
}

}  // namespace uvio
