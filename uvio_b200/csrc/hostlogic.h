// The reference's sequential, order-dependent bookkeeping, restated for the GPU tracker's host side.  No CUDA here.
//
//   occupancy pass          TrackKLT.cpp:409-464 (mono), :545-600 (stereo left), :722-780 (stereo right)
//   cell selection          TrackKLT.cpp:476-492, :611-630, :787-806
//   per-cell top-K + mask   Grider_GRID.h:84-156  (std::sort with Grider_FAST::compare_response, Grider_FAST.h:57)
//   min-distance reject     TrackKLT.cpp:497-512, :635-650, :811-825
#pragma once
#include <math.h>
#include <stdint.h>
#include <string.h>

#include <algorithm>
#include <utility>
#include <vector>

#include "engine.h"

namespace uvio {

// Small open-addressing table id -> first index it was inserted with (the id sets / maps of TrackKLT.cpp:314-320, :754-759).
struct IdTable {
  std::vector<size_t> keys;
  std::vector<int> vals;   // -1 = empty slot
  size_t mask = 0;
  void reset(size_t n) {
    size_t cap = 16;
    while (cap < 2 * n + 2) cap <<= 1;
    keys.resize(cap);
    vals.assign(cap, -1);
    mask = cap - 1;
    used = 0;
  }
  static size_t mix(size_t k) {
    k *= 0x9E3779B97F4A7C15ull;
    return k ^ (k >> 29);
  }
  void grow() {
    std::vector<size_t> ok;
    std::vector<int> ov;
    ok.swap(keys);
    ov.swap(vals);
    reset(ok.size());
    for (size_t i = 0; i < ok.size(); i++)
      if (ov[i] >= 0) insert(ok[i], ov[i]);
  }
  size_t used = 0;
  // keeps the value of the first insertion of a key; returns true when the key was new
  bool insert(size_t key, int val) {
    if (2 * (used + 1) > mask + 1) {
      grow();
    }
    size_t at = mix(key) & mask;
    while (vals[at] >= 0) {
      if (keys[at] == key) return false;
      at = (at + 1) & mask;
    }
    keys[at] = key;
    vals[at] = val;
    used++;
    return true;
  }
  int find(size_t key) const {   // -1 when absent
    if (vals.empty()) return -1;
    size_t at = mix(key) & mask;
    while (vals[at] >= 0) {
      if (keys[at] == key) return vals[at];
      at = (at + 1) & mask;
    }
    return -1;
  }
  void fill(const size_t *ids, size_t n) {
    reset(n);
    for (size_t i = 0; i < n; i++) insert(ids[i], (int)i);
  }
};

struct OccState {
  int cols = 0, rows = 0, mpd = 0, gx = 0, gy = 0;
  int cw = 0, ch = 0;                 // size_close
  float size_x = 0, size_y = 0;       // float cell size of the coarse grid (TrackKLT.cpp:404-405)
  std::vector<uint8_t> close, grid;
  std::vector<Pt> pts;                // kept points, in order
  std::vector<size_t> ids;
  std::vector<std::pair<int, int>> boxes;     // centres of the (2*mpd+1)^2 boxes of mask_updated (:457-461)
  // the boxes binned on an mpd-sized grid (built on demand): a pixel can only lie in boxes centred in the 3x3 bins around it
  std::vector<int> bin_head, box_next;
  int bw = 0, bh = 0;
  size_t binned = 0;
};

inline void occ_init(OccState &o, int cols, int rows, int mpd, int gx, int gy) {
  o.cols = cols; o.rows = rows; o.mpd = mpd; o.gx = gx; o.gy = gy;
  o.cw = (int)((float)cols / (float)mpd);
  o.ch = (int)((float)rows / (float)mpd);
  o.size_x = (float)cols / (float)gx;
  o.size_y = (float)rows / (float)gy;
  o.close.assign((size_t)o.cw * o.ch, 0);
  o.grid.assign((size_t)gx * gy, 0);
  o.pts.clear();
  o.ids.clear();
  o.boxes.clear();
  o.binned = 0;
  o.bin_head.clear();
}

// Runs the erase-while-iterating loop over n more points (kept ones are appended to o.pts / o.ids).
// mask: the image mask tested at :446/:762 (null = all zeros).  The 255-boxes of mask_updated (:457-461) are only recorded
// here; occ_boxed() answers the mask_updated lookups of a detection from the recorded boxes (most frames need none).
// stereo_ids: right pass only, ids exempt from the closeness erase (:754-759).
inline void occ_feed(OccState &o, const Pt *pts, const size_t *ids, size_t n, const uint8_t *mask,
                     const IdTable *stereo_ids) {
  const int cols = o.cols, rows = o.rows, mpd = o.mpd;
  for (size_t i = 0; i < n; i++) {
    const float px = pts[i].x, py = pts[i].y;
    const int x = (int)px, y = (int)py;
    const int edge = 10;
    if (x < edge || x >= cols - edge || y < edge || y >= rows - edge) continue;
    const int x_close = (int)(px / (float)mpd), y_close = (int)(py / (float)mpd);
    if (x_close < 0 || x_close >= o.cw || y_close < 0 || y_close >= o.ch) continue;
    const int x_grid = (int)floorf(px / o.size_x), y_grid = (int)floorf(py / o.size_y);
    if (x_grid < 0 || x_grid >= o.gx || y_grid < 0 || y_grid >= o.gy) continue;
    const bool occupied = o.close[(size_t)y_close * o.cw + x_close] > 127;
    if (stereo_ids) {
      if (occupied && stereo_ids->find(ids[i]) < 0) continue;
    } else if (occupied) {
      continue;
    }
    if (mask && mask[(size_t)y * cols + x] > 127) continue;
    o.close[(size_t)y_close * o.cw + x_close] = 255;
    uint8_t &g = o.grid[(size_t)y_grid * o.gx + x_grid];
    if (g < 255) g += 1;
    if (x - mpd >= 0 && x + mpd < cols && y - mpd >= 0 && y + mpd < rows) o.boxes.emplace_back(x, y);
    o.pts.push_back(pts[i]);
    o.ids.push_back(ids[i]);
  }
}

// mask_updated minus the original mask: is pixel `at` (linear index y*cols + x, as the reference's .at<uint8_t>(y, x)
// resolves it, including x == cols) inside one of the 255-boxes recorded so far?
inline bool occ_boxed(OccState &o, size_t at) {
  if (o.boxes.empty()) return false;
  const int mpd = o.mpd;
  if (o.bin_head.empty()) {
    o.bw = o.cols / mpd + 1;
    o.bh = o.rows / mpd + 2;
    o.bin_head.assign((size_t)o.bw * o.bh, -1);
    o.box_next.clear();
    o.binned = 0;
  }
  for (; o.binned < o.boxes.size(); o.binned++) {
    int b = (o.boxes[o.binned].second / mpd) * o.bw + o.boxes[o.binned].first / mpd;
    o.box_next.push_back(o.bin_head[b]);
    o.bin_head[b] = (int)o.binned;
  }
  const int y = (int)(at / (size_t)o.cols), x = (int)(at - (size_t)y * o.cols);
  const int bx = x / mpd, by = y / mpd;
  for (int yy = by - 1; yy <= by + 1; yy++) {
    if (yy < 0 || yy >= o.bh) continue;
    for (int xx = bx - 1; xx <= bx + 1; xx++) {
      if (xx < 0 || xx >= o.bw) continue;
      for (int i = o.bin_head[(size_t)yy * o.bw + xx]; i >= 0; i = o.box_next[i]) {
        int dx = x - o.boxes[i].first, dy = y - o.boxes[i].second;
        if (dx >= -mpd && dx <= mpd && dy >= -mpd && dy <= mpd) return true;
      }
    }
  }
  return false;
}

// cv::resize(mask, mask_grid, Size(gx,gy), 0, 0, INTER_NEAREST) sample position along one axis
inline int resize_nn_src(int d, int dst_len, int src_len) {
  double f = (double)dst_len / (double)src_len;
  double inv = 1. / f;
  int s = (int)floor(d * inv);
  return s < src_len - 1 ? s : src_len - 1;
}

// Cells to extract from, x-outer / y-inner (TrackKLT.cpp:483-492)
inline void valid_cells(const OccState &o, const uint8_t *mask, int num_features, std::vector<std::pair<int, int>> &out) {
  out.clear();
  const int nfg = (int)((double)num_features / (double)(o.gx * o.gy)) + 1;
  int req = (int)(0.5 * nfg);
  if (req < 1) req = 1;
  for (int x = 0; x < o.gx; x++)
    for (int y = 0; y < o.gy; y++) {
      int m = 0;
      if (mask) m = mask[(size_t)resize_nn_src(y, o.gy, o.rows) * o.cols + resize_nn_src(x, o.gx, o.cols)];
      if ((int)o.grid[(size_t)y * o.gx + x] < req && m != 255) out.emplace_back(x, y);
    }
}

// Geometry Grider_GRID::perform_griding derives from (num_features, grid) (Grider_GRID.h:88-100)
struct GridGeom {
  int gx, gy, size_x, size_y, per_cell;
};
inline GridGeom grider_geom(int cols, int rows, int num_features, int grid_x, int grid_y) {
  GridGeom g;
  if (num_features < grid_x * grid_y) {
    double ratio = (double)grid_x / (double)grid_y;
    grid_y = (int)ceil(sqrt(num_features / ratio));
    grid_x = (int)ceil(grid_y * ratio);
  }
  g.gx = grid_x;
  g.gy = grid_y;
  g.per_cell = (int)((double)num_features / (double)(grid_x * grid_y)) + 1;
  g.size_x = cols / grid_x;
  g.size_y = rows / grid_y;
  return g;
}

struct KeyPointLite {   // stands in for cv::KeyPoint; passed by value to the comparator like the reference does
  float x, y, size, angle, response;
  int octave, class_id;
};
inline bool compare_response(KeyPointLite first, KeyPointLite second) { return first.response > second.response; }

// The first K positions of what libstdc++'s std::sort(first, last, comp) leaves behind, without sorting the rest.
//
// std::sort is introsort: a loop that partitions [first,last) around a median-of-three pivot, recurses into the right part
// and continues with the left one, followed by one insertion-sort sweep (bits/stl_algo.h: __introsort_loop,
// __unguarded_partition_pivot, __final_insertion_sort).  Parts never exchange elements, and after a partition no element
// of the right part compares before an element of the left part, so the insertion sweep never carries an element of a
// part that starts at position >= K into the first K positions.  Skipping those parts (their recursion and their share
// of the sweep) therefore leaves the first K positions exactly as the full sort does, ties included.  The reference's
// per-cell top-K (Grider_GRID.h:128-141) reads only those.
template <class Comp>
struct StdSortPrefix {
  uint32_t *base;
  long keep;      // positions [0, keep) must end up as std::sort leaves them
  long bound;     // first position the insertion sweep may skip from
  Comp comp;
  static long lg(long n) { long k = 0; while (n > 1) { n >>= 1; k++; } return k; }
  void median_to_first(uint32_t *result, uint32_t *a, uint32_t *b, uint32_t *c) {
    if (comp(*a, *b)) {
      if (comp(*b, *c)) std::iter_swap(result, b);
      else if (comp(*a, *c)) std::iter_swap(result, c);
      else std::iter_swap(result, a);
    } else if (comp(*a, *c)) std::iter_swap(result, a);
    else if (comp(*b, *c)) std::iter_swap(result, c);
    else std::iter_swap(result, b);
  }
  uint32_t *partition(uint32_t *first, uint32_t *last, uint32_t *pivot) {
    while (true) {
      while (comp(*first, *pivot)) ++first;
      --last;
      while (comp(*pivot, *last)) --last;
      if (!(first < last)) return first;
      std::iter_swap(first, last);
      ++first;
    }
  }
  void loop(uint32_t *first, uint32_t *last, long depth_limit) {
    while (last - first > 16) {
      if (depth_limit == 0) {   // std::__partial_sort(first, last, last): heap sort of the range
        std::make_heap(first, last, comp);
        std::sort_heap(first, last, comp);
        return;
      }
      --depth_limit;
      uint32_t *mid = first + (last - first) / 2;
      median_to_first(first, first + 1, mid, last - 1);
      uint32_t *cut = partition(first + 1, last, first);
      if (cut - base < keep) loop(cut, last, depth_limit);
      else if (cut - base < bound) bound = cut - base;
      last = cut;
    }
  }
  void linear_insert(uint32_t *last) {
    uint32_t val = *last;
    uint32_t *next = last - 1;
    while (comp(val, *next)) {
      *last = *next;
      last = next;
      --next;
    }
    *last = val;
  }
  void run(uint32_t *first, long n, long k) {
    base = first;
    keep = k < n ? k : n;
    bound = n;
    if (n <= 0) return;
    loop(first, first + n, lg(n) * 2);
    // __final_insertion_sort: guarded insertion for the first 16 positions, unguarded after them
    const long guarded = n > 16 ? 16 : n;
    for (long i = 1; i < guarded && i < bound; i++) {
      if (comp(first[i], first[0])) {
        uint32_t val = first[i];
        memmove(first + 1, first, (size_t)i * sizeof(uint32_t));
        first[0] = val;
      } else {
        linear_insert(first + i);
      }
    }
    for (long i = 16; i < bound; i++) linear_insert(first + i);
  }
};

// One cell: packed corner words (any order) -> the first K of std::sort-by-response applied to cv::FAST's row-major
// emission order, shifted to image coordinates (Grider_GRID.h:125-141). No mask test here.
inline void cell_top_k(const uint32_t *words, unsigned count, int cell_w, int cell_h, int cell_x0, int cell_y0, int K,
                       std::vector<uint32_t> &tmp, std::vector<uint32_t> &tmp2, std::vector<KeyPointLite> &out) {
  // The GPU appends corners in no particular order: restore cv::FAST's row-major order first.  Positions are unique, so
  // any sort does; a two-pass byte radix sort on y*cell_w + x has no data-dependent branches.
  (void)cell_h;
  if (count > 32 && (size_t)cell_w * cell_h <= (1u << 16)) {
    tmp.resize(count);
    tmp2.resize(count);
    unsigned h0[256] = {0}, h1[256] = {0};
    for (unsigned i = 0; i < count; i++) {
      uint32_t w = words[i], key = ((w >> 12) & 4095u) * (uint32_t)cell_w + (w & 4095u);
      h0[key & 255]++;
      h1[key >> 8]++;
    }
    unsigned s0 = 0, s1 = 0;
    for (int i = 0; i < 256; i++) {
      unsigned c0 = h0[i], c1 = h1[i];
      h0[i] = s0;
      h1[i] = s1;
      s0 += c0;
      s1 += c1;
    }
    for (unsigned i = 0; i < count; i++) {
      uint32_t w = words[i], key = ((w >> 12) & 4095u) * (uint32_t)cell_w + (w & 4095u);
      tmp2[h0[key & 255]++] = w;
    }
    for (unsigned i = 0; i < count; i++) {
      uint32_t w = tmp2[i], key = ((w >> 12) & 4095u) * (uint32_t)cell_w + (w & 4095u);
      tmp[h1[key >> 8]++] = w;
    }
  } else {
    tmp.assign(words, words + count);
    std::sort(tmp.begin(), tmp.end(), [](uint32_t a, uint32_t b) { return (a & 0xFFFFFFu) < (b & 0xFFFFFFu); });
  }
  // std::sort's permutation depends only on the element count, the starting order and the comparator's answers, not on
  // the element type: the packed words under "response greater" alone go through the same moves as the reference's
  // std::sort(kpts.begin(), kpts.end(), Grider_FAST::compare_response), ties included; only its first K places are used.
  auto by_response = [](uint32_t a, uint32_t b) { return (a >> 24) > (b >> 24); };
  StdSortPrefix<decltype(by_response)> prefix{nullptr, 0, 0, by_response};
  prefix.run(tmp.data(), (long)count, (long)K);
  out.clear();
  for (int i = 0; i < K && i < (int)count; i++) {
    uint32_t w = tmp[i];
    out.push_back(KeyPointLite{(float)(w & 4095u) + (float)cell_x0, (float)((w >> 12) & 4095u) + (float)cell_y0, 7.f, -1.f,
                               (float)(w >> 24), 0, -1});
  }
}

// Min-distance reject of refined new points (TrackKLT.cpp:497-512); returns kept indices, marks the close grid.
inline void close_reject(OccState &o, const Pt *pts, size_t n, std::vector<int> &keep) {
  keep.clear();
  for (size_t i = 0; i < n; i++) {
    int xg = (int)(pts[i].x / (float)o.mpd), yg = (int)(pts[i].y / (float)o.mpd);
    if (xg < 0 || xg >= o.cw || yg < 0 || yg >= o.ch) continue;
    uint8_t &c = o.close[(size_t)yg * o.cw + xg];
    if (c > 127) continue;
    c = 255;
    keep.push_back((int)i);
  }
}

}  // namespace uvio
