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
#include <unordered_set>
#include <utility>
#include <vector>

#include "engine.h"

namespace uvio {

struct OccState {
  int cols = 0, rows = 0, mpd = 0, gx = 0, gy = 0;
  int cw = 0, ch = 0;                 // size_close
  float size_x = 0, size_y = 0;       // float cell size of the coarse grid (TrackKLT.cpp:404-405)
  std::vector<uint8_t> close, grid;
  std::vector<Pt> pts;                // kept points, in order
  std::vector<size_t> ids;
  std::vector<std::pair<int, int>> boxes;     // centres of the (2*mpd+1)^2 boxes of mask_updated (:457-461)
  size_t painted = 0;                         // how many of them are currently painted into the scratch image
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
  o.painted = 0;
}

// Runs the erase-while-iterating loop over n more points (kept ones are appended to o.pts / o.ids).
// mask: the image mask tested at :446/:762 (null = all zeros).  The 255-boxes of mask_updated (:457-461) are only recorded
// here; occ_paint() rasterises them when a detection actually needs mask_updated (most frames need none).
// stereo_ids: right pass only, ids exempt from the closeness erase (:754-759).
inline void occ_feed(OccState &o, const Pt *pts, const size_t *ids, size_t n, const uint8_t *mask,
                     const std::unordered_set<size_t> *stereo_ids) {
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
      if (occupied && !stereo_ids->count(ids[i])) continue;
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

inline void occ_fill(OccState &o, uint8_t *paint, size_t from, size_t to, int value) {
  for (size_t i = from; i < to; i++)
    for (int yy = o.boxes[i].second - o.mpd; yy <= o.boxes[i].second + o.mpd; yy++)
      memset(paint + (size_t)yy * o.cols + (o.boxes[i].first - o.mpd), value, (size_t)(2 * o.mpd + 1));
}
// paint = (mask_updated minus the original mask): all recorded boxes set to 255 in the (otherwise zero) scratch image
inline void occ_paint(OccState &o, uint8_t *paint) {
  occ_fill(o, paint, o.painted, o.boxes.size(), 255);
  o.painted = o.boxes.size();
}
inline void occ_unpaint(OccState &o, uint8_t *paint) {
  occ_fill(o, paint, 0, o.painted, 0);
  o.painted = 0;
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

// One cell: packed corner words (any order) -> the first K of std::sort-by-response applied to cv::FAST's row-major
// emission order, shifted to image coordinates (Grider_GRID.h:125-141). No mask test here.
inline void cell_top_k(const uint32_t *words, unsigned count, int cell_x0, int cell_y0, int K, std::vector<uint32_t> &tmp,
                       std::vector<KeyPointLite> &kps, std::vector<KeyPointLite> &out) {
  tmp.assign(words, words + count);
  std::sort(tmp.begin(), tmp.end(), [](uint32_t a, uint32_t b) { return (a & 0xFFFFFFu) < (b & 0xFFFFFFu); });
  kps.resize(count);
  for (unsigned i = 0; i < count; i++) {
    uint32_t w = tmp[i];
    kps[i] = KeyPointLite{(float)(w & 4095u), (float)((w >> 12) & 4095u), 7.f, -1.f, (float)(w >> 24), 0, -1};
  }
  std::sort(kps.begin(), kps.end(), compare_response);
  out.clear();
  for (int i = 0; i < K && i < (int)count; i++) {
    KeyPointLite k = kps[i];
    k.x += (float)cell_x0;
    k.y += (float)cell_y0;
    out.push_back(k);
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
