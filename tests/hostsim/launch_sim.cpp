// TEST SCAFFOLDING ONLY — see tests/hostsim/cuda_runtime.h. CPU stand-ins for the launch_* wrappers of kernels.cu: the
// same *_body.h code, one lane (GroupSeq), looping over what would be thread blocks / warps.
#include <stdio.h>
#include <stdlib.h>

#include <vector>

#include "../../uvio_b200/csrc/kernels.h"
#include "../../uvio_b200/csrc/lk_body.h"
#include "../../uvio_b200/csrc/subpix_body.h"
#include "../../uvio_b200/csrc/track_body.h"

namespace uvio {

cudaError_t launch_preprocess(cudaStream_t, const PreJob *jobs, int njobs, const PyrDesc *pyr_tab, uint8_t *,
                              int, int, int, int, int) {
  GroupSeq g;
  for (int j = 0; j < njobs; j++) {
    const ImgView I = jobs[j].src;
    const PyrDesc &P = pyr_tab[jobs[j].pyr];
    uint8_t *d0 = (uint8_t *)P.lv[0].p;
    if (jobs[j].method == 1) {
      unsigned hist[256] = {0};
      hist_rows(g, hist, I, 0, I.h);
      uint8_t lut[256];
      equalize_lut(hist, I.w * I.h, lut);
      lut_rows(g, lut, I, d0, P.lv[0].pitch, 0, I.h);
    } else if (jobs[j].method == 2) {
      ClaheGeom q = clahe_geom(I.w, I.h, 10.0, CLAHE_TILES, CLAHE_TILES);
      std::vector<uint8_t> luts(CLAHE_TILES * CLAHE_TILES * 256);
      for (int t = 0; t < CLAHE_TILES * CLAHE_TILES; t++) {
        unsigned hist[256] = {0};
        clahe_tile_hist(g, hist, I, q, t % CLAHE_TILES, t / CLAHE_TILES);
        clahe_tile_lut(hist, q, luts.data() + t * 256);
      }
      for (int y = 0; y < I.h; y++)
        for (int x = 0; x < I.w; x++) d0[(size_t)y * P.lv[0].pitch + x] = clahe_pixel(luts.data(), q, x, y, I.p[(size_t)y * I.pitch + x]);
    } else {
      for (int y = 0; y < I.h; y++) memcpy(d0 + (size_t)y * P.lv[0].pitch, I.p + (size_t)y * I.pitch, I.w);
    }
    for (int l = 1; l < P.nlv; l++)
      for (int oy = 0; oy < P.lv[l].h; oy += PD_ROWS)
        for (int ox = 0; ox < P.lv[l].w; ox += PD_COLS)
          pyrdown_strip(P.lv[l - 1], (uint8_t *)P.lv[l].p, P.lv[l].w, P.lv[l].h, P.lv[l].pitch, ox, oy);
  }
  return 0;
}

cudaError_t launch_downsample(cudaStream_t, const DownJob *jobs, int njobs, int, int) {
  for (int j = 0; j < njobs; j++) {
    const ImgView S = jobs[j].src, D = jobs[j].dst;
    for (int oy = 0; oy < D.h; oy += PD_ROWS)
      for (int ox = 0; ox < D.w; ox += PD_COLS) pyrdown_strip(S, (uint8_t *)D.p, D.w, D.h, D.pitch, ox, oy);
  }
  return 0;
}

cudaError_t launch_fast(cudaStream_t, const FastJob *jobs, int njobs, int, int, const PyrDesc *pyr_tab,
                        uint32_t *, unsigned *, unsigned *, uint32_t *out_host, unsigned *count_host) {
  GroupSeq g;
  FastSmem s;
  for (int j = 0; j < njobs; j++) {
    const FastJob job = jobs[j];
    unsigned count = 0;
    for (int ty = 0; ty < job.h; ty += FT_H)
      for (int tx = 0; tx < job.w; tx += FT_W) fast_tile(g, s, pyr_tab[job.img].lv[0], job, tx, ty, job.threshold, out_host, &count);
    count_host[j] = count;
  }
  return 0;
}

static float g_mask[SP_N];
cudaError_t upload_subpix_mask(float *mask_dev, cudaStream_t) {
  for (int i = 0; i < SP_W; i++) {
    float y = (float)(i - SP_WIN) / SP_WIN;
    float vy = expf(-y * y);
    for (int j = 0; j < SP_W; j++) {
      float x = (float)(j - SP_WIN) / SP_WIN;
      mask_dev[i * SP_W + j] = (float)(vy * expf(-x * x));
    }
  }
  return 0;
}

cudaError_t launch_subpix(cudaStream_t, const SubpixItem *items, int n, const PyrDesc *pyr_tab, const float *mask, float *out_xy,
                          float *out_xy_dev, const int *n_dev) {
  GroupSeq g;
  SubpixSmem s;
  if (n_dev) n = *n_dev;
  for (int i = 0; i < n; i++) {
    if (items[i].pyr < 0) continue;   // an unused slot of the device-side pipeline
    float x = items[i].x, y = items[i].y;
    subpix_point(g, s, pyr_tab[items[i].pyr].lv[0], mask, 20, 0.001 * 0.001, x, y);
    const int at = n_dev ? items[i].pad : i;
    out_xy[2 * at] = x;
    out_xy[2 * at + 1] = y;
    if (out_xy_dev) { out_xy_dev[2 * at] = x; out_xy_dev[2 * at + 1] = y; }
  }
  return 0;
}

cudaError_t launch_lk(cudaStream_t, const LkItem *items, int n, const PyrDesc *pyr_tab, const float *init_from_dev, int max_count,
                      double eps, LkOut *out_host, uint8_t *status_host, LkOut *out_dev, uint8_t *status_dev, bool, const int *out_idx,
                      const int *n_dev) {
  GroupSeq g;
  LkSmem s;
  if (n_dev) n = *n_dev;
  for (int i = 0; i < n; i++) {
    const LkItem it = items[i];
    if (it.prev_pyr < 0) continue;    // an unused slot of the device-side pipeline
    float px = it.px, py = it.py, nx = it.nx, ny = it.ny;
    const int at0 = out_idx ? out_idx[i] : i;
    if (init_from_dev) { px = nx = init_from_dev[2 * at0]; py = ny = init_from_dev[2 * at0 + 1]; }
    const PyrDesc &P0 = pyr_tab[it.prev_pyr], &P1 = pyr_tab[it.next_pyr];
    LkParams prm{P0.nlv - 1, max_count, eps * eps, 1e-4};
    uint8_t st;
    lk_track_point(g, s, P0.lv, P1.lv, prm, px, py, nx, ny, st);
    const int at = out_idx ? out_idx[i] : i;
    out_host[at] = out_dev[at] = LkOut{nx, ny};
    status_host[at] = status_dev[at] = st;
  }
  return 0;
}

cudaError_t read_lk_work(unsigned long long out[4], int) { out[0] = out[1] = out[2] = out[3] = 0; return 0; }

cudaError_t launch_ransac(cudaStream_t, const RansacJob *jobs, int njobs, int max_n, const LkItem *items, const LkOut *lk,
                          const uint8_t *status, const CamModel *cams, float *p0n, float *p1n, float *p1n_host, uint8_t *mask_host,
                          uint8_t *mask_dev, bool) {
  GroupSeq g;
  std::vector<float> buf((size_t)4 * max_n + 4);
  for (int j = 0; j < njobs; j++) {
    const RansacJob job = jobs[j];
    for (int i = 0; i < job.n; i++) {
      int k = job.off + i;
      undistort_point(cams[job.cam0], items[k].px, items[k].py, p0n[2 * k], p0n[2 * k + 1]);
      undistort_point(cams[job.cam1], lk[k].x, lk[k].y, p1n[2 * k], p1n[2 * k + 1]);
      p1n_host[2 * k] = p1n[2 * k];
      p1n_host[2 * k + 1] = p1n[2 * k + 1];
    }
    bool run = job.run && job.n >= 7;
    if (run) {
      RansacSmem s;
      s.x0 = buf.data(); s.y0 = s.x0 + job.n; s.x1 = s.y0 + job.n; s.y1 = s.x1 + job.n;
      ransac_job(g, s, p0n + 2 * job.off, p1n + 2 * job.off, job.n, job.thr, 0.999, 1000, mask_dev + job.off);
    }
    for (int i = 0; i < job.n; i++) mask_host[job.off + i] = run ? (uint8_t)(status[job.off + i] && mask_dev[job.off + i]) : 0;
  }
  return 0;
}

// ---- device-side bookkeeping (track_body.h) with one lane ----------------------------------------------------------------------
struct SimSmem {
  std::vector<unsigned long long> a, b;
  std::vector<float> p;
  DvSmem get(int P) {
    a.assign((size_t)P, 0xA5A5A5A5A5A5A5A5ull);
    b.assign((size_t)P, 0xA5A5A5A5A5A5A5A5ull);
    p.assign(2 * (size_t)P, -1e30f);
    return DvSmem{a.data(), b.data(), p.data()};
  }
};
static DvUnit sim_unit(const DvUnit *units, int u, unsigned bits) {
  DvUnit U = units[u];
  dv_apply_feed(U, bits);
  return U;
}
cudaError_t launch_dv_occupancy(cudaStream_t, const DvUnit *units, const DvHead *heads, int n_heads, const DvGlobals &X, int max_p,
                                const DvFeedBits &bits) {
  if (getenv("UVIO_DV_TRACE")) fprintf(stderr, "[sim] device-side pipeline: %d trackers\n", n_heads);
  SimSmem sm;
  for (int k = 0; k < 4; k++) X.n_items[k] = 0;
  for (int h = 0; h < n_heads; h++)
    for (int u = heads[h].first; u < heads[h].first + heads[h].count; u++) {
      X.feed_bits[u] = bits.v[u];
      dv_occupancy(GroupSeq(), sim_unit(units, u, bits.v[u]), X, sm.get(max_p));
    }
  return 0;
}
cudaError_t launch_dv_fast(cudaStream_t, const DvGlobals &X, const PyrDesc *pyr_tab, int, int, int) {
  GroupSeq g;
  FastSmem s;
  for (int a = 0; a < *X.fast_n_active; a++) {
    const int slot = X.fast_active[a];
    const FastJob job = X.fast_jobs[slot];
    for (int ty = 0; ty < job.h; ty += FT_H)
      for (int tx = 0; tx < job.w; tx += FT_W) fast_tile(g, s, pyr_tab[job.img].lv[0], job, tx, ty, job.threshold, X.fast_out, &X.fast_cnt[slot]);
  }
  return 0;
}
cudaError_t launch_dv_topk(cudaStream_t, const DvUnit *units, int n_units, int, const DvGlobals &X) {
  std::vector<uint32_t> buf(64);   // smaller than most cells: the in-place (global) sort path is exercised too
  for (int u = 0; u < n_units; u++)
    for (int side = 0; side < (units[u].stereo ? 2 : 1); side++)
      for (int r = 0; r < units[u].C; r++) dv_topk(GroupSeq(), sim_unit(units, u, X.feed_bits[u]), X, side, r, buf.data(), (int)buf.size());
  return 0;
}
cudaError_t launch_dv_select(cudaStream_t, const DvUnit *units, const DvHead *heads, int n_heads, const DvGlobals &X, int max_p) {
  SimSmem sm;
  *X.fast_n_active = 0;
  for (int h = 0; h < n_heads; h++)
    for (int u = heads[h].first; u < heads[h].first + heads[h].count; u++) dv_select(GroupSeq(), sim_unit(units, u, X.feed_bits[u]), X, sm.get(max_p));
  return 0;
}
cudaError_t launch_dv_assign(cudaStream_t, const DvUnit *units, const DvHead *heads, int n_heads, const DvGlobals &X, int max_p) {
  SimSmem sm;
  for (int h = 0; h < n_heads; h++)
    for (int u = heads[h].first; u < heads[h].first + heads[h].count; u++) dv_assign(GroupSeq(), sim_unit(units, u, X.feed_bits[u]), X, sm.get(max_p));
  return 0;
}
cudaError_t launch_dv_merge(cudaStream_t, const DvUnit *units, const DvHead *heads, int n_heads, const DvGlobals &X, int max_p) {
  SimSmem sm;
  for (int h = 0; h < n_heads; h++)
    for (int u = heads[h].first; u < heads[h].first + heads[h].count; u++) dv_merge(GroupSeq(), sim_unit(units, u, X.feed_bits[u]), X, sm.get(max_p));
  return 0;
}

}  // namespace uvio
