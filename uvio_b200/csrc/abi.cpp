// C ABI (include/uvio_klt.h) over the engine.  Nothing throws across this boundary.
#include "../../include/uvio_klt.h"

#include <string.h>

#include <algorithm>
#include <functional>
#include <new>
#include <vector>

#include "engine.h"

using namespace uvio;

struct uvio_klt {
  Rig *rig;
};

#define GUARD_BEGIN try {
#define GUARD_END                      \
  }                                    \
  catch (const std::bad_alloc &) {     \
    return UVIO_E_ARG;                 \
  }                                    \
  catch (...) {                        \
    return UVIO_E_ARG;                 \
  }

extern "C" {

void uvio_klt_default_params(UvioKltParams *p) {
  if (!p) return;
  p->num_features = 200;
  p->num_aruco = 0;
  p->use_stereo = 1;
  p->histogram_method = UVIO_HIST_HISTOGRAM;
  p->fast_threshold = 20;
  p->grid_x = 5;
  p->grid_y = 5;
  p->min_px_dist = 10;
  p->pyr_levels = 5;
  p->win_size = 15;
  p->lk_max_iters = 30;
  p->lk_eps = 0.01;
  p->downsample_cameras = 0;
}

int uvio_klt_create(const UvioKltParams *p, const UvioCamCalib *cams, int n_cams, int cuda_device, uvio_klt **out) {
  GUARD_BEGIN
  if (!p || !cams || n_cams <= 0 || !out) return UVIO_E_ARG;
  if (p->num_features <= 0 || p->grid_x <= 0 || p->grid_y <= 0 || p->min_px_dist <= 0 || p->pyr_levels < 0 || p->win_size != 15 ||
      p->histogram_method < 0 || p->histogram_method > 2 || p->lk_max_iters != 30 || p->lk_eps != 0.01)
    return UVIO_E_ARG;   // the LK termination criteria are the reference's constants (TrackKLT.cpp:670,857); others are refused
  *out = nullptr;
  int err = 0;
  Context *ctx = context_for_device(cuda_device, &err);
  if (!ctx) return UVIO_E_CUDA;   // no device, no tracker: there is deliberately no CPU path behind this ABI
  Params q;
  q.num_features = p->num_features;
  q.num_aruco = p->num_aruco;
  q.use_stereo = p->use_stereo;
  q.histogram_method = p->histogram_method;
  q.fast_threshold = p->fast_threshold;
  q.grid_x = p->grid_x;
  q.grid_y = p->grid_y;
  q.min_px_dist = p->min_px_dist;
  q.pyr_levels = p->pyr_levels;
  q.win_size = p->win_size;
  q.lk_max_iters = p->lk_max_iters;
  q.lk_eps = p->lk_eps;
  q.downsample_cameras = p->downsample_cameras != 0;
  std::vector<CamCalib> cc(n_cams);
  for (int i = 0; i < n_cams; i++) {
    if (cams[i].model != UVIO_CAM_RADTAN && cams[i].model != UVIO_CAM_EQUI) return UVIO_E_ARG;
    for (int k = 0; k < i; k++)
      if (cams[k].cam_id == cams[i].cam_id) return UVIO_E_ARG;
    cc[i].cam_id = cams[i].cam_id;
    cc[i].model.model = cams[i].model;
    for (int k = 0; k < 4; k++) {
      cc[i].model.k[k] = cams[i].intrinsics[k];
      cc[i].model.d[k] = cams[i].distortion[k];
    }
  }
  uvio_klt *h = new uvio_klt;
  h->rig = new Rig(ctx, q, cc.data(), n_cams);
  *out = h;
  return UVIO_OK;
  GUARD_END
}

void uvio_klt_destroy(uvio_klt *h) {
  if (!h) return;
  delete h->rig;   // frees the pyramids of its cameras; their upload buffers are recycled by the device context
  delete h;
}

// feed_batch and submit_batch share everything but the engine call
static int run_batch(uvio_klt *const *hs, UvioFrame *frames, int n, bool announce) {
  GUARD_BEGIN
  if (!hs || !frames || n <= 0) return UVIO_E_ARG;
  Context *ctx = hs[0] ? hs[0]->rig->ctx : nullptr;
  std::vector<FeedRequest> reqs(n);
  std::vector<std::vector<HostImage>> imgs(n), masks(n);
  std::vector<std::unique_lock<std::mutex>> locks;
  for (int i = 0; i < n; i++) {
    if (!hs[i] || hs[i]->rig->ctx != ctx) return UVIO_E_ARG;
    for (int k = 0; k < i; k++)
      if (hs[k] == hs[i]) return UVIO_E_ARG;
  }
  {
    // the handles' locks are taken in address order, whatever order the caller lists them in: two threads feeding
    // overlapping sets ([A, B] against [B, A]) cannot deadlock
    std::vector<Rig *> order(n);
    for (int i = 0; i < n; i++) order[i] = hs[i]->rig;
    std::sort(order.begin(), order.end(), [](Rig *a, Rig *b) { return std::less<Rig *>()(a, b); });
    for (Rig *r : order) locks.emplace_back(r->mtx);
  }
  for (int i = 0; i < n; i++) {
    const UvioFrame &f = frames[i];
    FeedRequest &r = reqs[i];
    r.rig = hs[i]->rig;
    r.timestamp = f.timestamp;
    r.n = f.n_img;
    r.sensor_ids = f.sensor_ids;
    r.status = UVIO_OK;
    bool ptrs = f.n_img > 0 && f.sensor_ids && f.img && f.img_stride && f.width && f.height;
    if (!ptrs) {
      r.n = 0;   // rejected as a malformed message inside the engine
      r.images = r.masks = nullptr;
      continue;
    }
    imgs[i].resize(f.n_img);
    masks[i].resize(f.n_img);
    for (int k = 0; k < f.n_img; k++) {
      imgs[i][k] = HostImage{f.img[k], f.img_stride[k], f.width[k], f.height[k]};
      const uint8_t *m = f.mask ? f.mask[k] : nullptr;
      masks[i][k] = HostImage{m, (m && f.mask_stride) ? f.mask_stride[k] : (size_t)f.width[k], f.width[k], f.height[k]};
    }
    r.images = imgs[i].data();
    r.masks = masks[i].data();
  }
  int rc = announce ? submit_batch(ctx, reqs.data(), n) : feed_batch(ctx, reqs.data(), n);
  for (int i = 0; i < n; i++) {
    frames[i].status = reqs[i].status;
    if (rc == UVIO_E_CUDA) {
      frames[i].status = UVIO_E_CUDA;
      hs[i]->rig->last_error = UVIO_E_CUDA;
      hs[i]->rig->cuda_error = context_cuda_error(ctx);
    }
  }
  return rc;
  GUARD_END
}

int uvio_klt_feed_batch(uvio_klt *const *hs, UvioFrame *frames, int n) { return run_batch(hs, frames, n, false); }

int uvio_klt_submit_batch(uvio_klt *const *hs, UvioFrame *frames, int n) { return run_batch(hs, frames, n, true); }

int uvio_klt_submit(uvio_klt *h, const UvioFrame *frame) {
  if (!h || !frame) return UVIO_E_ARG;
  UvioFrame f = *frame;
  uvio_klt *hs[1] = {h};
  int rc = run_batch(hs, &f, 1, true);
  return rc != UVIO_OK ? rc : f.status;
}

int uvio_klt_feed(uvio_klt *h, const UvioFrame *frame) {
  if (!h || !frame) return UVIO_E_ARG;
  UvioFrame f = *frame;
  uvio_klt *hs[1] = {h};
  int rc = uvio_klt_feed_batch(hs, &f, 1);
  return rc != UVIO_OK ? rc : f.status;
}

int uvio_klt_get_new_observations(uvio_klt *h, UvioObs *out, int cap, int *n) {
  if (!h || !n) return UVIO_E_ARG;
  std::lock_guard<std::mutex> lk(h->rig->mtx);
  if (!h->rig->dv_obs.empty()) {   // device-side pipeline: straight out of the mapped memory the merge kernel wrote
    static_assert(sizeof(Obs) == sizeof(UvioObs), "same bytes");
    int total = 0;
    for (auto &part : h->rig->dv_obs) {
      const int take = out ? std::max(0, std::min(part.second, cap - total)) : 0;
      if (take) memcpy(out + total, part.first, sizeof(UvioObs) * (size_t)take);
      total += part.second;
    }
    *n = total;
    return UVIO_OK;
  }
  const std::vector<Obs> &o = h->rig->obs;
  *n = (int)o.size();
  for (int i = 0; out && i < cap && i < (int)o.size(); i++) out[i] = UvioObs{o[i].id, o[i].cam_id, o[i].u, o[i].v, o[i].un, o[i].vn};
  return UVIO_OK;
}

int uvio_klt_get_last(uvio_klt *h, size_t cam_id, size_t *ids, float *uv, int cap, int *n) {
  if (!h || !n) return UVIO_E_ARG;
  std::lock_guard<std::mutex> lk(h->rig->mtx);
  dv_sync_mirror(*h->rig);
  CamState *c = h->rig->find_cam(cam_id);
  if (!c) return UVIO_E_CAMID;
  *n = (int)c->pts_last.size();
  for (int i = 0; i < cap && i < *n; i++) {
    if (ids) ids[i] = c->ids_last[i];
    if (uv) {
      uv[2 * i] = c->pts_last[i].x;
      uv[2 * i + 1] = c->pts_last[i].y;
    }
  }
  return UVIO_OK;
}

int uvio_klt_set_num_features(uvio_klt *h, int n) {
  if (!h || n <= 0) return UVIO_E_ARG;
  std::lock_guard<std::mutex> lk(h->rig->mtx);
  context_quiesce(h->rig->ctx);   // a background detection of the last feed may be reading num_features / currid / ids_last
  h->rig->prm.num_features = n;
  h->rig->epoch++;
  return UVIO_OK;
}

int uvio_klt_change_feat_id(uvio_klt *h, size_t id_old, size_t id_new) {
  if (!h) return UVIO_E_ARG;
  std::lock_guard<std::mutex> lk(h->rig->mtx);
  context_quiesce(h->rig->ctx);
  dv_sync_mirror(*h->rig);
  for (auto &c : h->rig->cams)                       // TrackBase.cpp:241-247
    for (auto &id : c.ids_last)
      if (id == id_old) id = id_new;
  h->rig->dv_dirty = true;   // the device-side pipeline re-uploads ids_last before the next feed
  h->rig->epoch++;
  return UVIO_OK;
}

int uvio_klt_get_timing(uvio_klt *h, float us[5]) {
  if (!h || !us) return UVIO_E_ARG;
  std::lock_guard<std::mutex> lk(h->rig->mtx);
  memcpy(us, h->rig->stage_us, sizeof(float) * 5);
  return UVIO_OK;
}

int uvio_klt_last_error(uvio_klt *h) { return h ? h->rig->last_error : UVIO_E_ARG; }
int uvio_klt_cuda_error(uvio_klt *h) { return h ? context_cuda_error(h->rig->ctx) : 0; }
const char *uvio_klt_version(void) { return "uvio_klt 0.1 (sm_100a)"; }

int uvio_klt_device_stats(int cuda_device, unsigned long long out[12]) {
  if (!out) return UVIO_E_ARG;
  int err = 0;
  Context *ctx = context_for_device(cuda_device, &err);
  if (!ctx) return UVIO_E_CUDA;
  context_stats(ctx, out);
  return UVIO_OK;
}
int uvio_klt_timer_start(int cuda_device) {
  int err = 0;
  Context *ctx = context_for_device(cuda_device, &err);
  return ctx ? context_timer(ctx, 0, nullptr) : UVIO_E_CUDA;
}
int uvio_klt_timer_stop(int cuda_device, float *ms) {
  if (!ms) return UVIO_E_ARG;
  int err = 0;
  Context *ctx = context_for_device(cuda_device, &err);
  return ctx ? context_timer(ctx, 1, ms) : UVIO_E_CUDA;
}

int uvio_klt_host_profile(int cuda_device, int reset, double us[8]) {
  int err = 0;
  Context *ctx = context_for_device(cuda_device, &err);
  if (!ctx) return UVIO_E_CUDA;
  context_host_profile(ctx, reset, us);
  return UVIO_OK;
}

int uvio_klt_profile(int cuda_device, int enable, double ms[16], unsigned long long launches[16]) {
  int err = 0;
  Context *ctx = context_for_device(cuda_device, &err);
  if (!ctx) return UVIO_E_CUDA;
  double m[K_COUNT];
  unsigned long long l[K_COUNT];
  int rc = context_profile(ctx, enable, m, l);
  for (int i = 0; i < 16; i++) {
    if (ms) ms[i] = i < K_COUNT ? m[i] : 0.0;
    if (launches) launches[i] = i < K_COUNT ? l[i] : 0ull;
  }
  return rc;
}

int uvio_klt_get_new_observations_batch(uvio_klt *const *hs, int n, UvioObs *out, int cap_each, int *counts) {
  if (!hs || n <= 0 || !out || !counts || cap_each < 0) return UVIO_E_ARG;
  for (int i = 0; i < n; i++) {
    int rc = uvio_klt_get_new_observations(hs[i], out + (size_t)i * cap_each, cap_each, &counts[i]);
    if (rc != UVIO_OK) return rc;
  }
  return UVIO_OK;
}

int uvio_klt_stage_get_level(uvio_klt *h, size_t cam_id, int slot, int level, uint8_t *out, size_t out_stride, int *w, int *hgt,
                             int *n_levels) {
  GUARD_BEGIN
  if (!h) return UVIO_E_ARG;
  return stage_get_level(h->rig, cam_id, slot, level, out, out_stride, w, hgt, n_levels);
  GUARD_END
}
int uvio_klt_stage_fast(uvio_klt *h, size_t cam_id, int slot, int x0, int y0, int cw, int ch, int threshold, int *xy, float *resp,
                        int cap, int *n) {
  GUARD_BEGIN
  if (!h || !xy || !resp || !n) return UVIO_E_ARG;
  return stage_fast(h->rig, cam_id, slot, x0, y0, cw, ch, threshold, xy, resp, cap, n);
  GUARD_END
}
int uvio_klt_stage_subpix(uvio_klt *h, size_t cam_id, int slot, const float *pts, int n, float *out) {
  GUARD_BEGIN
  if (!h || !pts || !out || n <= 0) return UVIO_E_ARG;
  return stage_subpix(h->rig, cam_id, slot, pts, n, out);
  GUARD_END
}
int uvio_klt_stage_lk(uvio_klt *h, size_t cam0, int slot0, size_t cam1, int slot1, const float *pts0, const float *pts1_init, int n,
                      float *out, uint8_t *status, uint8_t *mask, float *p1n) {
  GUARD_BEGIN
  if (!h || !pts0 || !pts1_init || !out || !status || n <= 0) return UVIO_E_ARG;
  return stage_lk(h->rig, cam0, slot0, cam1, slot1, pts0, pts1_init, n, out, status, mask, p1n);
  GUARD_END
}
int uvio_klt_stage_ransac(uvio_klt *h, size_t cam0, size_t cam1, const float *pts0, const float *pts1, int n, uint8_t *mask,
                          float *p0n, float *p1n) {
  GUARD_BEGIN
  if (!h || !pts0 || !pts1 || !mask || n <= 0) return UVIO_E_ARG;
  return stage_ransac(h->rig, cam0, cam1, pts0, pts1, n, mask, p0n, p1n);
  GUARD_END
}

}  // extern "C"
