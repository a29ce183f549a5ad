// Stage entry points (uvio_klt_stage_*): one kernel at a time on caller-supplied inputs, for the per-stage parity tests.
#include "engine_internal.h"

namespace uvio {

// ---------------------------------------------------------------------------------------------------------------------
// Stage entry points: run one kernel on the pyramids a rig already holds (parity tests and micro-benchmarks).
// slot: 0 = the frame fed last ("curr"), 1 = the one before it ("last").
// ---------------------------------------------------------------------------------------------------------------------
static int slot_pyr(Rig *rig, size_t cam_id, int slot, CamState **out) {
  CamState *c = rig->find_cam(cam_id);
  if (!c || c->raw_in[0] < 0) return -1;
  *out = c;
  return c->pyr[slot ? c->last() : c->cur];
}

int stage_get_level(Rig *rig, size_t cam_id, int slot, int level, uint8_t *out, size_t out_stride, int *w, int *h, int *nlv) {
  Context *dev = rig->ctx;
  std::lock_guard<std::mutex> guard(dev->mtx);
  dev->background_wait();
  Lane *ctx = dev->lanes[0].get();
  CamState *c;
  int p = slot_pyr(rig, cam_id, slot, &c);
  if (p < 0) return E_CAMID;
  const PyrDesc &d = dev->pyr_tab[p];
  if (nlv) *nlv = d.nlv;
  if (level < 0 || level >= d.nlv) return E_ARG;
  if (w) *w = d.lv[level].w;
  if (h) *h = d.lv[level].h;
  if (out) {
    if (cudaSetDevice(dev->device) != cudaSuccess) return E_CUDA;
    CU_TRY(cudaMemcpy2D(out, out_stride, d.lv[level].p, d.lv[level].pitch, d.lv[level].w, d.lv[level].h, cudaMemcpyDeviceToHost));
  }
  return OK;
}

int stage_fast(Rig *rig, size_t cam_id, int slot, int x0, int y0, int cw, int ch, int threshold, int *xy, float *resp, int cap,
               int *n_out) {
  Context *dev = rig->ctx;
  std::lock_guard<std::mutex> guard(dev->mtx);
  dev->background_wait();
  Lane *ctx = dev->lanes[0].get();
  CamState *c;
  int p = slot_pyr(rig, cam_id, slot, &c);
  if (p < 0) return E_CAMID;
  if (x0 < 0 || y0 < 0 || cw <= 0 || ch <= 0 || x0 + cw > c->w || y0 + ch > c->h) return E_ARG;
  if (cudaSetDevice(dev->device) != cudaSuccess) return E_CUDA;
  if (dev->upload_tables() != OK) return E_CUDA;
  size_t words = (size_t)((cw + 1) / 2) * ((ch + 1) / 2);
  CU_TRY(ctx->fast_jobs.reserve(1));
  CU_TRY(ctx->fast_cnt.reserve(1));
  CU_TRY(ctx->fast_out.reserve(words));
  CU_TRY(ctx->fast_out_dev.reserve(words));
  CU_TRY(ctx->fast_cnt_dev.reserve(1, true));
  CU_TRY(ctx->fast_ticket.reserve(1, true));
  cancel_ahead(*ctx);   // the lane's FAST buffers are about to be reused
  ctx->fast_jobs.h[0] = FastJob{p, x0, y0, cw, ch, 0, (int)words, threshold};
  CU_TRY(ctx->fast_jobs.upload(ctx->stream, 1));
  CU_TRY(launch_fast(ctx->stream, ctx->fast_jobs.d, 1, cw, ch, dev->pyr_dev.d, ctx->fast_out_dev.d,
                     ctx->fast_cnt_dev.d, ctx->fast_ticket.d, ctx->fast_out.d, ctx->fast_cnt.d));
  if (ctx->sync() != OK) return E_CUDA;
  unsigned cnt = ctx->fast_cnt.h[0];
  std::vector<uint32_t> v(ctx->fast_out.h, ctx->fast_out.h + cnt);
  std::sort(v.begin(), v.end(), [](uint32_t a, uint32_t b) { return (a & 0xFFFFFFu) < (b & 0xFFFFFFu); });
  for (unsigned i = 0; i < cnt && (int)i < cap; i++) {
    xy[2 * i] = (int)(v[i] & 4095u);
    xy[2 * i + 1] = (int)((v[i] >> 12) & 4095u);
    resp[i] = (float)(v[i] >> 24);
  }
  *n_out = (int)cnt;
  return OK;
}

int stage_subpix(Rig *rig, size_t cam_id, int slot, const float *pts, int n, float *out) {
  Context *dev = rig->ctx;
  std::lock_guard<std::mutex> guard(dev->mtx);
  dev->background_wait();
  Lane *ctx = dev->lanes[0].get();
  CamState *c;
  int p = slot_pyr(rig, cam_id, slot, &c);
  if (p < 0) return E_CAMID;
  if (cudaSetDevice(dev->device) != cudaSuccess) return E_CUDA;
  if (dev->upload_tables() != OK) return E_CUDA;
  if (!ctx->sp_mask_ready) {
    CU_TRY(ctx->sp_mask.reserve(SP_N));
    CU_TRY(upload_subpix_mask(ctx->sp_mask.d, ctx->stream));
    ctx->sp_mask_ready = true;
  }
  CU_TRY(ctx->sp_items.reserve(n));
  CU_TRY(ctx->sp_out.reserve(2 * (size_t)n));
  for (int i = 0; i < n; i++) ctx->sp_items.h[i] = SubpixItem{pts[2 * i], pts[2 * i + 1], p, 0};
  CU_TRY(ctx->sp_items.upload(ctx->stream, (size_t)n));
  CU_TRY(launch_subpix(ctx->stream, ctx->sp_items.d, n, dev->pyr_dev.d, ctx->sp_mask.d, ctx->sp_out.d, nullptr));
  if (ctx->sync() != OK) return E_CUDA;
  memcpy(out, ctx->sp_out.h, sizeof(float) * 2 * (size_t)n);
  return OK;
}

// LK from (cam0, slot0) to (cam1, slot1); optionally followed by undistort + RANSAC on the result (mask/p1n may be null)
int stage_lk(Rig *rig, size_t cam0, int slot0, size_t cam1, int slot1, const float *pts0, const float *pts1_init, int n,
             float *out, uint8_t *status, uint8_t *mask, float *p1n) {
  Context *dev = rig->ctx;
  std::lock_guard<std::mutex> guard(dev->mtx);
  dev->background_wait();
  Lane *ctx = dev->lanes[0].get();
  CamState *c0, *c1;
  int p0 = slot_pyr(rig, cam0, slot0, &c0), p1 = slot_pyr(rig, cam1, slot1, &c1);
  if (p0 < 0 || p1 < 0) return E_CAMID;
  if (dev->pyr_tab[p0].nlv != dev->pyr_tab[p1].nlv || c0->w != c1->w || c0->h != c1->h) return E_SIZE;
  if (cudaSetDevice(dev->device) != cudaSuccess) return E_CUDA;
  if (dev->upload_tables() != OK) return E_CUDA;
  size_t N = (size_t)n;
  CU_TRY(ctx->lk_items.reserve(N));
  CU_TRY(ctx->lk_out.reserve(N));
  CU_TRY(ctx->lk_status.reserve(N));
  CU_TRY(ctx->lk_out_dev.reserve(N));
  CU_TRY(ctx->lk_status_dev.reserve(N));
  for (int i = 0; i < n; i++)
    ctx->lk_items.h[i] = LkItem{pts0[2 * i], pts0[2 * i + 1], pts1_init[2 * i], pts1_init[2 * i + 1], p0, p1};
  CU_TRY(ctx->lk_items.upload(ctx->stream, (size_t)n));
  CU_TRY(launch_lk(ctx->stream, ctx->lk_items.d, n, dev->pyr_dev.d, nullptr, rig->prm.lk_max_iters, rig->prm.lk_eps, ctx->lk_out.d,
                   ctx->lk_status.d, ctx->lk_out_dev.d, ctx->lk_status_dev.d));
  if (mask) {
    CU_TRY(ctx->rs_jobs.reserve(1));
    CU_TRY(ctx->p0n_dev.reserve(2 * N));
    CU_TRY(ctx->p1n_dev.reserve(2 * N));
    CU_TRY(ctx->p1n.reserve(2 * N));
    CU_TRY(ctx->rs_mask.reserve(N));
    CU_TRY(ctx->rs_mask_dev.reserve(N));
    RansacJob j;
    j.off = 0;
    j.n = n;
    j.cam0 = c0->cam_model_idx;
    j.cam1 = c1->cam_model_idx;
    j.thr = 2.0 / std::max(c0->max_focal, c1->max_focal);
    j.run = n >= 10;
    ctx->rs_jobs.h[0] = j;
    CU_TRY(ctx->rs_jobs.upload(ctx->stream, 1));
    CU_TRY(launch_ransac(ctx->stream, ctx->rs_jobs.d, 1, n, ctx->lk_items.d, ctx->lk_out_dev.d, ctx->lk_status_dev.d, dev->cam_dev.d,
                         ctx->p0n_dev.d, ctx->p1n_dev.d, ctx->p1n.d, ctx->rs_mask.d, ctx->rs_mask_dev.d));
  }
  if (ctx->sync() != OK) return E_CUDA;
  for (int i = 0; i < n; i++) {
    out[2 * i] = ctx->lk_out.h[i].x;
    out[2 * i + 1] = ctx->lk_out.h[i].y;
    status[i] = ctx->lk_status.h[i];
  }
  if (mask) {
    memcpy(mask, ctx->rs_mask.h, N);
    if (p1n) memcpy(p1n, ctx->p1n.h, sizeof(float) * 2 * N);
  }
  return OK;
}

// undistort + RANSAC on caller-supplied pixel correspondences (no LK): mask = inliers, p0n/p1n = normalised points
int stage_ransac(Rig *rig, size_t cam0, size_t cam1, const float *pts0, const float *pts1, int n, uint8_t *mask, float *p0n,
                 float *p1n) {
  Context *dev = rig->ctx;
  std::lock_guard<std::mutex> guard(dev->mtx);
  dev->background_wait();
  Lane *ctx = dev->lanes[0].get();
  CamState *c0 = rig->find_cam(cam0), *c1 = rig->find_cam(cam1);
  if (!c0 || !c1) return E_CAMID;
  if (cudaSetDevice(dev->device) != cudaSuccess) return E_CUDA;
  if (dev->upload_tables() != OK) return E_CUDA;
  size_t N = (size_t)n;
  CU_TRY(ctx->lk_items.reserve(N));
  CU_TRY(ctx->lk_out_dev.reserve(N));
  CU_TRY(ctx->lk_status_dev.reserve(N));
  CU_TRY(ctx->rs_jobs.reserve(1));
  CU_TRY(ctx->p0n_dev.reserve(2 * N));
  CU_TRY(ctx->p1n_dev.reserve(2 * N));
  CU_TRY(ctx->p1n.reserve(2 * N));
  CU_TRY(ctx->rs_mask.reserve(N));
  CU_TRY(ctx->rs_mask_dev.reserve(N));
  std::vector<uint8_t> ones(N, 1);
  for (int i = 0; i < n; i++) ctx->lk_items.h[i] = LkItem{pts0[2 * i], pts0[2 * i + 1], 0, 0, 0, 0};
  CU_TRY(cudaMemcpyAsync(ctx->lk_out_dev.d, pts1, sizeof(float) * 2 * N, cudaMemcpyHostToDevice, ctx->stream));
  CU_TRY(cudaMemcpyAsync(ctx->lk_status_dev.d, ones.data(), N, cudaMemcpyHostToDevice, ctx->stream));
  RansacJob j;
  j.off = 0;
  j.n = n;
  j.cam0 = c0->cam_model_idx;
  j.cam1 = c1->cam_model_idx;
  j.thr = 2.0 / std::max(c0->max_focal, c1->max_focal);
  j.run = n >= 10;
  ctx->rs_jobs.h[0] = j;
  CU_TRY(ctx->lk_items.upload(ctx->stream, (size_t)n));
  CU_TRY(ctx->rs_jobs.upload(ctx->stream, 1));
  CU_TRY(launch_ransac(ctx->stream, ctx->rs_jobs.d, 1, n, ctx->lk_items.d, ctx->lk_out_dev.d, ctx->lk_status_dev.d, dev->cam_dev.d,
                       ctx->p0n_dev.d, ctx->p1n_dev.d, ctx->p1n.d, ctx->rs_mask.d, ctx->rs_mask_dev.d));
  if (ctx->sync() != OK) return E_CUDA;
  memcpy(mask, ctx->rs_mask.h, N);
  if (p1n) memcpy(p1n, ctx->p1n.h, sizeof(float) * 2 * N);
  if (p0n) CU_TRY(cudaMemcpy(p0n, ctx->p0n_dev.d, sizeof(float) * 2 * N, cudaMemcpyDeviceToHost));
  return OK;
}


}  // namespace uvio
