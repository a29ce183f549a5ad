// Host side and C ABI (include/uvio_desc.h) of the descriptor tracker: stands in for ov_core::TrackDescriptor
// (ov_core/src/track/TrackDescriptor.cpp).  The host validates the message, uploads the frames, assembles the tables of the
// launch and issues one fixed chain of kernels per (batched) feed:
//
//   k_preprocess (equalise, one-level pyramid) -> k_fast (every grid cell, Grider_FAST.h:105-147) -> k_desc_topk ->
//   k_desc_gather -> k_subpix (:158-172) -> k_desc_orb (cv::ORB::compute) -> k_desc_detect (L->R robust_match + ids, or the
//   mono occupancy reject + ids) -> k_desc_temporal (last->new robust_match, both cameras) -> k_desc_merge
//
// followed by a single stream synchronisation.  Which branch of the reference a frame takes (first frame, lost track, too few
// matches) is decided on the device from the counts there; the host learns the outcome from four words in mapped memory.
#include "../../include/uvio_desc.h"

#include <string.h>

#include <algorithm>
#include <functional>
#include <mutex>
#include <new>
#include <unordered_map>
#include <vector>

#include "desc.h"
#include "nvtx_ranges.h"
#include "subpix_body.h"

using namespace uvio;

namespace {

inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

struct DescCtx;

struct DescCam {                      // per camera id: TrackBase.h:180-189 img_last / pts_last / ids_last, TrackDescriptor.h:170 desc_last
  size_t cam_id = 0;
  CamModel model{};
  double max_focal = 0;
  int cam_slot = -1;                  // index in the context's CamModel table
  int w = 0, h = 0;                   // fixed by the first frame
  int n_last = 0;                     // host copy of the device's count after the last feed
  // geometry of the extraction grid for the current num_features (Grider_FAST.h:83-103)
  int K = 0, size_x = 0, size_y = 0, ct_cols = 0, ct_rows = 0, ncells = 0, cell_cap = 0, P = 0;
  // device memory
  void *base = nullptr;
  uint8_t *eq = nullptr, *mask = nullptr;   // (the uploaded frame itself lives in the launch arena: it is consumed by k_preprocess)
  int raw_pitch = 0, eq_pitch = 0;
  bool mask_zero = false;
  DescSide side{};                    // the pointers of this camera's lists (everything but the per-launch indices)
  std::vector<FastJob> jobs;          // its cells (out_off relative to the camera's region of the launch's corner arena)
};

struct DescScratch {                  // what one unit of a feed needs besides its cameras' lists
  void *base = nullptr;
  int P = 0, grid_words = 0;
  int *grid = nullptr, *m_q = nullptr, *m_t = nullptr, *m_n = nullptr, *fwd = nullptr, *bwd = nullptr;
  float *p0n = nullptr, *p1n = nullptr;
  uint8_t *rs_mask = nullptr;
  DvObs *obs = nullptr;               // mapped host memory
  int *head = nullptr;
  void *host_base = nullptr;
};

struct DescRig {
  DescCtx *ctx = nullptr;
  UvioDescParams prm{};
  std::vector<DescCam> cams;
  std::unordered_map<size_t, int> cam_index;
  std::vector<DescScratch> scratch;   // one per camera: a feed has at most that many units
  unsigned long long *currid_dev = nullptr;
  unsigned long long currid = 0;      // host mirror of TrackBase::currid (every feed ends with a synchronisation)
  std::vector<UvioObs> obs;
  int cuda_error = 0;
  float last_us = 0.f;
  std::mutex mtx;
};

struct DescCtx {                      // one per device
  int device = 0;
  std::mutex mtx;                     // one batch at a time
  cudaStream_t st = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  float *sp_mask = nullptr;
  CamModel *cams_dev = nullptr;
  std::vector<int> cam_free;
  int cam_used = 0;
  static constexpr int CAM_CAP = 4096;
  // transient arena of a launch (grows)
  void *arena = nullptr;
  size_t arena_bytes = 0;
  void *stage = nullptr;              // pinned host staging of the launch tables
  size_t stage_bytes = 0;
  int cuda_error = 0;
};

std::mutex g_ctx_mtx;
std::unordered_map<int, DescCtx *> g_ctx;

#define CK(expr)                                  \
  do {                                            \
    cudaError_t e_ = (expr);                      \
    if (e_ != cudaSuccess) return (int)e_;        \
  } while (0)

DescCtx *ctx_for_device(int device) {
  std::lock_guard<std::mutex> lk(g_ctx_mtx);
  auto it = g_ctx.find(device);
  if (it != g_ctx.end()) return it->second;
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || device < 0 || device >= n) return nullptr;
  if (cudaSetDevice(device) != cudaSuccess) return nullptr;
  DescCtx *c = new DescCtx;
  c->device = device;
  bool ok = cudaStreamCreateWithFlags(&c->st, cudaStreamNonBlocking) == cudaSuccess && cudaEventCreate(&c->ev0) == cudaSuccess &&
            cudaEventCreate(&c->ev1) == cudaSuccess && cudaMalloc((void **)&c->sp_mask, sizeof(float) * SP_N) == cudaSuccess &&
            cudaMalloc((void **)&c->cams_dev, sizeof(CamModel) * DescCtx::CAM_CAP) == cudaSuccess &&
            upload_subpix_mask(c->sp_mask, c->st) == cudaSuccess && cudaStreamSynchronize(c->st) == cudaSuccess;
  if (!ok) {
    delete c;
    return nullptr;
  }
  g_ctx[device] = c;
  return c;
}

// Grider_FAST.h:83-103 for one camera
void cam_geometry(DescCam &c, const UvioDescParams &p) {
  int gx = p.grid_x, gy = p.grid_y;
  if (p.num_features < gx * gy) {                                         // :83-87
    double ratio = (double)gx / (double)gy;
    gy = (int)ceil(sqrt(p.num_features / ratio));
    gx = (int)ceil(gy * ratio);
  }
  c.K = (int)((double)p.num_features / (double)(gx * gy)) + 1;            // :88
  c.size_x = c.w / gx;                                                    // :94-95
  c.size_y = c.h / gy;
  c.ct_cols = c.size_x > 0 ? c.w / c.size_x : 0;                          // :102-103
  c.ct_rows = c.size_y > 0 ? c.h / c.size_y : 0;
  c.ncells = c.ct_cols * c.ct_rows;
  int cap = ((c.size_x + 1) / 2) * ((c.size_y + 1) / 2);                  // NMS worst case
  int pow2 = 1;
  while (pow2 < cap) pow2 <<= 1;
  c.cell_cap = pow2;
  c.jobs.assign(c.ncells, FastJob{});
  for (int r = 0; r < c.ncells; r++) {                                    // :105-117
    FastJob &j = c.jobs[r];
    const int x = r % c.ct_cols * c.size_x, y = r / c.ct_cols * c.size_y;
    j.img = 0;
    j.x0 = x;
    j.y0 = y;
    j.w = (x + c.size_x > c.w || y + c.size_y > c.h) ? 0 : c.size_x;      // :112 (w = 0: skipped)
    j.h = j.w ? c.size_y : 0;
    j.out_off = r * c.cell_cap;
    j.out_cap = c.cell_cap;
    j.threshold = p.fast_threshold;
  }
}

struct Carver {                       // lays arrays out in one allocation
  size_t off = 0;
  template <class T>
  size_t take(size_t n) {
    off = align_up(off, 256);
    size_t at = off;
    off += n * sizeof(T);
    return at;
  }
};

// (re)allocates the camera's device memory for its image size and the current P; the last lists survive a re-allocation
int cam_alloc(DescRig &rig, DescCam &c, int P) {
  DescCtx &ctx = *rig.ctx;
  const int oldP = c.P;
  void *old = c.base;
  DescSide olds = c.side;
  Carver cv;
  c.raw_pitch = (int)align_up(c.w, 16);
  c.eq_pitch = (int)align_up(c.w, 128);
  const size_t o_eq = cv.take<uint8_t>((size_t)c.eq_pitch * c.h), o_mask = cv.take<uint8_t>((size_t)c.w * c.h);
  const size_t o_top = cv.take<uint32_t>((size_t)c.ncells * c.K), o_cn = cv.take<int>(c.ncells), o_cnt = cv.take<int>(8);
  const size_t o_kxy = cv.take<float>(2 * (size_t)P), o_kd = cv.take<uint32_t>(8 * (size_t)P);
  const size_t o_nxy = cv.take<float>(2 * (size_t)P), o_nd = cv.take<uint32_t>(8 * (size_t)P), o_nid = cv.take<unsigned long long>(P);
  const size_t o_lxy = cv.take<float>(2 * (size_t)P), o_ld = cv.take<uint32_t>(8 * (size_t)P), o_lid = cv.take<unsigned long long>(P);
  char *b = nullptr;
  CK(cudaMalloc((void **)&b, cv.off + 256));
  CK(cudaMemsetAsync(b, 0, cv.off + 256, ctx.st));
  c.base = b;
  c.eq = (uint8_t *)(b + o_eq);
  c.mask = (uint8_t *)(b + o_mask);
  c.mask_zero = true;
  DescSide &s = c.side;
  s.img = ImgView{c.eq, c.w, c.h, c.eq_pitch};
  s.mask = c.mask;
  s.cam_id = c.cam_id;
  s.cam = c.cam_slot;
  s.w = c.w;
  s.h = c.h;
  s.ncells = c.ncells;
  s.cell_top = (uint32_t *)(b + o_top);
  s.cell_n = (int *)(b + o_cn);
  s.cnt = (int *)(b + o_cnt);
  s.kp_xy = (float *)(b + o_kxy);
  s.kp_desc = (uint32_t *)(b + o_kd);
  s.nw_xy = (float *)(b + o_nxy);
  s.nw_desc = (uint32_t *)(b + o_nd);
  s.nw_ids = (unsigned long long *)(b + o_nid);
  s.la_xy = (float *)(b + o_lxy);
  s.la_desc = (uint32_t *)(b + o_ld);
  s.la_ids = (unsigned long long *)(b + o_lid);
  if (old) {                          // carry pts_last / desc_last / ids_last over
    const int n = c.n_last < oldP ? c.n_last : oldP;
    if (n > 0) {
      CK(cudaMemcpyAsync(s.la_xy, olds.la_xy, sizeof(float) * 2 * n, cudaMemcpyDefault, ctx.st));
      CK(cudaMemcpyAsync(s.la_desc, olds.la_desc, sizeof(uint32_t) * 8 * n, cudaMemcpyDefault, ctx.st));
      CK(cudaMemcpyAsync(s.la_ids, olds.la_ids, sizeof(unsigned long long) * n, cudaMemcpyDefault, ctx.st));
      CK(cudaMemcpyAsync(s.cnt + 3, &c.n_last, sizeof(int), cudaMemcpyHostToDevice, ctx.st));
    }
    CK(cudaStreamSynchronize(ctx.st));
    cudaFree(old);
  }
  c.P = P;
  return 0;
}

int scratch_alloc(DescRig &rig, DescScratch &sc, int P, int grid_words) {
  DescCtx &ctx = *rig.ctx;
  if (sc.base) cudaFree(sc.base);
  if (sc.host_base) cudaFreeHost(sc.host_base);
  sc.base = sc.host_base = nullptr;
  Carver cv;
  const size_t o_grid = cv.take<int>(grid_words), o_q = cv.take<int>(3 * (size_t)P), o_t = cv.take<int>(3 * (size_t)P), o_n = cv.take<int>(4);
  const size_t o_f = cv.take<int>(3 * (size_t)P), o_b = cv.take<int>(3 * (size_t)P);
  const size_t o_p0 = cv.take<float>(6 * (size_t)P + 2), o_p1 = cv.take<float>(6 * (size_t)P + 2), o_m = cv.take<uint8_t>(3 * (size_t)P);
  char *b = nullptr;
  CK(cudaMalloc((void **)&b, cv.off + 256));
  CK(cudaMemsetAsync(b, 0, cv.off + 256, ctx.st));
  sc.base = b;
  sc.grid = (int *)(b + o_grid);
  sc.m_q = (int *)(b + o_q);
  sc.m_t = (int *)(b + o_t);
  sc.m_n = (int *)(b + o_n);
  sc.fwd = (int *)(b + o_f);
  sc.bwd = (int *)(b + o_b);
  sc.p0n = (float *)(b + o_p0);
  sc.p1n = (float *)(b + o_p1);
  sc.rs_mask = (uint8_t *)(b + o_m);
  std::vector<int> inf(grid_words, DV_INF);
  CK(cudaMemcpyAsync(sc.grid, inf.data(), sizeof(int) * grid_words, cudaMemcpyHostToDevice, ctx.st));
  CK(cudaStreamSynchronize(ctx.st));
  char *hb = nullptr;
  const size_t obs_bytes = align_up(sizeof(DvObs) * 2 * (size_t)P, 256);
  CK(cudaHostAlloc((void **)&hb, obs_bytes + 256, cudaHostAllocMapped));
  memset(hb, 0, obs_bytes + 256);
  sc.host_base = hb;
  void *dv = nullptr;
  CK(cudaHostGetDevicePointer(&dv, hb, 0));
  sc.obs = (DvObs *)dv;
  sc.head = (int *)((char *)dv + obs_bytes);
  sc.P = P;
  sc.grid_words = grid_words;
  return 0;
}

struct UnitPlan {                     // one unit of a feed
  DescRig *rig;
  int cam[2];
  int n_cam;
  int scratch;
};

struct Req {
  DescRig *rig;
  const UvioFrame *f;
  int status;
};

bool is_device_ptr(const void *p) {
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return a.type == cudaMemoryTypeDevice;
}

// TrackDescriptor::feed_new_camera for each request; returns the worst status
int feed_batch(DescCtx &ctx, Req *reqs, int n) {
  UVIO_NVTX("uvio_desc_feed_batch");
  std::lock_guard<std::mutex> lk(ctx.mtx);
  if (cudaSetDevice(ctx.device) != cudaSuccess) return UVIO_E_CUDA;
  std::vector<UnitPlan> plan;
  int worst = UVIO_OK;
  auto fail_cuda = [&](int e) {
    ctx.cuda_error = e;
    for (int i = 0; i < n; i++) {
      reqs[i].rig->cuda_error = e;
      reqs[i].status = UVIO_E_CUDA;
    }
    cudaStreamSynchronize(ctx.st);
    return UVIO_E_CUDA;
  };
  // ---- validate, fix sizes, plan units --------------------------------------------------------------------------------------------
  for (int i = 0; i < n; i++) {
    Req &r = reqs[i];
    const UvioFrame &f = *r.f;
    DescRig &rig = *r.rig;
    r.status = UVIO_OK;
    rig.obs.clear();
    if (f.n_img <= 0 || !f.sensor_ids || !f.img || !f.img_stride || !f.width || !f.height) r.status = UVIO_E_BADMSG;   // :36-42
    else if (rig.prm.use_stereo && f.n_img > 2) r.status = UVIO_E_BADMSG;                                            // :57-58
    else if (f.n_img > 4) r.status = UVIO_E_ARG;                         // (a non-stereo tracker takes up to four cameras per message here)
    if (r.status == UVIO_OK)
      for (int k = 0; k < f.n_img; k++) {
        auto it = rig.cam_index.find(f.sensor_ids[k]);
        if (it == rig.cam_index.end()) { r.status = UVIO_E_CAMID; break; }
        if (!f.img[k] || f.width[k] <= 0 || f.height[k] <= 0 || f.img_stride[k] < (size_t)f.width[k]) { r.status = UVIO_E_BADMSG; break; }
        for (int q = 0; q < k; q++)
          if (f.sensor_ids[q] == f.sensor_ids[k]) r.status = UVIO_E_BADMSG;
        DescCam &c = rig.cams[it->second];
        if (c.w && (c.w != f.width[k] || c.h != f.height[k])) { r.status = UVIO_E_SIZE; break; }
        if (f.width[k] >= 4096 * 4 || f.height[k] >= 4096 * 4) { r.status = UVIO_E_ARG; break; }
      }
    if (r.status != UVIO_OK) {
      worst = std::max(worst, r.status);
      continue;
    }
    // geometry and memory of the cameras in this message
    for (int k = 0; k < f.n_img; k++) {
      DescCam &c = rig.cams[rig.cam_index[f.sensor_ids[k]]];
      const bool fresh = c.w == 0;
      c.w = f.width[k];
      c.h = f.height[k];
      const int oldK = c.K, old_cells = c.ncells, old_sx = c.size_x, old_sy = c.size_y;
      cam_geometry(c, rig.prm);
      if (c.size_x <= 0 || c.size_y <= 0 || c.size_x >= 4096 || c.size_y >= 4096) { r.status = UVIO_E_ARG; break; }   // Grider_FAST.h:98-99 asserts
      int P = c.ncells * c.K;
      if (P < 16) P = 16;
      if (P > DESC_MAX_P) { r.status = UVIO_E_ARG; break; }
      if (fresh || P > c.P || c.K != oldK || c.ncells != old_cells || c.size_x != old_sx || c.size_y != old_sy) {
        int e = cam_alloc(rig, c, std::max(P, c.P));
        if (e) return fail_cuda(e);
      }
    }
    if (r.status != UVIO_OK) {
      worst = std::max(worst, r.status);
      continue;
    }
    const bool stereo = f.n_img == 2 && rig.prm.use_stereo;                                                         // :47-55
    const int n_units = stereo ? 1 : f.n_img;
    for (int u = 0; u < n_units; u++) {
      UnitPlan p;
      p.rig = &rig;
      p.n_cam = stereo ? 2 : 1;
      p.cam[0] = rig.cam_index[f.sensor_ids[stereo ? 0 : u]];
      p.cam[1] = stereo ? rig.cam_index[f.sensor_ids[1]] : -1;
      p.scratch = u;
      // both cameras of a pair share one capacity
      int P = rig.cams[p.cam[0]].P;
      if (stereo) {
        int P1 = rig.cams[p.cam[1]].P;
        if (P1 != P) {
          P = std::max(P, P1);
          for (int k = 0; k < 2; k++)
            if (rig.cams[p.cam[k]].P != P) {
              int e = cam_alloc(rig, rig.cams[p.cam[k]], P);
              if (e) return fail_cuda(e);
            }
        }
      }
      const DescCam &c0 = rig.cams[p.cam[0]];
      const int gw = (int)((float)c0.w / (float)rig.prm.min_px_dist), gh = (int)((float)c0.h / (float)rig.prm.min_px_dist);
      DescScratch &sc = rig.scratch[u];
      if (sc.P < P || sc.grid_words < gw * gh + 1) {
        int e = scratch_alloc(rig, sc, P, gw * gh + 1);
        if (e) return fail_cuda(e);
      }
      plan.push_back(p);
    }
  }
  if (plan.empty()) return worst;
  // ---- lay the launch tables out ------------------------------------------------------------------------------------------------------
  const int n_units = (int)plan.size();
  int n_sides = 0, n_jobs = 0, slots = 0, max_cells = 0, max_p = 0, max_cw = 0, max_ch = 0, max_band = 0, any_clahe = 0;
  size_t corner_words = 0, raw_bytes = 0;
  for (const UnitPlan &p : plan)
    for (int k = 0; k < p.n_cam; k++) {
      const DescCam &c = p.rig->cams[p.cam[k]];
      n_sides++;
      n_jobs += c.ncells;
      slots += c.P;
      corner_words += (size_t)c.ncells * c.cell_cap;
      raw_bytes += (size_t)c.raw_pitch * c.h;
      max_cells = std::max(max_cells, c.ncells);
      max_p = std::max(max_p, c.P);
      max_cw = std::max(max_cw, c.size_x);
      max_ch = std::max(max_ch, c.size_y);
      max_band = std::max(max_band, ((c.h + 7) / 8 + 1) * c.w);
      if (p.rig->prm.histogram_method == UVIO_HIST_CLAHE) any_clahe = 1;
    }
  Carver tab;                         // tables uploaded from the staging buffer: one copy
  const size_t o_pre = tab.take<PreJob>(n_sides), o_pyr = tab.take<PyrDesc>(n_sides), o_jobs = tab.take<FastJob>(n_jobs), o_units = tab.take<DescUnit>(n_units);
  const size_t tab_bytes = align_up(tab.off, 256);
  Carver ar;
  ar.off = tab_bytes;
  const size_t o_cnt = ar.take<unsigned>(2 * (size_t)n_jobs + 4);         // FAST counters + tickets + n_sp: zeroed every feed
  const size_t zero_bytes = ar.off - o_cnt;
  const size_t o_corners = ar.take<uint32_t>(corner_words), o_items = ar.take<SubpixItem>(slots), o_spout = ar.take<float>(2 * (size_t)slots);
  const size_t o_clahe = ar.take<uint8_t>(any_clahe ? (size_t)n_sides * 64 * 256 : 1);
  const size_t o_raw = ar.take<uint8_t>(raw_bytes);      // the uploaded frames, in batch order: adjacent host frames go up as one copy
  if (ar.off + 256 > ctx.arena_bytes) {
    if (ctx.arena) cudaFree(ctx.arena);
    ctx.arena = nullptr;
    ctx.arena_bytes = 0;
    const size_t want = (ar.off + 256) * 3 / 2;
    cudaError_t e = cudaMalloc(&ctx.arena, want);
    if (e != cudaSuccess) return fail_cuda((int)e);
    ctx.arena_bytes = want;
  }
  if (tab_bytes > ctx.stage_bytes) {
    if (ctx.stage) cudaFreeHost(ctx.stage);
    ctx.stage = nullptr;
    ctx.stage_bytes = 0;
    cudaError_t e = cudaHostAlloc(&ctx.stage, tab_bytes * 2, cudaHostAllocDefault);
    if (e != cudaSuccess) return fail_cuda((int)e);
    ctx.stage_bytes = tab_bytes * 2;
  }
  char *A = (char *)ctx.arena, *H = (char *)ctx.stage;
  memset(H, 0, tab_bytes);
  PreJob *h_pre = (PreJob *)(H + o_pre);
  PyrDesc *h_pyr = (PyrDesc *)(H + o_pyr);
  FastJob *h_jobs = (FastJob *)(H + o_jobs);
  DescUnit *h_units = (DescUnit *)(H + o_units);
  DescGlobals X;
  X.fast_jobs = (const FastJob *)(A + o_jobs);
  X.fast_cnt = (unsigned *)(A + o_cnt);
  unsigned *ticket = X.fast_cnt + n_jobs;
  X.n_sp = (int *)(X.fast_cnt + 2 * n_jobs);
  X.fast_out = (uint32_t *)(A + o_corners);
  X.sp_items = (SubpixItem *)(A + o_items);
  X.sp_out = (float *)(A + o_spout);
  X.cams = ctx.cams_dev;
  cudaStream_t st = ctx.st;
#define CKF(expr)                                        \
  do {                                                   \
    cudaError_t e_ = (expr);                             \
    if (e_ != cudaSuccess) return fail_cuda((int)e_);    \
  } while (0)
  CKF(cudaEventRecord(ctx.ev0, st));
  // ---- uploads and tables --------------------------------------------------------------------------------------------------------------
  int side_i = 0, job_i = 0, slot_i = 0;
  size_t corner_i = 0, raw_i = 0;
  // copies whose source and destination both continue the previous one are merged (a batch of tightly packed frames is one copy)
  uint8_t *pend_dst = nullptr;
  const uint8_t *pend_src = nullptr;
  size_t pend_bytes = 0;
  auto flush = [&]() -> cudaError_t {
    cudaError_t e = pend_bytes ? cudaMemcpyAsync(pend_dst, pend_src, pend_bytes, cudaMemcpyHostToDevice, st) : cudaSuccess;
    pend_bytes = 0;
    return e;
  };
  std::vector<const UvioFrame *> unit_frame(n_units);
  {
    int u = 0;
    for (int i = 0; i < n; i++) {
      if (reqs[i].status != UVIO_OK) continue;
      const bool stereo = reqs[i].f->n_img == 2 && reqs[i].rig->prm.use_stereo;
      const int k = stereo ? 1 : reqs[i].f->n_img;
      for (int q = 0; q < k; q++) unit_frame[u++] = reqs[i].f;
    }
  }
  for (int u = 0; u < n_units; u++) {
    const UnitPlan &p = plan[u];
    DescRig &rig = *p.rig;
    const UvioFrame &f = *unit_frame[u];
    DescScratch &sc = rig.scratch[p.scratch];
    DescUnit &U = h_units[u];
    U.stereo = p.n_cam == 2;
    U.P = rig.cams[p.cam[0]].P;
    U.K = rig.cams[p.cam[0]].K;
    U.mpd = rig.prm.min_px_dist;
    U.knn_ratio = rig.prm.knn_ratio;
    U.currid = rig.currid_dev;
    U.id_base = rig.currid;
    U.last_of_rig = (u + 1 == n_units || plan[u + 1].rig != p.rig) ? 1 : 0;
    {
      int ns = 0;
      for (int j = u - 1; j >= 0 && plan[j].rig == p.rig && ns < 3; j--) ns++;          // units of this tracker before this one
      for (int k = 0; k < 3; k++) U.sib_cnt[k] = k < ns ? p.rig->cams[plan[u - ns + k].cam[0]].side.cnt : nullptr;
    }
    U.grid = sc.grid;
    U.m_q = sc.m_q;
    U.m_t = sc.m_t;
    U.m_n = sc.m_n;
    U.fwd = sc.fwd;
    U.bwd = sc.bwd;
    U.p0n = sc.p0n;
    U.p1n = sc.p1n;
    U.rs_mask = sc.rs_mask;
    U.obs = sc.obs;
    U.head = sc.head;
    for (int k = 0; k < p.n_cam; k++) {
      DescCam &c = rig.cams[p.cam[k]];
      // which image of the message is this camera's
      int mi = 0;
      for (int q = 0; q < f.n_img; q++)
        if (f.sensor_ids[q] == c.cam_id) mi = q;
      // frame
      ImgView src;
      if (is_device_ptr(f.img[mi])) {
        src = ImgView{f.img[mi], c.w, c.h, (int)f.img_stride[mi]};
      } else {
        uint8_t *raw = (uint8_t *)(A + o_raw) + raw_i;
        const size_t span = (size_t)c.raw_pitch * c.h;
        if ((size_t)c.raw_pitch == f.img_stride[mi] && c.raw_pitch == c.w) {
          if (pend_bytes && pend_src + pend_bytes == f.img[mi] && pend_dst + pend_bytes == raw) {
            pend_bytes += span;
          } else {
            CKF(flush());
            pend_dst = raw;
            pend_src = f.img[mi];
            pend_bytes = span;
          }
        } else {
          CKF(flush());
          CKF(cudaMemcpy2DAsync(raw, c.raw_pitch, f.img[mi], f.img_stride[mi], c.w, c.h, cudaMemcpyHostToDevice, st));
        }
        src = ImgView{raw, c.w, c.h, c.raw_pitch};
        raw_i += span;
      }
      // mask (TrackDescriptor.cpp:83, :204-205: the message's mask is the one Grider_FAST tests)
      const uint8_t *m = f.mask ? f.mask[mi] : nullptr;
      if (m) {
        const size_t ms = f.mask_stride ? f.mask_stride[mi] : (size_t)c.w;
        CKF(cudaMemcpy2DAsync(c.mask, c.w, m, ms, c.w, c.h, cudaMemcpyDefault, st));
        c.mask_zero = false;
      } else if (!c.mask_zero) {
        CKF(cudaMemsetAsync(c.mask, 0, (size_t)c.w * c.h, st));
        c.mask_zero = true;
      }
      h_pre[side_i].src = src;
      h_pre[side_i].pyr = side_i;
      h_pre[side_i].method = rig.prm.histogram_method;
      h_pyr[side_i].lv[0] = c.side.img;
      h_pyr[side_i].nlv = 1;
      DescSide &S = U.s[k];
      S = c.side;
      S.pyr = side_i;
      S.job0 = job_i;
      S.slot0 = slot_i;
      for (int r = 0; r < c.ncells; r++) {
        FastJob j = c.jobs[r];
        j.img = side_i;
        j.out_off = (int)(corner_i + (size_t)r * c.cell_cap);
        j.threshold = rig.prm.fast_threshold;
        h_jobs[job_i + r] = j;
      }
      job_i += c.ncells;
      slot_i += c.P;
      corner_i += (size_t)c.ncells * c.cell_cap;
      side_i++;
    }
    const DescCam &c0 = rig.cams[p.cam[0]];
    const DescCam &c1 = rig.cams[p.cam[p.n_cam == 2 ? 1 : 0]];
    U.gw[0] = (int)((float)c0.w / (float)U.mpd);                           // :374 / :436-439
    U.gh[0] = (int)((float)c0.h / (float)U.mpd);
    U.gw[1] = (int)((float)c1.w / (float)U.mpd);
    U.gh[1] = (int)((float)c1.h / (float)U.mpd);
    U.thr[0] = 1.0 / std::max(c0.max_focal, c1.max_focal);                 // :530-533
    U.thr[1] = 1.0 / c0.max_focal;
    U.thr[2] = 1.0 / c1.max_focal;
  }
  CKF(flush());
  if (corner_i >= (size_t)0x7fffffff) return fail_cuda((int)cudaErrorInvalidValue);
  CKF(cudaMemcpyAsync(A, H, tab_bytes, cudaMemcpyHostToDevice, st));
  CKF(cudaMemsetAsync(A + o_cnt, 0, zero_bytes, st));
  const PreJob *d_pre = (const PreJob *)(A + o_pre);
  const PyrDesc *d_pyr = (const PyrDesc *)(A + o_pyr);
  const DescUnit *d_units = (const DescUnit *)(A + o_units);
  // ---- the chain -------------------------------------------------------------------------------------------------------------------------
  CKF(launch_preprocess(st, d_pre, n_sides, d_pyr, (uint8_t *)(A + o_clahe), any_clahe, max_band));
  // k_fast hands each cell's list and count to a second pair of pointers (mapped host memory in the KLT host pipeline): here both
  // pairs name the same device arrays, the lists stay where they were appended and the count is rewritten in place
  CKF(launch_fast(st, X.fast_jobs, n_jobs, max_cw, max_ch, d_pyr, X.fast_out, X.fast_cnt, ticket, X.fast_out, X.fast_cnt));
  CKF(launch_desc_topk(st, d_units, n_units, max_cells, X));
  CKF(launch_desc_gather(st, d_units, n_units, X));
  CKF(launch_subpix(st, X.sp_items, slots, d_pyr, ctx.sp_mask, X.sp_out, nullptr, X.n_sp));
  CKF(launch_desc_orb(st, d_units, n_units, max_p, X));
  CKF(launch_desc_detect(st, d_units, n_units, max_p, X));
  CKF(launch_desc_temporal(st, d_units, n_units, max_p, X));
  CKF(launch_desc_merge(st, d_units, n_units, max_p, X));
  CKF(cudaEventRecord(ctx.ev1, st));
  CKF(cudaStreamSynchronize(st));
  float ms = 0.f;
  cudaEventElapsedTime(&ms, ctx.ev0, ctx.ev1);
  // ---- results ----------------------------------------------------------------------------------------------------------------------------
  for (int u = 0; u < n_units; u++) {
    const UnitPlan &p = plan[u];
    DescRig &rig = *p.rig;
    DescScratch &sc = rig.scratch[p.scratch];
    const int *head = (const int *)((const char *)sc.host_base + align_up(sizeof(DvObs) * 2 * (size_t)sc.P, 256));
    const DvObs *ob = (const DvObs *)sc.host_base;
    const int n_obs = head[0];
    for (int k = 0; k < p.n_cam; k++) rig.cams[p.cam[k]].n_last = head[1];
    if (u + 1 == n_units || plan[u + 1].rig != p.rig)
      rig.currid = (unsigned long long)(unsigned)head[2] | ((unsigned long long)(unsigned)head[3] << 32);   // (the last unit's: all of the tracker's new ids)
    for (int i = 0; i < n_obs; i++) {
      UvioObs o;
      o.id = (size_t)ob[i].id;
      o.cam_id = (size_t)ob[i].cam_id;
      o.u = ob[i].u;
      o.v = ob[i].v;
      o.un = ob[i].un;
      o.vn = ob[i].vn;
      rig.obs.push_back(o);
    }
    rig.last_us = ms * 1e3f;
  }
  return worst;
}

}  // namespace

struct uvio_desc {
  DescRig *rig;
};

#define GUARD_BEGIN try {
#define GUARD_END                      \
  }                                    \
  catch (...) {                        \
    return UVIO_E_ARG;                 \
  }

extern "C" {

void uvio_desc_default_params(UvioDescParams *p) {
  if (!p) return;
  p->num_features = 200;
  p->num_aruco = 0;
  p->use_stereo = 1;
  p->histogram_method = UVIO_HIST_HISTOGRAM;
  p->fast_threshold = 20;
  p->grid_x = 5;
  p->grid_y = 5;
  p->min_px_dist = 10;
  p->knn_ratio = 0.70;
}

int uvio_desc_create(const UvioDescParams *p, const UvioCamCalib *cams, int n_cams, int cuda_device, uvio_desc **out) {
  GUARD_BEGIN
  if (!p || !cams || n_cams <= 0 || !out) return UVIO_E_ARG;
  if (p->num_features <= 0 || p->grid_x <= 0 || p->grid_y <= 0 || p->min_px_dist <= 0 || p->histogram_method < 0 || p->histogram_method > 2 ||
      !(p->knn_ratio > 0))
    return UVIO_E_ARG;
  *out = nullptr;
  for (int i = 0; i < n_cams; i++) {
    if (cams[i].model != UVIO_CAM_RADTAN && cams[i].model != UVIO_CAM_EQUI) return UVIO_E_ARG;
    for (int k = 0; k < i; k++)
      if (cams[k].cam_id == cams[i].cam_id) return UVIO_E_ARG;
  }
  DescCtx *ctx = ctx_for_device(cuda_device);
  if (!ctx) return UVIO_E_CUDA;       // no device, no tracker: there is deliberately no CPU path behind this ABI
  std::lock_guard<std::mutex> lk(ctx->mtx);
  if (cudaSetDevice(ctx->device) != cudaSuccess) return UVIO_E_CUDA;
  DescRig *rig = new DescRig;
  rig->ctx = ctx;
  rig->prm = *p;
  rig->cams.resize(n_cams);
  rig->scratch.resize(n_cams);
  for (int i = 0; i < n_cams; i++) {
    DescCam &c = rig->cams[i];
    c.cam_id = cams[i].cam_id;
    c.model.model = cams[i].model;
    for (int k = 0; k < 4; k++) {
      c.model.k[k] = cams[i].intrinsics[k];
      c.model.d[k] = cams[i].distortion[k];
    }
    c.max_focal = std::max(cams[i].intrinsics[0], cams[i].intrinsics[1]);
    if (!ctx->cam_free.empty()) {
      c.cam_slot = ctx->cam_free.back();
      ctx->cam_free.pop_back();
    } else if (ctx->cam_used < DescCtx::CAM_CAP) {
      c.cam_slot = ctx->cam_used++;
    } else {
      delete rig;
      return UVIO_E_ARG;
    }
    if (cudaMemcpy(ctx->cams_dev + c.cam_slot, &c.model, sizeof(CamModel), cudaMemcpyHostToDevice) != cudaSuccess) {
      delete rig;
      return UVIO_E_CUDA;
    }
    rig->cam_index[c.cam_id] = i;
  }
  const unsigned long long id0 = 4ull * (unsigned long long)p->num_aruco + 1;   // TrackBase.cpp:34
  rig->currid = id0;
  if (cudaMalloc((void **)&rig->currid_dev, 256) != cudaSuccess ||
      cudaMemcpy(rig->currid_dev, &id0, sizeof(id0), cudaMemcpyHostToDevice) != cudaSuccess) {
    delete rig;
    return UVIO_E_CUDA;
  }
  uvio_desc *h = new uvio_desc;
  h->rig = rig;
  *out = h;
  return UVIO_OK;
  GUARD_END
}

void uvio_desc_destroy(uvio_desc *h) {
  if (!h) return;
  DescRig *rig = h->rig;
  {
    std::lock_guard<std::mutex> lk(rig->ctx->mtx);
    cudaSetDevice(rig->ctx->device);
    for (DescCam &c : rig->cams) {
      if (c.base) cudaFree(c.base);
      if (c.cam_slot >= 0) rig->ctx->cam_free.push_back(c.cam_slot);
    }
    for (DescScratch &s : rig->scratch) {
      if (s.base) cudaFree(s.base);
      if (s.host_base) cudaFreeHost(s.host_base);
    }
    if (rig->currid_dev) cudaFree(rig->currid_dev);
  }
  delete rig;
  delete h;
}

int uvio_desc_feed_batch(uvio_desc *const *hs, UvioFrame *frames, int n) {
  GUARD_BEGIN
  if (!hs || !frames || n <= 0) return UVIO_E_ARG;
  DescCtx *ctx = hs[0] ? hs[0]->rig->ctx : nullptr;
  for (int i = 0; i < n; i++) {
    if (!hs[i] || hs[i]->rig->ctx != ctx) return UVIO_E_ARG;
    for (int k = 0; k < i; k++)
      if (hs[k] == hs[i]) return UVIO_E_ARG;
  }
  std::vector<DescRig *> order(n);
  for (int i = 0; i < n; i++) order[i] = hs[i]->rig;
  std::sort(order.begin(), order.end(), [](DescRig *a, DescRig *b) { return std::less<DescRig *>()(a, b); });   // address order: no ABBA
  std::vector<std::unique_lock<std::mutex>> locks;
  for (DescRig *r : order) locks.emplace_back(r->mtx);
  std::vector<Req> reqs(n);
  for (int i = 0; i < n; i++) reqs[i] = Req{hs[i]->rig, &frames[i], UVIO_OK};
  int rc = feed_batch(*ctx, reqs.data(), n);
  for (int i = 0; i < n; i++) frames[i].status = reqs[i].status;
  return rc;
  GUARD_END
}

int uvio_desc_feed(uvio_desc *h, const UvioFrame *frame) {
  GUARD_BEGIN
  if (!h || !frame) return UVIO_E_ARG;
  std::lock_guard<std::mutex> lk(h->rig->mtx);
  Req r{h->rig, frame, UVIO_OK};
  feed_batch(*h->rig->ctx, &r, 1);
  return r.status;
  GUARD_END
}

int uvio_desc_get_new_observations(uvio_desc *h, UvioObs *out, int cap, int *n) {
  GUARD_BEGIN
  if (!h || !n || cap < 0 || (cap > 0 && !out)) return UVIO_E_ARG;
  std::lock_guard<std::mutex> lk(h->rig->mtx);
  const int have = (int)h->rig->obs.size();
  *n = have;
  if (have > cap) return UVIO_E_SIZE;
  if (have) memcpy(out, h->rig->obs.data(), sizeof(UvioObs) * have);
  return UVIO_OK;
  GUARD_END
}

static int get_last_common(uvio_desc *h, size_t cam_id, size_t *ids, float *uv, uint8_t *desc, int cap, int *n) {
  if (!h || !n || cap < 0) return UVIO_E_ARG;
  DescRig &rig = *h->rig;
  std::lock_guard<std::mutex> lk(rig.mtx);
  auto it = rig.cam_index.find(cam_id);
  if (it == rig.cam_index.end()) return UVIO_E_CAMID;
  DescCam &c = rig.cams[it->second];
  *n = c.n_last;
  if (c.n_last > cap) return UVIO_E_SIZE;
  if (c.n_last == 0) return UVIO_OK;
  std::lock_guard<std::mutex> lc(rig.ctx->mtx);
  if (cudaSetDevice(rig.ctx->device) != cudaSuccess) return UVIO_E_CUDA;
  cudaError_t e = cudaSuccess;
  if (ids) {
    std::vector<unsigned long long> tmp(c.n_last);
    e = cudaMemcpy(tmp.data(), c.side.la_ids, sizeof(unsigned long long) * c.n_last, cudaMemcpyDeviceToHost);
    for (int i = 0; i < c.n_last; i++) ids[i] = (size_t)tmp[i];
  }
  if (e == cudaSuccess && uv) e = cudaMemcpy(uv, c.side.la_xy, sizeof(float) * 2 * c.n_last, cudaMemcpyDeviceToHost);
  if (e == cudaSuccess && desc) e = cudaMemcpy(desc, c.side.la_desc, 32 * (size_t)c.n_last, cudaMemcpyDeviceToHost);
  if (e != cudaSuccess) {
    rig.cuda_error = (int)e;
    return UVIO_E_CUDA;
  }
  return UVIO_OK;
}

int uvio_desc_get_last(uvio_desc *h, size_t cam_id, size_t *ids, float *uv, int cap, int *n) {
  GUARD_BEGIN
  return get_last_common(h, cam_id, ids, uv, nullptr, cap, n);
  GUARD_END
}

int uvio_desc_get_last_descriptors(uvio_desc *h, size_t cam_id, uint8_t *desc, int cap, int *n) {
  GUARD_BEGIN
  return get_last_common(h, cam_id, nullptr, nullptr, desc, cap, n);
  GUARD_END
}

int uvio_desc_set_num_features(uvio_desc *h, int n) {
  GUARD_BEGIN
  if (!h || n <= 0) return UVIO_E_ARG;
  std::lock_guard<std::mutex> lk(h->rig->mtx);
  h->rig->prm.num_features = n;       // the extraction grid is re-derived at the next feed (Grider_FAST.h:83-103)
  return UVIO_OK;
  GUARD_END
}

int uvio_desc_change_feat_id(uvio_desc *h, size_t id_old, size_t id_new) {
  GUARD_BEGIN
  if (!h) return UVIO_E_ARG;
  DescRig &rig = *h->rig;
  std::lock_guard<std::mutex> lk(rig.mtx);
  std::lock_guard<std::mutex> lc(rig.ctx->mtx);
  if (cudaSetDevice(rig.ctx->device) != cudaSuccess) return UVIO_E_CUDA;
  for (DescCam &c : rig.cams) {                       // TrackBase.cpp:241-247: every camera's ids_last
    if (c.n_last <= 0 || !c.base) continue;
    std::vector<unsigned long long> ids(c.n_last);
    cudaError_t e = cudaMemcpy(ids.data(), c.side.la_ids, sizeof(unsigned long long) * c.n_last, cudaMemcpyDeviceToHost);
    bool any = false;
    for (auto &v : ids)
      if (v == (unsigned long long)id_old) { v = (unsigned long long)id_new; any = true; }
    if (e == cudaSuccess && any) e = cudaMemcpy(c.side.la_ids, ids.data(), sizeof(unsigned long long) * c.n_last, cudaMemcpyHostToDevice);
    if (e != cudaSuccess) { rig.cuda_error = (int)e; return UVIO_E_CUDA; }
  }
  return UVIO_OK;
  GUARD_END
}

int uvio_desc_cuda_error(uvio_desc *h) { return h ? h->rig->cuda_error : 0; }

int uvio_desc_get_timing(uvio_desc *h, float *us) {
  if (!h || !us) return UVIO_E_ARG;
  *us = h->rig->last_us;
  return UVIO_OK;
}

}  // extern "C"
