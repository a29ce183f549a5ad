// The device-side pipeline (default): a feed is one chain of launches with the bookkeeping in track_body.h kernels and a
// single synchronisation at its end.  Device state of a tracker (DvRig), the per-feed descriptors, issue and collect.
#include "engine_internal.h"

namespace uvio {

void dv_destroy(DvRig *d) { delete d; }

bool device_pipeline_enabled() {   // UVIO_KLT_PIPELINE=host selects the round-1 path (bookkeeping on host threads)
  static const bool v = [] {
    const char *e = getenv("UVIO_KLT_PIPELINE");
    return !(e && strcmp(e, "host") == 0);
  }();
  return v;
}

template <class T>
static T *dv_carve(char *&p, size_t n) {
  T *r = (T *)p;
  p += (n * sizeof(T) + 255) / 256 * 256;
  return r;
}

// (re)builds the tracker's device memory when its feature budget or image sizes changed; uploads pts_last / ids_last / currid
// when the host copy is newer (first use, set_num_features, change_feat_id, a feed through the host path)
int dv_prepare(Context *ctx, Rig &rig) {
  const Params &P = rig.prm;
  if (rig.dv_mirror_stale && (rig.dv_dirty || rig.dv == nullptr || rig.dv->nf != P.num_features)) dv_sync_mirror(rig);
  bool rebuild = rig.dv == nullptr || rig.dv->nf != P.num_features || rig.dv->cam_w.size() != rig.cams.size();
  if (!rebuild)
    for (size_t k = 0; k < rig.cams.size(); k++) rebuild |= rig.dv->cam_w[k] != rig.cams[k].w || rig.dv->cam_h[k] != rig.cams[k].h;
  if (rebuild) {
    CU_TRY(cudaDeviceSynchronize());   // nothing of this tracker may be in flight while its memory moves
    delete rig.dv;
    rig.dv = new DvRig;
    DvRig &D = *rig.dv;
    D.nf = P.num_features;
    D.C = P.grid_x * P.grid_y;
    int max_w = 1, max_h = 1;
    for (auto &c : rig.cams) { max_w = std::max(max_w, c.w); max_h = std::max(max_h, c.h); D.cam_w.push_back(c.w); D.cam_h.push_back(c.h); }
    const GridGeom geom = grider_geom(max_w, max_h, P.num_features, P.grid_x, P.grid_y);
    D.K = geom.per_cell;
    D.size_x = geom.size_x;   // (per camera at use; the largest here sizes the FAST output)
    D.size_y = geom.size_y;
    int cap = std::max(1, ((geom.size_x + 1) / 2) * ((geom.size_y + 1) / 2));
    D.cell_cap = 1;
    while (D.cell_cap < cap) D.cell_cap <<= 1;   // a power of two: a cell's top-K block may bitonic-sort in place
    D.P = P.num_features + 2 * D.C * D.K + 64;
    D.obs_cap = 2 * D.P;
    const size_t ncam = rig.cams.size(), scr = (size_t)std::max(D.P, D.C * D.K) + D.C + 16;
    D.cam.resize(ncam);
    D.unit.resize(ncam);
    // the same walk lays the two slabs out twice: once from address 0 to size them, once over the real allocations
    auto layout = [&](char *p, char *hp, ptrdiff_t to_dev, size_t &dev_bytes, size_t &host_bytes) {
      char *p0 = p, *hp0 = hp;
      D.currid = dv_carve<unsigned long long>(p, 1);
      D.h_currid = dv_carve<unsigned long long>(hp, 1);
      D.m_currid = (unsigned long long *)((char *)D.h_currid + to_dev);
      for (size_t k = 0; k < ncam; k++) {
        DvCamMem &m = D.cam[k];
        const CamState &c = rig.cams[k];
        const size_t cells = (size_t)(c.w / std::max(1, P.min_px_dist) + 1) * (size_t)(c.h / std::max(1, P.min_px_dist) + 1);
        m.pts_last = dv_carve<float>(p, 2 * (size_t)D.P);
        m.ids_last = dv_carve<unsigned long long>(p, D.P);
        m.n_last = dv_carve<int>(p, 1);
        m.pts = dv_carve<float>(p, 2 * (size_t)D.P);
        m.ids = dv_carve<unsigned long long>(p, D.P);
        m.close = dv_carve<uint8_t>(p, cells);
        m.first = dv_carve<int>(p, cells);
        m.grid = dv_carve<int>(p, D.C);
        m.cells = dv_carve<int>(p, D.C);
        m.top = dv_carve<uint32_t>(p, (size_t)D.C * D.K);
        m.top_n = dv_carve<int>(p, D.C);
        m.ctl = dv_carve<int>(p, 16);
        m.h_pts_last = dv_carve<float>(hp, 2 * (size_t)D.P);
        m.h_ids_last = dv_carve<unsigned long long>(hp, D.P);
        m.h_n_last = dv_carve<int>(hp, 1);
        m.m_pts_last = (float *)((char *)m.h_pts_last + to_dev);
        m.m_ids_last = (unsigned long long *)((char *)m.h_ids_last + to_dev);
        m.m_n_last = (int *)((char *)m.h_n_last + to_dev);
        for (int q = 0; q < PYR_SLOTS; q++) m.mask[q] = dv_carve<uint8_t>(p, (size_t)c.w * c.h);
      }
      for (size_t k = 0; k < ncam; k++) {
        DvUnitMem &u = D.unit[k];
        for (int q = 0; q < 4; q++) u.scr[q] = dv_carve<int>(p, scr);
        u.part = dv_carve<int>(p, 1024);
        u.uctl = dv_carve<int>(p, 16);
        u.tmp_pts = dv_carve<float>(p, 2 * (size_t)D.P);
        u.tmp_ids = dv_carve<unsigned long long>(p, D.P);
        u.d_obs = dv_carve<DvObs>(p, D.obs_cap);
        u.h_obs = dv_carve<DvObs>(hp, D.obs_cap);
        u.h_n_obs = dv_carve<int>(hp, 1);
        u.m_obs = (DvObs *)((char *)u.h_obs + to_dev);
        u.m_n_obs = (int *)((char *)u.h_n_obs + to_dev);
      }
      dev_bytes = (size_t)(p - p0);
      host_bytes = (size_t)(hp - hp0);
    };
    size_t dev_bytes = 0, host_bytes = 0, chk_d = 0, chk_h = 0;
    layout(nullptr, nullptr, 0, dev_bytes, host_bytes);
    CU_TRY(cudaMalloc(&D.slab, dev_bytes + 256));
    CU_TRY(cudaMemset(D.slab, 0, dev_bytes + 256));
    CU_TRY(cudaHostAlloc(&D.host_slab, host_bytes + 256, cudaHostAllocMapped));
    memset(D.host_slab, 0, host_bytes + 256);
    char *hd = nullptr;
    CU_TRY(cudaHostGetDevicePointer((void **)&hd, D.host_slab, 0));
    layout((char *)D.slab, (char *)D.host_slab, hd - (char *)D.host_slab, chk_d, chk_h);
    rig.dv_dirty = true;
  }
  if (rig.dv_dirty) {
    DvRig &D = *rig.dv;
    CU_TRY(cudaDeviceSynchronize());
    for (auto &c : rig.cams) c.dv_n_last = -1;     // the host vectors are the truth now (dv_sync_mirror ran before they were edited)
    const unsigned long long id = (unsigned long long)rig.currid;
    CU_TRY(cudaMemcpy(D.currid, &id, sizeof(id), cudaMemcpyHostToDevice));
    for (size_t k = 0; k < rig.cams.size(); k++) {
      const CamState &c = rig.cams[k];
      const int n = (int)std::min(c.pts_last.size(), (size_t)D.P);
      if (n) {
        CU_TRY(cudaMemcpy(D.cam[k].pts_last, c.pts_last.data(), sizeof(Pt) * (size_t)n, cudaMemcpyHostToDevice));
        std::vector<unsigned long long> ids(c.ids_last.begin(), c.ids_last.begin() + n);
        CU_TRY(cudaMemcpy(D.cam[k].ids_last, ids.data(), 8 * (size_t)n, cudaMemcpyHostToDevice));
      }
      CU_TRY(cudaMemcpy(D.cam[k].n_last, &n, sizeof(int), cudaMemcpyHostToDevice));
    }
    rig.dv_dirty = false;
  }
  return OK;
}

// ---------------------------------------------------------------------------------------------------------------------
// The device-side pipeline: one feed_batch call = one chain of launches on the lane's stream, bookkeeping included
// (track_body.h), one synchronisation at the end.  All trackers of the call go through one lane.
//   up_stream : H2D of the new frames -> k_preprocess                                  (announced frames: already under way)
//   stream    : k_dv_occupancy -> k_dv_fast -> k_dv_topk -> k_dv_select -> k_subpix -> k_lk (stereo) -> k_dv_assign
//               -> [event: new pyramids] -> k_lk (temporal) -> k_ransac -> k_dv_merge
// The host then copies the observation lists and the pts_last / ids_last mirrors out of mapped memory.
// ---------------------------------------------------------------------------------------------------------------------
int Pass::device_feed() {
  UVIO_NVTX("device_feed: descriptors + launch chain");
  Context *dev = ctx->dev;
  int e = stage0_book();
  if (e != OK) return e;
  cudaStream_t st = ctx->stream;
  // ---- per-tracker device memory; unit descriptors ---------------------------------------------------------------------------
  for (int i = 0; i < n; i++) {
    e = dv_prepare(dev, *reqs[i]->rig);
    if (e != OK) { ctx->cuda_err = dev->cuda_err; return e; }
  }
  const int nu = (int)units.size();
  if (nu > DV_MAX_UNITS) return E_ARG;   // feed_batch cuts bigger batches into more groups
  DvFeedBits feed_bits;
  {
    const DvUnit *du = ctx->dv_units.d;
    const DvHead *dh = ctx->dv_heads.d;
    CU_TRY(ctx->dv_units.reserve((size_t)std::max(nu, 1)));
    CU_TRY(ctx->dv_heads.reserve((size_t)std::max(n, 1)));
    if (du != ctx->dv_units.d) ctx->dv_units_sent.clear();   // the device copy moved: whatever it held is gone
    if (dh != ctx->dv_heads.d) ctx->dv_heads_sent.clear();
  }
  int job_slots = 0, sp_slots = 0, lk_slots = 0, max_c = 1, max_cw = 1, max_ch = 1, max_p = 1, nheads = 0;
  size_t fast_words = 0;
  bool any_first = false;
  {
    int k = 0;
    Rig *prev = nullptr;
    for (auto &u : units) {
      Rig &rig = *u.rig;
      DvRig &D = *rig.dv;
      const Params &P = rig.prm;
      DvUnit &U = ctx->dv_units.h[k];
      memset(&U, 0, sizeof(U));
      const int unit_idx = u.stereo ? 0 : u.L.msg;
      U.stereo = u.stereo;
      unsigned fb = (unsigned)(u.first && !u.failed) | ((unsigned)rig.dv_wants_mirror << 1);
      any_first |= u.first;
      U.num_features = P.num_features;
      U.gx = P.grid_x; U.gy = P.grid_y; U.mpd = P.min_px_dist; U.fast_threshold = P.fast_threshold;
      const GridGeom geom = grider_geom(u.L.cam->w, u.L.cam->h, P.num_features, P.grid_x, P.grid_y);
      U.size_x = geom.size_x; U.size_y = geom.size_y; U.per_cell = geom.per_cell;
      U.P = D.P; U.C = D.C; U.K = D.K;
      U.job_base = job_slots; U.fast_base = (int)fast_words; U.cell_cap = D.cell_cap;
      U.sp_base = sp_slots; U.spr_base = -1 - sp_slots; U.lk_base = lk_slots; U.lkn_base = 2 * sp_slots; U.rs_base = 2 * k;   // spr_base is fixed up below
      U.currid = D.currid;
      U.h_currid = D.m_currid;
      for (int side = 0; side < (u.stereo ? 2 : 1); side++) {
        Side &s = side ? u.R : u.L;
        CamState &c = *s.cam;
        const int ci = (int)(s.cam - rig.cams.data());
        DvCamMem &m = D.cam[ci];
        DvSide &S = U.s[side];
        S.w = c.w; S.h = c.h;
        const int det_slot = u.first ? c.cur : c.last();
        for (int q = 0; q < PYR_SLOTS; q++) { S.pyr_slot[q] = c.pyr[q]; S.mask_slot[q] = m.mask[q]; }
        static_assert(PYR_SLOTS == 3, "dv_apply_feed derives the last slot as (cur + 2) % 3");
        fb |= dv_side_bits(det_slot, c.cur, c.mask_nonzero[det_slot], c.mask_nonzero[c.cur]) << (2 + 8 * side);
        S.cam_model = c.cam_model_idx; S.cam_id = c.cam_id; S.thr = 2.0 / c.max_focal;
        S.pts_last = m.pts_last; S.ids_last = m.ids_last; S.n_last = m.n_last;
        S.pts = m.pts; S.ids = m.ids; S.close = m.close; S.first = m.first; S.grid = m.grid; S.cells = m.cells;
        S.top = m.top; S.top_n = m.top_n; S.ctl = m.ctl;
        S.h_pts_last = m.m_pts_last; S.h_ids_last = m.m_ids_last; S.h_n_last = m.m_n_last;
        max_cw = std::max(max_cw, geom.size_x);
        max_ch = std::max(max_ch, geom.size_y);
      }
      if (!u.stereo) U.s[1] = U.s[0];   // never used; keeps every pointer valid
      DvUnitMem &um = D.unit[(size_t)unit_idx];
      U.scr_a = um.scr[0]; U.scr_b = um.scr[1]; U.scr_c = um.scr[2]; U.scr_d = um.scr[3];
      U.part = um.part; U.tmp_pts = um.tmp_pts; U.tmp_ids = um.tmp_ids; U.uctl = um.uctl;
      U.obs = um.m_obs; U.obs_dev = um.d_obs; U.n_obs = um.m_n_obs; U.obs_cap = D.obs_cap;
      *um.h_n_obs = 0;
      job_slots += 2 * D.C;
      fast_words += 2 * (size_t)D.C * D.cell_cap;
      sp_slots += D.C * D.K;
      lk_slots += 2 * D.P;
      max_c = std::max(max_c, D.C);
      max_p = std::max(max_p, D.P);
      feed_bits.v[k] = fb;
      if (&rig != prev) {
        ctx->dv_heads.h[nheads++] = DvHead{k, 1};
        prev = &rig;
      } else {
        ctx->dv_heads.h[nheads - 1].count++;                 // the mono units of one rig share its id counter: same block, in order
      }
      k++;
    }
  }
  // sub-pixel slots: the left corners of every unit first (they are also the stereo-LK slots), then the right corners
  for (int k = 0; k < nu; k++) ctx->dv_units.h[k].spr_base = sp_slots + (-1 - ctx->dv_units.h[k].spr_base);
  const int spl_slots = sp_slots;
  sp_slots *= 2;
  // masks with a non-zero pixel are needed on the device (occupancy, cell selection, merge)
  for (int i = 0; i < n; i++) {
    FeedRequest &rq = *reqs[i];
    for (int k = 0; k < rq.n; k++) {
      CamState &c = *rq.rig->find_cam(rq.sensor_ids[k]);
      if (!c.mask_nonzero[c.cur]) continue;
      const int ci = (int)(&c - rq.rig->cams.data());
      CU_TRY(cudaMemcpyAsync(rq.rig->dv->cam[ci].mask[c.cur], c.mask[c.cur].data(), (size_t)c.w * c.h, cudaMemcpyHostToDevice, st));
    }
  }
  CU_TRY(ctx->dv_fast_jobs.reserve((size_t)job_slots));
  CU_TRY(ctx->dv_fast_cnt.reserve((size_t)job_slots, true));
  CU_TRY(ctx->dv_fast_out.reserve(fast_words));
  CU_TRY(ctx->dv_fast_active.reserve((size_t)job_slots + 1, true));
  CU_TRY(ctx->dv_sp_items.reserve((size_t)sp_slots));
  CU_TRY(ctx->dv_sp_out.reserve(2 * (size_t)sp_slots));
  CU_TRY(ctx->dv_lk1_items.reserve((size_t)spl_slots));
  CU_TRY(ctx->dv_lk1_out.reserve((size_t)spl_slots));
  CU_TRY(ctx->dv_lk1_status.reserve((size_t)spl_slots));
  CU_TRY(ctx->dv_lk2_items.reserve((size_t)lk_slots));
  CU_TRY(ctx->dv_lk2n_items.reserve((size_t)sp_slots));
  CU_TRY(ctx->dv_lk2n_out.reserve((size_t)sp_slots));
  CU_TRY(ctx->dv_lk1_idx.reserve((size_t)spl_slots));
  CU_TRY(ctx->dv_n_items.reserve(8, true));
  CU_TRY(ctx->dv_lk2_out.reserve((size_t)lk_slots));
  CU_TRY(ctx->dv_lk2_status.reserve((size_t)lk_slots));
  CU_TRY(ctx->dv_rs_mask.reserve((size_t)lk_slots));
  CU_TRY(ctx->dv_rs_mask_raw.reserve((size_t)lk_slots));
  CU_TRY(ctx->dv_rs_jobs.reserve(2 * (size_t)std::max(nu, 1)));
  CU_TRY(ctx->dv_p0n.reserve(2 * (size_t)lk_slots));
  CU_TRY(ctx->dv_p1n.reserve(2 * (size_t)lk_slots));
  DvGlobals X;
  X.fast_jobs = ctx->dv_fast_jobs.d; X.fast_out = ctx->dv_fast_out.d; X.fast_cnt = ctx->dv_fast_cnt.d;
  X.fast_active = ctx->dv_fast_active.d; X.fast_n_active = ctx->dv_fast_active.d + ctx->dv_fast_active.cap - 1;
  X.sp_items = ctx->dv_sp_items.d; X.sp_out = ctx->dv_sp_out.d;
  X.lk1_items = ctx->dv_lk1_items.d; X.lk1_out = ctx->dv_lk1_idx.d; X.lk1_res = ctx->dv_lk1_out.d; X.lk1_status = ctx->dv_lk1_status.d;
  X.n_items = ctx->dv_n_items.d;
  X.lk2_items = ctx->dv_lk2_items.d; X.lk2_out = ctx->dv_lk2_out.d;
  X.lk2n_items = ctx->dv_lk2n_items.d; X.lk2n_out = ctx->dv_lk2n_out.d;
  X.rs_jobs = ctx->dv_rs_jobs.d; X.p1n = ctx->dv_p1n.d; X.rs_mask = ctx->dv_rs_mask.d;
  // the descriptors live in device memory and go up only when they changed (a new batch shape, reallocated buffers): in
  // steady state a feed uploads nothing but its frames -- what varies per feed travels in the occupancy kernel's parameters
  if ((int)ctx->dv_units_sent.size() != nu || memcmp(ctx->dv_units_sent.data(), ctx->dv_units.h, sizeof(DvUnit) * (size_t)nu) != 0) {
    CU_TRY(ctx->dv_units.upload(st, (size_t)nu));
    ctx->dv_units_sent.assign(ctx->dv_units.h, ctx->dv_units.h + nu);
  }
  if ((int)ctx->dv_heads_sent.size() != nheads || memcmp(ctx->dv_heads_sent.data(), ctx->dv_heads.h, sizeof(DvHead) * (size_t)nheads) != 0) {
    CU_TRY(ctx->dv_heads.upload(st, (size_t)nheads));
    ctx->dv_heads_sent.assign(ctx->dv_heads.h, ctx->dv_heads.h + nheads);
  }
  CU_TRY(ctx->dv_feed_bits.reserve((size_t)DV_MAX_UNITS));
  X.feed_bits = ctx->dv_feed_bits.d;
  // ---- the new frames go up and are pre-processed on the upload stream (frames announced by submit_batch are on their way) ----
  npre = (int)ctx->pre_items.size();
  {
    const PreBuffers B{&ctx->pre_jobs, &ctx->down_jobs, &ctx->clahe_lut};
    e = issue_preprocess(dev, ctx->up_stream, B, ctx->pre_items, ctx->stats);
    if (e != OK) { ctx->cuda_err = dev->cuda_err; return e; }
    CU_TRY(cudaEventRecord(ctx->up_done, ctx->up_stream));
  }
  seg(0);
  if (nu > 0) {
    if (any_first) CU_TRY(wait_new_frames());   // first frames detect on the frame that is being uploaded (TrackKLT.cpp:110, :221)
    const DvUnit *U = ctx->dv_units.d;
    const DvHead *H = ctx->dv_heads.d;
    CU_TRY(launch_dv_occupancy(st, U, H, nheads, X, max_p, feed_bits));
    // The temporal LK of the points that survived the occupancy pass (the bulk of the feed's instructions) starts now, on the
    // lane's low-priority stream, beside the latency-bound detection chain; only the new points wait for the detection.
    CU_TRY(cudaEventRecord(ctx->occ_done, st));
    CU_TRY(cudaStreamWaitEvent(ctx->lk_stream, ctx->occ_done, 0));
    if (!any_first) {
      CU_TRY(cudaStreamWaitEvent(ctx->lk_stream, ctx->up_done, 0));   // it reads the new pyramids
      for (int i = 0; i < Context::PF_RING; i++)
        if ((wait_evts >> i) & 1u) CU_TRY(cudaStreamWaitEvent(ctx->lk_stream, dev->pf_evt[i], 0));
    }
    CU_TRY(launch_lk(ctx->lk_stream, ctx->dv_lk2_items.d, lk_slots, dev->pyr_dev.d, nullptr, 30, 0.01, ctx->dv_lk2_out.d, ctx->dv_lk2_status.d,
                     ctx->dv_lk2_out.d, ctx->dv_lk2_status.d));
    CU_TRY(cudaEventRecord(ctx->lk_done, ctx->lk_stream));
    // The rest of the chain is two runs of kernels with nothing external between them: the detection (fast .. assign) and the
    // tail (ransac, merge).  Once a batch shape has been seen twice they are replayed as CUDA graphs (one launch call each,
    // no per-kernel launch gaps); what varies from feed to feed -- unit descriptors, item lists, counters -- lives in memory
    // whose addresses do not change.  Shapes seen for the first time, first frames and profiled runs use plain launches.
    auto detect_run = [&]() -> cudaError_t {
      cudaError_t ce;
      if ((ce = launch_dv_fast(st, X, dev->pyr_dev.d, job_slots, max_cw, max_ch)) != cudaSuccess) return ce;
      if ((ce = launch_dv_topk(st, U, nu, max_c, X)) != cudaSuccess) return ce;
      if ((ce = launch_dv_select(st, U, H, nheads, X, max_p)) != cudaSuccess) return ce;
      if ((ce = launch_subpix(st, ctx->dv_sp_items.d, sp_slots, dev->pyr_dev.d, ctx->sp_mask.d, ctx->dv_sp_out.d, nullptr, ctx->dv_n_items.d)) != cudaSuccess) return ce;
      if ((ce = launch_lk(st, ctx->dv_lk1_items.d, spl_slots, dev->pyr_dev.d, ctx->dv_sp_out.d, 30, 0.01, ctx->dv_lk1_out.d, ctx->dv_lk1_status.d,
                          ctx->dv_lk1_out.d, ctx->dv_lk1_status.d, false, ctx->dv_lk1_idx.d, ctx->dv_n_items.d + 1)) != cudaSuccess) return ce;
      return launch_dv_assign(st, U, H, nheads, X, max_p);
    };
    auto tail_run = [&]() -> cudaError_t {
      cudaError_t ce = launch_ransac(st, ctx->dv_rs_jobs.d, 2 * nu, max_p, ctx->dv_lk2_items.d, ctx->dv_lk2_out.d, ctx->dv_lk2_status.d, dev->cam_dev.d,
                                     ctx->dv_p0n.d, ctx->dv_p1n.d, ctx->dv_p1n.d, ctx->dv_rs_mask.d, ctx->dv_rs_mask_raw.d);
      if (ce != cudaSuccess) return ce;
      return launch_dv_merge(st, U, H, nheads, X, max_p);
    };
    bool use_graphs = false;
#ifndef UVIO_NO_GRAPHS
    static const bool graphs_on = [] {
      const char *e = getenv("UVIO_KLT_GRAPHS");
      return !(e && e[0] == '0');
    }();
    if (graphs_on && !ctx->prof_on && !any_first) {
      unsigned long long sig = 1469598103934665603ull;
      auto mix = [&](unsigned long long v) { sig = (sig ^ v) * 1099511628211ull; };
      for (unsigned long long v : {(unsigned long long)nu, (unsigned long long)nheads, (unsigned long long)job_slots, (unsigned long long)sp_slots,
                                   (unsigned long long)lk_slots, (unsigned long long)max_c, (unsigned long long)max_p, (unsigned long long)max_cw,
                                   (unsigned long long)max_ch, (unsigned long long)(size_t)U, (unsigned long long)(size_t)H,
                                   (unsigned long long)(size_t)X.fast_jobs, (unsigned long long)(size_t)X.fast_out, (unsigned long long)(size_t)X.fast_cnt,
                                   (unsigned long long)(size_t)X.fast_active, (unsigned long long)(size_t)X.fast_n_active, (unsigned long long)(size_t)X.sp_items,
                                   (unsigned long long)(size_t)X.sp_out, (unsigned long long)(size_t)X.lk1_items, (unsigned long long)(size_t)X.lk1_out,
                                   (unsigned long long)(size_t)X.lk1_res, (unsigned long long)(size_t)X.lk1_status, (unsigned long long)(size_t)X.lk2_items,
                                   (unsigned long long)(size_t)X.lk2n_items, (unsigned long long)(size_t)X.lk2n_out, (unsigned long long)(size_t)X.lk2_out,
                                   (unsigned long long)(size_t)X.rs_jobs, (unsigned long long)(size_t)X.p1n, (unsigned long long)(size_t)X.rs_mask,
                                   (unsigned long long)(size_t)X.n_items, (unsigned long long)(size_t)X.feed_bits, (unsigned long long)(size_t)ctx->dv_lk2_status.d, (unsigned long long)(size_t)ctx->dv_p0n.d,
                                   (unsigned long long)(size_t)ctx->dv_rs_mask_raw.d, (unsigned long long)(size_t)dev->pyr_dev.d,
                                   (unsigned long long)(size_t)dev->cam_dev.d, (unsigned long long)(size_t)ctx->sp_mask.d})
        mix(v);
      if (sig == ctx->graph_sig && ctx->graph_detect && ctx->graph_tail) {
        use_graphs = true;
      } else if (sig == ctx->seen_sig) {       // second feed of this shape: capture (the plain launches of the first warmed everything up)
        auto capture = [&](const std::function<cudaError_t()> &run, cudaGraphExec_t &exec) -> bool {
          if (exec) { cudaGraphExecDestroy(exec); exec = nullptr; }
          cudaGraph_t graph = nullptr;
          if (cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal) != cudaSuccess) { cudaGetLastError(); return false; }
          const cudaError_t ce = run();
          const cudaError_t ee = cudaStreamEndCapture(st, &graph);
          bool ok = ce == cudaSuccess && ee == cudaSuccess && graph != nullptr;
          if (ok) ok = cudaGraphInstantiate(&exec, graph, 0) == cudaSuccess;
          if (graph) cudaGraphDestroy(graph);
          if (!ok) { cudaGetLastError(); exec = nullptr; }
          return ok;
        };
        if (capture(detect_run, ctx->graph_detect) && capture(tail_run, ctx->graph_tail)) {
          ctx->graph_sig = sig;
          use_graphs = true;
        } else {
          ctx->graph_sig = 0;
        }
      }
      ctx->seen_sig = sig;
    }
#endif
    if (use_graphs) {
#ifndef UVIO_NO_GRAPHS
      CU_TRY(cudaGraphLaunch(ctx->graph_detect, st));
#endif
    } else {
      CU_TRY(detect_run());
    }
    if (!any_first) CU_TRY(wait_new_frames());
    CU_TRY(launch_lk(st, ctx->dv_lk2n_items.d, sp_slots, dev->pyr_dev.d, nullptr, 30, 0.01, ctx->dv_lk2_out.d, ctx->dv_lk2_status.d,
                     ctx->dv_lk2_out.d, ctx->dv_lk2_status.d, false, ctx->dv_lk2n_out.d, ctx->dv_n_items.d + 2));
    CU_TRY(cudaStreamWaitEvent(st, ctx->lk_done, 0));
    if (use_graphs) {
#ifndef UVIO_NO_GRAPHS
      CU_TRY(cudaGraphLaunch(ctx->graph_tail, st));
#endif
    } else {
      CU_TRY(tail_run());
    }
    ctx->stats[1] += 11;
  } else {
    CU_TRY(wait_new_frames());
  }
  seg(2);
  return OK;
}

// second half of the device-side feed: wait for the chain, then point the trackers at what the GPU left in mapped memory
// (observation lists, pts_last / ids_last mirrors); nothing is copied until somebody asks for it
int Pass::device_collect() {
  UVIO_NVTX("device_collect: wait + results");
  if (ctx->sync() != OK) return E_CUDA;
  seg(5);
  ctx->prof_fold();
  t1 = t2 = t3 = now_us();
  for (auto &u : units) {
    Rig &rig = *u.rig;
    DvRig &D = *rig.dv;
    const int unit_idx = u.stereo ? 0 : u.L.msg;
    DvUnitMem &um = D.unit[(size_t)unit_idx];
    const int no = std::min(*um.h_n_obs, D.obs_cap);
    static_assert(sizeof(Obs) == sizeof(DvObs), "the observation structs are the same bytes");
    if (no > 0) rig.dv_obs.emplace_back((const Obs *)um.h_obs, no);
    ctx->stats[3] += sizeof(Obs) * (size_t)no;
    for (int side = 0; side < (u.stereo ? 2 : 1); side++) {
      CamState &c = *(side ? u.R : u.L).cam;
      DvCamMem &m = D.cam[(size_t)(&c - rig.cams.data())];
      c.dv_n_last = std::min(*m.h_n_last, D.P);
      c.have_last = true;
      ctx->stats[3] += 16ull * (size_t)c.dv_n_last;
    }
    rig.dv_mirror_stale = true;
    rig.dv_mirror_published = rig.dv_wants_mirror;
    rig.currid = (size_t)*D.h_currid;   // the id counter's mirror (written by k_dv_assign)
  }
  ctx->stats[0] += (unsigned long long)n;
  const double t4 = now_us();
  seg(6);
  ctx->seg_us[7] += 1;
  ctx->prof_end();
  for (int i = 0; i < n; i++) {
    float *sus = reqs[i]->rig->stage_us;
    sus[0] = (float)(t1 - t0);
    sus[1] = sus[2] = sus[3] = 0.f;
    sus[4] = (float)(t4 - t3);
  }
  return OK;
}

void dv_sync_mirror(Rig &rig) {
  rig.dv_wants_mirror = true;   // whoever reads the last points once will do so again: later feeds publish them
  if (!rig.dv_mirror_stale || !rig.dv) return;
  DvRig &D = *rig.dv;
  for (size_t k = 0; k < rig.cams.size(); k++) {
    CamState &c = rig.cams[k];
    if (c.dv_n_last < 0) continue;
    const DvCamMem &m = D.cam[k];
    const int nl = c.dv_n_last;
    c.pts_last.resize((size_t)nl);
    c.ids_last.resize((size_t)nl);
    if (!nl) continue;
    if (rig.dv_mirror_published) {
      memcpy(c.pts_last.data(), m.h_pts_last, sizeof(Pt) * (size_t)nl);
      for (int i = 0; i < nl; i++) c.ids_last[(size_t)i] = (size_t)m.h_ids_last[i];
    } else {                      // the feed did not publish: fetch from the device state (the feed that wrote it has been waited for)
      std::vector<unsigned long long> ids((size_t)nl);
      if (cudaSetDevice(rig.ctx->device) != cudaSuccess ||
          cudaMemcpy(c.pts_last.data(), m.pts_last, sizeof(Pt) * (size_t)nl, cudaMemcpyDeviceToHost) != cudaSuccess ||
          cudaMemcpy(ids.data(), m.ids_last, 8 * (size_t)nl, cudaMemcpyDeviceToHost) != cudaSuccess) {
        rig.last_error = E_CUDA;
        c.pts_last.clear();
        c.ids_last.clear();
        continue;
      }
      for (int i = 0; i < nl; i++) c.ids_last[(size_t)i] = (size_t)ids[(size_t)i];
    }
  }
  rig.dv_mirror_stale = false;
}


}  // namespace uvio
