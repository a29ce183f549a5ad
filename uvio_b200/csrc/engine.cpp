// Per-device context of the GPU TrackKLT, tracker objects, and the batched entry points (feed_batch, submit_batch).
// See engine.h for the public side and engine_internal.h for how the engine is cut into translation units:
//   engine_device.cpp  the default pipeline: a feed is one chain of launches, bookkeeping on the device (track_body.h)
//   engine_host.cpp    round 1's pipeline (UVIO_KLT_PIPELINE=host): three GPU rounds with host bookkeeping between them
//   engine_upload.cpp  frame upload and the image-wide passes of a batch
//   engine_stages.cpp  per-stage entry points for the parity tests
#include "engine_internal.h"

namespace uvio {

thread_local KernelProf *g_prof = nullptr;

void Context::background_wait() {
  if (pool) pool->finish();
}
void Context::uploader_drain() {
  std::unique_lock<std::mutex> lk(pf_mtx);
  pf_cv.wait(lk, [&] { return pf_issued == pf_queued; });
}
int Context::wait_issued(unsigned long long seq) {
  std::unique_lock<std::mutex> lk(pf_mtx);
  pf_cv.wait(lk, [&] { return pf_issued >= seq; });
  return pf_error;
}

Context::~Context() {
  g_contexts_gone.store(true);
  if (uploader.joinable()) {
    {
      std::lock_guard<std::mutex> lk(pf_mtx);
      pf_quit = true;
    }
    pf_cv.notify_all();
    uploader.join();
  }
  delete pool;
  lanes.clear();
  if (pf_stream) cudaStreamDestroy(pf_stream);
  if (pf_copy_stream) cudaStreamDestroy(pf_copy_stream);
  for (int i = 0; i < PF_RING; i++) {
    if (pf_evt[i]) cudaEventDestroy(pf_evt[i]);
    if (pf_copied[i]) cudaEventDestroy(pf_copied[i]);
  }
  for (void *p : allocs) cudaFree(p);
}

// Host threads per device context: UVIO_KLT_LANES, else half of the process's share of the logical cores (one process per
// GPU shares the host with its siblings), at most 8.  Every thread drives UVIO_KLT_DEPTH lanes alternately: 1 when there
// are 8 threads, else 2.  Measured on a 16-core host with one B200 (32 / 64 streams): 8 x 1 beats 4 x 2 by 3 % / 4 %
// resident and 0 % / 16 % end to end; 12 or 16 threads lose to the shrinking launches and the busy cores.
static int default_threads() {
  if (const char *e = getenv("UVIO_KLT_LANES")) {
    int v = atoi(e);
    if (v >= 1) return v > 16 ? 16 : v;
  }
  int hw = (int)std::thread::hardware_concurrency();
  int local_world = 1;
  if (const char *e = getenv("LOCAL_WORLD_SIZE")) local_world = atoi(e) > 0 ? atoi(e) : 1;
  int v = hw / local_world / 2;
  return v < 1 ? 1 : (v > 8 ? 8 : v);
}
static int default_depth(int threads) {
  if (const char *e = getenv("UVIO_KLT_DEPTH")) {
    int v = atoi(e);
    if (v >= 1) return v > 4 ? 4 : v;
  }
  return threads >= 8 ? 1 : 2;
}

// The context uses 2 streams per lane and one for announced frames.  With the driver's default of 8 hardware work queues
// several streams share a queue and falsely serialise: a lane's launch queued behind the 1 MiB pieces of a frame upload
// waits for all of them (seen as one lane per call finishing 0.5-1 ms late).  The variable only counts if it is set
// before the process creates its CUDA context, so a host that initialises CUDA first should export it itself.
static const int g_connections_hint = [] {
  setenv("CUDA_DEVICE_MAX_CONNECTIONS", "32", 0);
  return 0;
}();

Context *context_for_device(int device, int *err) {
  static std::mutex m;
  static std::map<int, std::unique_ptr<Context>> table;
  std::lock_guard<std::mutex> lk(m);
  auto it = table.find(device);
  if (it != table.end()) return it->second.get();
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || device < 0 || device >= count) {   // no CUDA device: there is no other way to run this tracker
    if (err) *err = e != cudaSuccess ? (int)e : (int)cudaErrorInvalidDevice;
    return nullptr;
  }
  std::unique_ptr<Context> c(new Context);
  c->device = device;
  if ((e = cudaSetDevice(device)) != cudaSuccess) {
    if (err) *err = (int)e;
    return nullptr;
  }
  c->threads = default_threads();
  int prio_lo = 0, prio_hi = 0;
  cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi);
  const int nl = c->threads * default_depth(c->threads);
  for (int i = 0; i < nl; i++) {
    std::unique_ptr<Lane> ln(new Lane);
    ln->dev = c.get();
    ln->index = i;
    // the rounds a feed waits for run ahead of the pre-processing of frames that are only needed later
    if ((e = cudaStreamCreateWithPriority(&ln->stream, cudaStreamNonBlocking, prio_hi)) != cudaSuccess ||
        (e = cudaStreamCreateWithPriority(&ln->up_stream, cudaStreamNonBlocking, prio_lo)) != cudaSuccess ||
        (e = cudaEventCreateWithFlags(&ln->up_done, cudaEventDisableTiming)) != cudaSuccess ||
        (e = cudaStreamCreateWithPriority(&ln->lk_stream, cudaStreamNonBlocking, prio_lo)) != cudaSuccess ||
        (e = cudaEventCreateWithFlags(&ln->lk_go, cudaEventDisableTiming)) != cudaSuccess ||
        (e = cudaEventCreateWithFlags(&ln->lk_done, cudaEventDisableTiming)) != cudaSuccess ||
        (e = cudaEventCreateWithFlags(&ln->occ_done, cudaEventDisableTiming)) != cudaSuccess) {
      if (err) *err = (int)e;
      return nullptr;
    }
    c->lanes.push_back(std::move(ln));
  }
  // the kernels of announced batches (equalisation, pyramids): UVIO_KLT_UP_PRIO=high lets them overtake the LK blocks of the
  // feed in flight, so that an announced frame is ready soon after its upload ends
  const char *up_prio = getenv("UVIO_KLT_UP_PRIO");
  const int pf_prio = (up_prio && !strcmp(up_prio, "low")) ? prio_lo : prio_hi;
  if ((e = cudaStreamCreateWithPriority(&c->pf_stream, cudaStreamNonBlocking, pf_prio)) != cudaSuccess ||
      (e = cudaStreamCreateWithPriority(&c->pf_copy_stream, cudaStreamNonBlocking, getenv("UVIO_KLT_COPY_PRIO_LOW") ? prio_lo : prio_hi)) != cudaSuccess) {
    if (err) *err = (int)e;
    return nullptr;
  }
  for (int i = 0; i < Context::PF_RING; i++)
    if ((e = cudaEventCreateWithFlags(&c->pf_evt[i], cudaEventDisableTiming)) != cudaSuccess ||
        (e = cudaEventCreateWithFlags(&c->pf_copied[i], cudaEventDisableTiming)) != cudaSuccess) {
      if (err) *err = (int)e;
      return nullptr;
    }
  c->pool = new Pool(c->threads);   // one worker per lane thread: the caller takes a share in run(), none in start()
  for (auto &a : c->pf_stats) a.store(0);
  c->uploader = std::thread([raw = c.get()] { raw->uploader_loop(); });
  Context *p = c.get();
  table[device] = std::move(c);
  return p;
}

Rig::Rig(Context *c, const Params &p, const CamCalib *cc, int n_cams) : ctx(c), prm(p) {
  static std::atomic<unsigned long long> next_serial{1};
  serial = next_serial.fetch_add(1);
  currid = 4 * (size_t)p.num_aruco + 1;   // TrackBase.cpp:34
  std::lock_guard<std::mutex> lk(c->mtx);
  for (int i = 0; i < n_cams; i++) {
    CamState s;
    s.cam_id = cc[i].cam_id;
    s.cam_model_idx = (int)c->cam_tab.size();
    s.max_focal = cc[i].model.k[0] > cc[i].model.k[1] ? cc[i].model.k[0] : cc[i].model.k[1];
    c->cam_tab.push_back(cc[i].model);
    cam_index[s.cam_id] = (int)cams.size();
    cams.push_back(std::move(s));
  }
  c->tables_dirty = true;
}
CamState *Rig::find_cam(size_t id) {
  auto it = cam_index.find(id);
  return it == cam_index.end() ? nullptr : &cams[it->second];
}

// Device memory of a destroyed tracker's camera: the pyramids go back to the driver, the upload buffers to the free list.
// (Its entries in the image tables stay; they are a few hundred bytes and are never referenced again.)
void Context::release_camera(CamState &c) {
  if (!c.dev_base) return;
  for (size_t k = 0; k < allocs.size(); k++)
    if (allocs[k] == c.dev_base) {
      allocs.erase(allocs.begin() + (long)k);
      break;
    }
  cudaFree(c.dev_base);
  c.dev_base = nullptr;
  raw_free.push_back(RawSlot{c.raw_bytes, {c.raw_ptr[0], c.raw_ptr[1]}});
}


Rig::~Rig() {
  if (!g_contexts_gone.load()) {
    std::lock_guard<std::mutex> lk(ctx->mtx);
    ctx->uploader_drain();
    ctx->background_wait();
    ctx->forget_rig(this);     // before the units go: the lanes' detections issued ahead point at them
    bool any = false;
    for (auto &c : cams) any |= c.dev_base != nullptr;
    if (any && cudaSetDevice(ctx->device) == cudaSuccess) {
      cudaDeviceSynchronize();   // a detection issued ahead or an announced frame of this tracker may still be running
      for (auto &c : cams) ctx->release_camera(c);
    }
  }
  for (Unit *u : unit_pool) delete u;
  if (dv && !g_contexts_gone.load()) dv_destroy(dv);
}
Unit *Rig::unit_at(size_t i) {
  while (unit_pool.size() <= i) unit_pool.push_back(new Unit);
  Unit *u = unit_pool[i];
  u->rig = this;
  u->failed = false;
  return u;
}

// The checks of TrackKLT::feed_new_camera on one message; new cameras get their device memory here.
int validate_request(Context *ctx, FeedRequest &rq, bool announcing) {
  Rig &rig = *rq.rig;
  const Params &P = rig.prm;
  if (rq.n <= 0 || !rq.sensor_ids || !rq.images || (!announcing && !rq.masks)) return E_BADMSG;   // TrackKLT.cpp:37-43
  if (P.use_stereo && rq.n > 2) return E_BADMSG;                                                  // :90-93
  if (P.win_size != LK_WIN) return E_ARG;
  for (int k = 0; k < rq.n; k++) {
    CamState *c = rig.find_cam(rq.sensor_ids[k]);
    if (!c) return E_CAMID;                                                                       // :53-54 (.at throws)
    const HostImage &im = rq.images[k];
    if (!im.p || im.w <= 0 || im.h <= 0 || im.w > 4095 || im.h > 4095) return E_BADMSG;
    if (!announcing && rq.masks[k].p && (rq.masks[k].w != im.w || rq.masks[k].h != im.h)) return E_BADMSG;
    for (int q = 0; q < k; q++)
      if (rq.sensor_ids[q] == rq.sensor_ids[k]) return E_BADMSG;
    if (c->raw_in[0] < 0) {
      ctx->uploader_drain();   // a new camera grows the tables a batch in flight may be reading
      ctx->background_wait();  // ... and so may the background part of the last feed
      int e = ctx->alloc_camera(*c, im.w, im.h, P.pyr_levels, P.win_size, P.downsample_cameras != 0);
      if (e != OK) return e;
    } else if (c->in_w != im.w || c->in_h != im.h) return E_SIZE;
    const unsigned ahead = c->n_sub - c->n_fed;   // frames announced and not yet fed
    if (announcing) {
      if (ahead >= 2) return E_ARG;               // the third slot is the only one a feed in flight does not read
    } else if (ahead > 0 && c->sub_src[(c->n_fed + 1) % PYR_SLOTS] != im.p) {
      return E_ARG;                               // the feed after a submit must be given the announced frame
    }
  }
  return OK;
}

// Issues the batches submit_batch queued: H2D pieces, the optional halving, equalisation + pyramids, then the batch's event.
void Context::uploader_loop() {
  Context *ctx = this;
  cudaSetDevice(device);
  while (true) {
    int q;
    {
      std::unique_lock<std::mutex> lk(pf_mtx);
      pf_cv.wait(lk, [&] { return pf_quit || !pf_queue.empty(); });
      if (pf_queue.empty()) return;   // quit
      q = pf_queue.front();
    }
    unsigned long long st[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    const PreBuffers B{&pf_pre[q], &pf_down[q], &pf_clahe[q]};
    if (pf_prof_on.load()) {
      if (!pf_prof.ready) {
        for (int i = 0; i < KernelProf::MAX_MARKS; i++) cudaEventCreate(&pf_prof.ev[i]);
        pf_prof.ready = true;
      }
      if (pf_prof.n > 0) {   // the batch before this one: wait for it and take its intervals
        cudaStreamSynchronize(pf_stream);
        std::lock_guard<std::mutex> lk(pf_mtx);
        fold_prof(pf_prof, pf_prof_ms, pf_prof_launches);
      }
      pf_prof.open = false;
      pf_prof.mask = pf_prof_mask;
      g_prof = &pf_prof;
    } else {
      g_prof = nullptr;
    }
    // The copies of this batch overwrite the upload buffers of the same frame parity: whatever read them two announcements
    // ago (or earlier: the kernel stream is in order) must be done.  The announcement before this one reads the other pair.
    const int q2 = (q + PF_RING - 2) % PF_RING;
    int e = OK;
    if (pf_recorded[q2] && cudaStreamWaitEvent(pf_copy_stream, pf_evt[q2], 0) != cudaSuccess) e = E_CUDA;
    if (e == OK) e = issue_preprocess(ctx, pf_stream, B, pf_items[q], st, pf_group, pf_copy_stream, pf_copied[q]);
    if (e == OK && cudaEventRecord(pf_evt[q], pf_stream) != cudaSuccess) e = E_CUDA;
    pf_recorded[q] = true;
    pf_stats[1] += st[1];
    pf_stats[2] += st[2];
    {
      std::lock_guard<std::mutex> lk(pf_mtx);
      pf_queue.erase(pf_queue.begin());
      pf_issued++;
      if (e != OK) pf_error = e;
    }
    pf_cv.notify_all();
  }
}

int submit_batch(Context *ctx, FeedRequest *reqs, int n) {
  UVIO_NVTX("uvio_klt_submit_batch");
  HOSTPROF(submit_total);
  std::lock_guard<std::mutex> guard(ctx->mtx);
  if (cudaSetDevice(ctx->device) != cudaSuccess) return E_CUDA;
  int worst = OK;
  const int q = (int)(ctx->pf_seq++ % Context::PF_RING);
  // the ring entry is about to be rewritten: its previous batch (8 calls old) must have been issued and have run
  if (ctx->wait_issued(ctx->pf_entry_seq[q]) != OK) return E_CUDA;
  if (ctx->pf_used[q]) CU_TRY(cudaEventSynchronize(ctx->pf_evt[q]));
  std::vector<PreItem> &items = ctx->pf_items[q];
  items.clear();
  for (int i = 0; i < n; i++) {
    FeedRequest &rq = reqs[i];
    rq.status = validate_request(ctx, rq, true);
    if (rq.status != OK) {
      rq.rig->last_error = rq.status;
      if (rq.status > worst) worst = rq.status;
      continue;
    }
    for (int k = 0; k < rq.n; k++) {
      CamState &c = *rq.rig->find_cam(rq.sensor_ids[k]);
      c.n_sub++;
      const int slot = (int)(c.n_sub % PYR_SLOTS);
      c.sub_src[slot] = rq.images[k].p;
      c.sub_evt[slot] = q;
      c.sub_seq[slot] = ctx->pf_queued_next();
      items.push_back(PreItem{&c, rq.images[k], slot, rq.rig->prm.histogram_method, rq.rig->prm.downsample_cameras != 0, (int)(c.n_sub & 1u)});
    }
  }
  if (ctx->upload_tables() != OK) return E_CUDA;
  static const int group = [] {
    // images per upload + pre-processing group of an announced batch; 0 = one group.  Measured (32 streams): groups of 8
    // or 16 images cost more (a cluster launch per group) than their earlier pyramids give back
    const char *e = getenv("UVIO_KLT_SUBMIT_GROUP");
    int v = e ? atoi(e) : 0;
    return v < 0 ? 0 : v;
  }();
  {
    std::lock_guard<std::mutex> lk(ctx->pf_mtx);
    ctx->pf_group = group;
    ctx->pf_queue.push_back(q);
    ctx->pf_entry_seq[q] = ++ctx->pf_queued;
  }
  ctx->pf_cv.notify_all();
  ctx->pf_used[q] = true;
  return worst;
}

void Context::forget_rig(Rig *rig) {
  for (auto &ln : lanes)
    for (Unit *u : ln->ahead.unit)
      if (ln->ahead.valid && u->rig == rig) {
        cancel_ahead(*ln);
        break;
      }
}

int feed_batch(Context *ctx, FeedRequest *reqs, int n) {
  UVIO_NVTX("uvio_klt_feed_batch");
  HOSTPROF(fb_total);
  std::lock_guard<std::mutex> guard(ctx->mtx);
  if (cudaSetDevice(ctx->device) != cudaSuccess) return E_CUDA;
  ctx->background_wait();   // the background part of the previous call (it may hold these trackers' detection)
  int worst = OK;
  auto fail = [&](FeedRequest &r, int code) {
    r.status = code;
    r.rig->last_error = code;
    if (code > worst) worst = code;
  };
  // ---- validate every message and give new cameras their device memory (serial: it edits the shared tables) ----------------
  std::vector<FeedRequest *> good;
  for (int i = 0; i < n; i++) {
    FeedRequest &rq = reqs[i];
    rq.status = OK;
    rq.rig->obs.clear();
    rq.rig->dv_obs.clear();
    int e = validate_request(ctx, rq, false);
    if (e != OK) fail(rq, e);
    else good.push_back(&rq);
  }
  if (ctx->upload_tables() != OK) return E_CUDA;
  if (device_pipeline_enabled()) {
    // The device-side pipeline: the trackers of the call are cut into a few groups, each a chain of launches on its own
    // stream (so that the latency-bound kernels of one group run under the LK of another), all issued by this thread;
    // then one wait per group.  No host thread pool, no host bookkeeping between kernels.
    static const int max_groups = [] {
      const char *e = getenv("UVIO_KLT_GROUPS");
      int v = e ? atoi(e) : 2;
      return v < 1 ? 1 : (v > 16 ? 16 : v);
    }();
    for (auto &ln : ctx->lanes) cancel_ahead(*ln);
    const int need_groups = ((int)good.size() * 2 + DV_MAX_UNITS - 1) / DV_MAX_UNITS;   // a launch takes DV_MAX_UNITS units, a tracker has at most 2... per message image
    const int ng = std::min((int)ctx->lanes.size(), std::max(1, std::max(need_groups, std::min(max_groups, ((int)good.size() + 3) / 4))));
    if (!ctx->bg_passes) ctx->bg_passes = new std::vector<Pass>;
    std::vector<Pass> &grp = *ctx->bg_passes;
    grp.clear();
    grp.resize((size_t)ng);
    int rc = OK;
    for (int g = 0; g < ng && !good.empty(); g++) {
      size_t b = good.size() * (size_t)g / ng, e = good.size() * (size_t)(g + 1) / ng;
      Lane &ln = *ctx->lanes[(size_t)g];
      grp[g].ctx = &ln;
      grp[g].reqs = good.data() + b;
      grp[g].n = (int)(e - b);
      g_prof = ln.prof_on ? &ln.prof : nullptr;
      int r = grp[g].device_feed();
      g_prof = nullptr;
      if (r != OK) { grp[g].rc = r; rc = std::max(rc, r); }
    }
    for (int g = 0; g < ng && !good.empty(); g++) {
      Lane &ln = *ctx->lanes[(size_t)g];
      g_prof = ln.prof_on ? &ln.prof : nullptr;
      int r = grp[g].rc == OK ? grp[g].device_collect() : (cudaStreamSynchronize(ln.stream), grp[g].rc);
      g_prof = nullptr;
      rc = std::max(rc, r);
      if (grp[g].worst > worst) worst = grp[g].worst;
      for (int k = 0; k < 8; k++) {
        ctx->stats[k] += ln.stats[k];
        ln.stats[k] = 0;
        ctx->seg_us[k] += ln.seg_us[k];
        ln.seg_us[k] = 0;
      }
      if (ln.cuda_err) ctx->cuda_err = ln.cuda_err;
      grp[g].reqs = nullptr;
    }
    if (rc > worst) worst = rc;
    ctx->stats[4] += 1;
    return worst;
  }
  // ---- split the trackers over the lanes (contiguous shares); every host thread drives `depth` lanes stage by stage ---------
  const int nl = std::max(1, std::min((int)ctx->lanes.size(), (int)good.size()));
  const int nthreads = std::max(1, std::min(ctx->threads, nl));
  if (!ctx->bg_passes) ctx->bg_passes = new std::vector<Pass>;
  std::vector<Pass> &passes = *ctx->bg_passes;   // they outlive the call: Pass::background runs after it
  passes.clear();
  passes.resize((size_t)nl);
  for (int l = 0; l < nl; l++) {
    size_t b = good.size() * (size_t)l / nl, e = good.size() * (size_t)(l + 1) / nl;
    passes[l].ctx = ctx->lanes[l].get();
    passes[l].reqs = good.data() + b;
    passes[l].n = (int)(e - b);
    // a tracker whose finished detection sits in another lane (the batch was composed differently last time): drop that
    // lane's result before anything of this call runs, so that the id counters are back where the reference has them
    for (size_t i = b; i < e; i++) {
      const int bl = good[i]->rig->bg_lane;
      if (bl >= 0 && bl != l) cancel_ahead(*ctx->lanes[(size_t)bl]);
    }
  }
  static const bool trace = getenv("UVIO_KLT_TRACE_LANES") != nullptr;   // development: per-lane wall clock of a few calls
  const double t_call = now_us();
  std::vector<double> t_start(nl, 0), t_end(nl, 0);
  {
  HOSTPROF(fb_pool_run);
  ctx->pool->run(nthreads, [&](int t) {
    HOSTPROF(fb_thread);
    if (trace) for (int l = t; l < nl; l += nthreads) t_start[l] = now_us();
    for (int stage = 0; stage < 4; stage++)
      for (int l = t; l < nl; l += nthreads) {
        passes[l].step();
        if (trace && stage == 3) t_end[l] = now_us();
      }
  });
  }
  if (trace && (ctx->stats[4] % 8) == 5) {
    fprintf(stderr, "[lanes] call %llu: pool done at +%.0f us; lane start/end:", ctx->stats[4], now_us() - t_call);
    for (int l = 0; l < nl; l++)
      fprintf(stderr, " [%d: s1@%.0f s2@%.0f s3@%.0f end@%.0f]", l, passes[l].t1 - t_call, passes[l].t2 - t_call, passes[l].t3 - t_call,
              t_end[l] - t_call);
    fprintf(stderr, "\n");
  }
  std::vector<int> rc(nl, OK);
  for (int l = 0; l < nl; l++) rc[l] = passes[l].result();
  for (int l = 0; l < nl; l++) {
    Lane &ln = *ctx->lanes[l];
    for (int k = 0; k < 8; k++) {
      ctx->stats[k] += ln.stats[k];
      ln.stats[k] = 0;
      ctx->seg_us[k] += ln.seg_us[k];
      ln.seg_us[k] = 0;
    }
    if (ln.cuda_err) ctx->cuda_err = ln.cuda_err;
    if (rc[l] > worst) worst = rc[l];
  }
  ctx->stats[4] += 1;
  // The call returns here; the pool's workers go on with rounds 1-2 of the next feed of the same trackers (Pass::background).
  if (worst != E_CUDA && background_enabled()) {
    for (auto &p : passes) p.reqs = nullptr;   // the requests die with the call; the background part never looks at them
    ctx->pool->start(nl, [ctx](int l) { (*ctx->bg_passes)[(size_t)l].background(); });
  }
  return worst;
}

int context_cuda_error(Context *ctx) { return ctx->cuda_err; }

void context_quiesce(Context *ctx) {
  std::lock_guard<std::mutex> guard(ctx->mtx);
  ctx->background_wait();
}

int context_profile(Context *ctx, int enable, double ms[K_COUNT], unsigned long long launches[K_COUNT]) {
  std::lock_guard<std::mutex> guard(ctx->mtx);
  ctx->background_wait();
  if (cudaSetDevice(ctx->device) != cudaSuccess) return E_CUDA;
  for (int k = 0; k < K_COUNT; k++) {
    if (ms) ms[k] = 0;
    if (launches) launches[k] = 0;
  }
  ctx->uploader_drain();
  {
    cudaStreamSynchronize(ctx->pf_stream);
    std::lock_guard<std::mutex> lk(ctx->pf_mtx);
    if (ctx->pf_prof.n > 0) fold_prof(ctx->pf_prof, ctx->pf_prof_ms, ctx->pf_prof_launches);
    for (int k = 0; k < K_COUNT; k++) {
      if (ms) ms[k] += ctx->pf_prof_ms[k];
      if (launches) launches[k] += ctx->pf_prof_launches[k];
    }
    if (enable >= 0) {
      ctx->pf_prof_on.store(enable != 0);
      ctx->pf_prof_mask = enable > 1 ? (unsigned)(enable >> 1) : ~0u;
      if (enable) {
        memset(ctx->pf_prof_ms, 0, sizeof(ctx->pf_prof_ms));
        memset(ctx->pf_prof_launches, 0, sizeof(ctx->pf_prof_launches));
      }
    }
  }
  for (auto &ln : ctx->lanes) {
    for (int k = 0; k < K_COUNT; k++) {
      if (ms) ms[k] += ln->prof_ms[k];
      if (launches) launches[k] += ln->prof_launches[k];
    }
    if (enable >= 0) {
      ln->prof_on = enable != 0;
      ln->prof.n = 0;
      ln->prof.open = false;
      ln->prof.mask = enable > 1 ? (unsigned)(enable >> 1) : ~0u;   // enable = 1: all kernels; 1 | mask<<1: only those
      if (enable) {
        memset(ln->prof_ms, 0, sizeof(ln->prof_ms));
        memset(ln->prof_launches, 0, sizeof(ln->prof_launches));
      }
    }
  }
  return OK;
}

void context_stats(Context *ctx, unsigned long long out[12]) {
  std::lock_guard<std::mutex> guard(ctx->mtx);
  ctx->background_wait();
  memcpy(out, ctx->stats, sizeof(ctx->stats));
  out[1] += ctx->pf_stats[1].load();
  out[2] += ctx->pf_stats[2].load();
  out[8] = out[9] = out[10] = out[11] = 0;
  unsigned long long w[4] = {0, 0, 0, 0};
  if (cudaSetDevice(ctx->device) == cudaSuccess && read_lk_work(w, 0) == cudaSuccess) {
    out[6] = w[0];   // LK point-levels
    out[7] = w[1];   // LK iterations
    out[8] = w[3];   // ... of which ran the ordered fp32 chains
  }
}

// CUDA-event stopwatch. The events are recorded on lane 0's stream; every feed is host-synchronous on all lanes, so the
// interval between the two records spans exactly the device work (and host gaps) of the feeds issued in between.
void context_host_profile(Context *ctx, int reset, double us[8]) {
  std::lock_guard<std::mutex> guard(ctx->mtx);
  ctx->background_wait();
  for (int k = 0; k < 8; k++) {
    if (us) us[k] = ctx->seg_us[k];
    if (reset) ctx->seg_us[k] = 0;
  }
}

int context_timer(Context *ctx, int stop, float *ms) {
  std::lock_guard<std::mutex> guard(ctx->mtx);
  ctx->background_wait();
  if (cudaSetDevice(ctx->device) != cudaSuccess) return E_CUDA;
  for (int i = 0; i < 2; i++)
    if (!ctx->ev[i]) CU_TRY(cudaEventCreate(&ctx->ev[i]));
  CU_TRY(cudaEventRecord(ctx->ev[stop ? 1 : 0], ctx->lanes[0]->stream));
  if (stop) {
    CU_TRY(cudaEventSynchronize(ctx->ev[1]));
    CU_TRY(cudaEventElapsedTime(ms, ctx->ev[0], ctx->ev[1]));
  }
  return OK;
}


}  // namespace uvio
