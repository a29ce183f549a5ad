// Internals shared by the translation units of the GPU TrackKLT engine (engine.cpp: context, batching and the public entry
// points; engine_upload.cpp: frame upload + image passes; engine_host.cpp: the round-1 pipeline with the bookkeeping on host
// threads; engine_device.cpp: the device-side pipeline; engine_stages.cpp: the per-stage test entry points).  Not installed.
#pragma once
#include "engine.h"
#include "nvtx_ranges.h"
#include <stdio.h>
#include <stdlib.h>
#include <chrono>
#include <condition_variable>
#include <functional>
#include <map>
#include <memory>
#include <atomic>
#include <thread>
#include "hostlogic.h"
#include "lk_body.h"
#include "subpix_body.h"
#include "track_body.h"

namespace uvio {

extern thread_local KernelProf *g_prof;

// Optional host-side scope timers (compile with -DUVIO_HOSTPROF; development aid, prints at exit).
#ifdef UVIO_HOSTPROF
struct HostProfTable {
  const char *name[32];
  std::atomic<long long> ns[32];
  std::atomic<long long> calls[32];
  int n = 0;
  std::mutex m;
  int slot(const char *nm) {
    std::lock_guard<std::mutex> lk(m);
    for (int i = 0; i < n; i++) if (name[i] == nm) return i;
    name[n] = nm;
    return n++;
  }
  ~HostProfTable() {
    for (int i = 0; i < n; i++)
      fprintf(stderr, "[hostprof] %-24s %10.1f us total %8lld calls %8.2f us/call\n", name[i], ns[i].load() / 1e3, calls[i].load(),
              ns[i].load() / 1e3 / (double)std::max(1ll, calls[i].load()));
  }
};
inline HostProfTable g_hostprof;
struct HostProfScope {
  int slot;
  std::chrono::steady_clock::time_point t0;
  explicit HostProfScope(int s) : slot(s), t0(std::chrono::steady_clock::now()) {}
  ~HostProfScope() {
    g_hostprof.ns[slot] += std::chrono::duration_cast<std::chrono::nanoseconds>(std::chrono::steady_clock::now() - t0).count();
    g_hostprof.calls[slot] += 1;
  }
};
#define HOSTPROF(id) static const int hp_slot_##id = g_hostprof.slot(#id); HostProfScope hp_scope_##id(hp_slot_##id)
#else
#define HOSTPROF(id) do { } while (0)
#endif

#define CU_TRY(expr)                                          \
  do {                                                        \
    cudaError_t e__ = (expr);                                 \
    if (e__ != cudaSuccess) { ctx->cuda_err = (int)e__; return E_CUDA; } \
  } while (0)

template <class T>
struct MappedVec {   // pinned host memory mapped into the device address space (job lists in, small results out)
  T *h = nullptr, *d = nullptr;
  size_t cap = 0;
  cudaError_t reserve(size_t want) {   // contents are not preserved; call only while the stream is idle
    if (want <= cap) return cudaSuccess;
    size_t ncap = want + want / 2 + 64;
    if (h) cudaFreeHost(h);
    h = d = nullptr;
    cap = 0;
    cudaError_t e = cudaHostAlloc((void **)&h, ncap * sizeof(T), cudaHostAllocMapped);
    if (e != cudaSuccess) return e;
    e = cudaHostGetDevicePointer((void **)&d, h, 0);
    if (e != cudaSuccess) return e;
    cap = ncap;
    return cudaSuccess;
  }
  ~MappedVec() { if (h) cudaFreeHost(h); }
};

// A job list written by the host and read by one kernel.  Two transports: "copy" keeps the list in pinned host memory and
// sends the used part to a device buffer with one asynchronous copy ahead of the kernel, so that the kernel's blocks
// read it from L2/HBM (lists read by many blocks: pre-processing, FAST); "mapped" lets the kernel read the pinned host
// memory across PCIe (lists read once per warp while the frames still occupy the copy engine: sub-pixel, LK, RANSAC).
// UVIO_KLT_JOBS=mapped|copy forces one transport for all lists.
inline bool jobs_mapped() {
  static const bool v = [] {
    const char *e = getenv("UVIO_KLT_JOBS");
    return e && strcmp(e, "mapped") == 0;
  }();
  return v;
}
inline bool pdl_on() {   // UVIO_KLT_PDL=0 turns the programmatic dependent launches off (A/B measurements)
  static const bool v = [] {
    const char *e = getenv("UVIO_KLT_PDL");
    return !(e && e[0] == '0');
  }();
  return v;
}
inline bool jobs_copied() {
  static const bool v = [] {
    const char *e = getenv("UVIO_KLT_JOBS");
    return e && strcmp(e, "copy") == 0;
  }();
  return v;
}
template <class T>
struct JobVec {
  T *h = nullptr, *d = nullptr;
  size_t cap = 0;
  bool mapped = false;
  bool prefer_mapped = false;   // lists issued while the frames are still crossing PCIe: a copy would queue behind them
  cudaError_t reserve(size_t want) {   // contents are not preserved; call only while the stream is idle
    if (want <= cap) return cudaSuccess;
    size_t ncap = want + want / 2 + 64;
    release();
    mapped = jobs_mapped() || (prefer_mapped && !jobs_copied());
    cudaError_t e = cudaHostAlloc((void **)&h, ncap * sizeof(T), mapped ? cudaHostAllocMapped : cudaHostAllocDefault);
    if (e != cudaSuccess) return e;
    if (mapped) e = cudaHostGetDevicePointer((void **)&d, h, 0);
    else e = cudaMalloc((void **)&d, ncap * sizeof(T));
    if (e != cudaSuccess) return e;
    cap = ncap;
    return cudaSuccess;
  }
  cudaError_t upload(cudaStream_t st, size_t n) {
    if (mapped || n == 0) return cudaSuccess;
    return cudaMemcpyAsync(d, h, n * sizeof(T), cudaMemcpyHostToDevice, st);
  }
  void release() {
    if (h) cudaFreeHost(h);
    if (d && !mapped) cudaFree(d);
    h = d = nullptr;
    cap = 0;
  }
  ~JobVec() { release(); }
};

template <class T>
struct DevVec {
  T *d = nullptr;
  size_t cap = 0;
  cudaError_t reserve(size_t want, bool zero = false) {
    if (want <= cap) return cudaSuccess;
    size_t ncap = want + want / 2 + 64;
    if (d) cudaFree(d);
    d = nullptr;
    cap = 0;
    cudaError_t e = cudaMalloc((void **)&d, ncap * sizeof(T));
    if (e != cudaSuccess) return e;
    if (zero) {
      // cudaMemset on device memory may return before it has run, and the lanes' streams are non-blocking: without the
      // wait a kernel of the lane could start on counters that are zeroed under it (seen as a rare lost corner)
      e = cudaMemset(d, 0, ncap * sizeof(T));
      if (e == cudaSuccess) e = cudaStreamSynchronize(nullptr);
    }
    cap = ncap;
    return e;
  }
  ~DevVec() { if (d) cudaFree(d); }
};

class Context;
inline std::atomic<bool> g_contexts_gone{false};   // set when process exit has torn the device contexts down

// after the stream was synchronised: every (kernel mark, next mark) interval is one kernel's duration
inline void fold_prof(KernelProf &prof, double *ms_sum, unsigned long long *launches) {
  for (int i = 0; i + 1 < prof.n; i++) {
    if (prof.id[i] < 0) continue;
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, prof.ev[i], prof.ev[i + 1]) == cudaSuccess) {
      ms_sum[prof.id[i]] += ms;
      launches[prof.id[i]] += 1;
    }
  }
  prof.n = 0;
}

// One frame to upload and pre-process: the shared front of a feed (frames that were not announced) and of submit_batch.
struct PreItem {
  CamState *cam;
  HostImage im;                  // by value: an announced batch is issued after submit_batch has returned
  int slot;                      // destination pyramid slot
  int method;                    // TrackBase::HistogramMethod
  bool downsample;
  int par;                       // parity of the frame number: which of the camera's two upload buffers it lands in
};
struct PreBuffers {              // job lists of one upload round (a lane's, or one entry of the context's submit ring)
  JobVec<PreJob> *pre;
  JobVec<DownJob> *down;
  DevVec<uint8_t> *clahe;
};

// One lane = one CUDA stream with its own job / result buffers.  A feed_batch call splits its trackers over the lanes and
// runs them on separate host threads, so the host bookkeeping of one group overlaps the GPU rounds of the others.
struct Lane {
  Lane() {
    sp_items.prefer_mapped = lk_items.prefer_mapped = rs_jobs.prefer_mapped = true;
    const char *e = getenv("UVIO_KLT_FAST_MAPPED");   // the FAST list is read by every tile block of a cell: copied (measured)
    fast_jobs.prefer_mapped = e && e[0] == '1';
  }
  Context *dev = nullptr;
  int index = 0;
  cudaStream_t stream = nullptr;      // detection and tracking rounds
  // The new frames travel and are pre-processed on a stream of their own: detection runs on the previous frame's pyramid
  // (TrackKLT.cpp:129, :245), so only the temporal LK of round 3 has to wait for them (up_done).
  cudaStream_t up_stream = nullptr;
  cudaEvent_t up_done = nullptr;
  int cuda_err = 0;
  unsigned long long stats[8] = {0, 0, 0, 0, 0, 0, 0, 0};   // accumulated into the context after every feed
  // host wall-clock per segment of lane_feed (us): issue r1, wait r1, host+issue r2, wait r2, host+issue r3, wait r3, host D, calls
  double seg_us[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  // per-kernel CUDA-event timing (off by default)
  KernelProf prof;
  bool prof_on = false;
  double prof_ms[K_COUNT] = {0};
  unsigned long long prof_launches[K_COUNT] = {0};
  void prof_begin() {
    if (!prof_on) return;
    if (!prof.ready) {
      for (int i = 0; i < KernelProf::MAX_MARKS; i++) cudaEventCreate(&prof.ev[i]);
      prof.ready = true;
    }
    g_prof = &prof;   // marks accumulate (a pass's background part records too) until prof_fold takes them
  }
  void prof_end() {
    if (g_prof == &prof) g_prof = nullptr;
  }
  void prof_fold() {   // after the stream was synchronised
    if (!prof_on || g_prof != &prof) return;
    fold_prof(prof, prof_ms, prof_launches);
  }
  // round 1
  JobVec<DownJob> down_jobs;     // downsample_cameras: frames (upload stream)
  JobVec<DownJob> mask_jobs;     // ... and non-zero masks (main stream, finished before the host bookkeeping starts)
  DevVec<uint8_t> mask_in;       // the full-size masks of this feed
  MappedVec<uint8_t> mask_out;   // their halved versions, written by the kernel into host memory
  JobVec<PreJob> pre_jobs;
  std::vector<PreItem> pre_items;
  std::vector<HostImage> zero_masks;
  // Detection issued ahead: when a feed ends, the occupancy pass and the FAST launch of the NEXT feed of the same trackers
  // depend only on state that is final by then (pts_last, the frame that will be "last", its mask), so they are issued
  // before the call returns and run while the caller reads the observations and prepares the next message.  The next
  // pass on this lane uses them if it is made of the same units and nothing they read has changed since.
  struct Ahead {
    bool valid = false;
    std::vector<unsigned long long> serial;   // per unit: rig serial, rig epoch, unit index in the rig, stereo flag
    std::vector<unsigned> epoch;
    std::vector<unsigned long long> feeds;    // the tracker's feed count when the detection was issued: a feed through
                                              // another lane in between (another batch composition) makes it stale
    std::vector<Unit *> unit;
    std::vector<char> stereo;
    // the cameras each unit held: units are indexed by message position, and a tracker may be fed cam0-only, then cam1-only
    // messages, or the same pair in another order (TrackBase.h:60-65; all state is keyed by camera id, TrackKLT.cpp:80-89)
    std::vector<CamState *> cam_l, cam_r;
    int njobs = 0;
    // level 1: FAST issued.  level 2: the whole detection (top-K, sub-pixel, stereo LK, ids) was finished in the background
    // after the feed returned (Pass::background); currid_before = the trackers' id counters before it, put back if the
    // result cannot be used.
    int level = 0;
    std::vector<size_t> currid_before;
    int nsp = 0, nlk = 0;
  } ahead;
  DevVec<uint8_t> clahe_lut;
  JobVec<FastJob> fast_jobs;
  DevVec<uint32_t> fast_out_dev;
  DevVec<unsigned> fast_cnt_dev, fast_ticket;
  MappedVec<uint32_t> fast_out;
  MappedVec<unsigned> fast_cnt;
  // round 2 / 3
  DevVec<float> sp_mask;
  bool sp_mask_ready = false;
  JobVec<SubpixItem> sp_items;   // prefer_mapped set in Lane()
  MappedVec<float> sp_out;
  DevVec<float> sp_out_dev;
  JobVec<LkItem> lk_items;
  MappedVec<LkOut> lk_out;
  MappedVec<uint8_t> lk_status;
  DevVec<LkOut> lk_out_dev;
  DevVec<uint8_t> lk_status_dev;
  JobVec<RansacJob> rs_jobs;
  DevVec<float> p0n_dev, p1n_dev;
  MappedVec<float> p1n;
  MappedVec<uint8_t> rs_mask;
  DevVec<uint8_t> rs_mask_dev;

  // device-side pipeline (Pass::device_feed): unit descriptors and the batch-wide job / result arrays, all in HBM
  JobVec<DvUnit> dv_units;
  JobVec<DvHead> dv_heads;
  std::vector<DvUnit> dv_units_sent;   // what the device copies hold: descriptors go up only when the batch changed shape
  std::vector<DvHead> dv_heads_sent;
  DevVec<unsigned> dv_feed_bits;
  DevVec<FastJob> dv_fast_jobs;
  DevVec<uint32_t> dv_fast_out;
  DevVec<unsigned> dv_fast_cnt;
  DevVec<int> dv_fast_active;    // [slots] + the counter in its last word
  cudaStream_t lk_stream = nullptr;   // the issue-bound LK launches of the device-side pipeline run at low priority, so that the
  cudaEvent_t lk_go = nullptr, lk_done = nullptr;   // latency-bound kernels of another group get SM slots as soon as LK blocks retire
  DevVec<SubpixItem> dv_sp_items;
  DevVec<float> dv_sp_out;
  DevVec<LkItem> dv_lk1_items, dv_lk2_items, dv_lk2n_items;
  DevVec<int> dv_lk2n_out, dv_lk1_idx, dv_n_items;
  cudaEvent_t occ_done = nullptr;
#ifndef UVIO_NO_GRAPHS
  // the two kernel runs of the chain that have no external dependency inside them, captured once per batch shape
  cudaGraphExec_t graph_detect = nullptr, graph_tail = nullptr;
  unsigned long long graph_sig = 0, seen_sig = 0;
#endif
  DevVec<LkOut> dv_lk1_out, dv_lk2_out;
  DevVec<uint8_t> dv_lk1_status, dv_lk2_status, dv_rs_mask, dv_rs_mask_raw;
  DevVec<RansacJob> dv_rs_jobs;
  DevVec<float> dv_p0n, dv_p1n;

  int sync() {
    cudaError_t e = cudaStreamSynchronize(stream);
    if (e != cudaSuccess) { cuda_err = (int)e; return E_CUDA; }
    return OK;
  }
  ~Lane() {
    if (stream) cudaStreamDestroy(stream);
    if (up_stream) cudaStreamDestroy(up_stream);
    if (up_done) cudaEventDestroy(up_done);
    if (lk_stream) cudaStreamDestroy(lk_stream);
    if (lk_go) cudaEventDestroy(lk_go);
    if (lk_done) cudaEventDestroy(lk_done);
    if (occ_done) cudaEventDestroy(occ_done);
#ifndef UVIO_NO_GRAPHS
    if (graph_detect) cudaGraphExecDestroy(graph_detect);
    if (graph_tail) cudaGraphExecDestroy(graph_tail);
#endif
  }
};

class Pool;

class Context {
 public:
  int device = 0;
  int cuda_err = 0;
  std::mutex mtx;
  // counters since creation: [0] frames fed, [1] kernel launches, [2] bytes uploaded, [3] bytes the GPU wrote to host memory,
  // [4] feed_batch calls, [5] pixels scanned by FAST
  unsigned long long stats[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  double seg_us[8] = {0, 0, 0, 0, 0, 0, 0, 0};   // sum over lanes of Lane::seg_us
  cudaEvent_t ev[2] = {nullptr, nullptr};
  std::vector<std::unique_ptr<Lane>> lanes;   // threads * depth of them
  int threads = 1;                            // host threads a feed_batch call uses (caller included)
  Pool *pool = nullptr;
  // tables (host copy + device copy, re-uploaded when dirty; read-only while lanes run)
  std::vector<PyrDesc> pyr_tab;
  std::vector<ImgView> raw_tab;
  std::vector<CamModel> cam_tab;
  DevVec<PyrDesc> pyr_dev;
  DevVec<CamModel> cam_dev;
  bool tables_dirty = true;
  std::vector<void *> allocs;   // image memory of all cameras
  struct RawSlot { size_t bytes; uint8_t *p[2]; };
  std::vector<struct Pass> *bg_passes = nullptr;   // the passes of the last feed_batch: their background part runs after it returned
  void forget_rig(Rig *rig);       // a tracker is going away: detections issued ahead that list it are dropped
  void background_wait();          // returns when no background part is running (pool->finish)
  std::vector<RawSlot> raw_free;   // upload buffers of destroyed trackers, reused by cameras of the same size
  void release_camera(CamState &c);
  // submit_batch: frames announced ahead of their feed go up and are pre-processed on a stream of their own, one event
  // per call out of a small ring; a feed makes its lanes wait for the events that cover its frames
  static constexpr int PF_RING = 8;
  cudaStream_t pf_stream = nullptr;        // kernels of the announced batches
  cudaStream_t pf_copy_stream = nullptr;   // their H2D copies: batch k+1 goes up while batch k is pre-processed
  cudaEvent_t pf_copied[PF_RING] = {nullptr};
  cudaEvent_t pf_evt[PF_RING] = {nullptr};
  bool pf_used[PF_RING] = {false};
  bool pf_recorded[PF_RING] = {false};             // (uploader thread only) the entry's event has been recorded at least once
  unsigned pf_seq = 0;
  JobVec<PreJob> pf_pre[PF_RING];
  JobVec<DownJob> pf_down[PF_RING];
  DevVec<uint8_t> pf_clahe[PF_RING];
  std::vector<PreItem> pf_items[PF_RING];
  // The queued work of submit_batch is issued by a thread of its own, so the caller gets on with the feed of the frame
  // before (the issue costs ~60 us of CUDA calls per batch).  pf_issued = batches whose event has been recorded.
  std::thread uploader;
  std::mutex pf_mtx;
  std::condition_variable pf_cv;
  std::vector<int> pf_queue;                       // ring entries waiting to be issued, in order
  unsigned long long pf_queued = 0, pf_issued = 0; // batches handed to / finished by the uploader (under pf_mtx)
  unsigned long long pf_entry_seq[PF_RING] = {0};  // batch number last issued from each ring entry
  bool pf_quit = false;
  int pf_error = OK;
  int pf_group = 0;
  std::atomic<unsigned long long> pf_stats[3];     // [1] launches, [2] bytes uploaded (indices as in stats[])
  // per-kernel timing of the announced batches (same switch as the lanes'): marks of a batch are folded before the next
  KernelProf pf_prof;
  std::atomic<bool> pf_prof_on{false};
  unsigned pf_prof_mask = ~0u;
  double pf_prof_ms[K_COUNT] = {0};
  unsigned long long pf_prof_launches[K_COUNT] = {0};
  void uploader_loop();
  void uploader_drain();                           // returns when everything queued has been issued
  int wait_issued(unsigned long long seq);         // returns once batch number `seq` has been issued
  unsigned long long pf_queued_next() {            // number the batch being assembled will get
    std::lock_guard<std::mutex> lk(pf_mtx);
    return pf_queued + 1;
  }
  uint8_t *raw_arena[2] = {nullptr, nullptr};   // next free byte of the current upload-buffer chunk (one arena per frame parity)
  size_t raw_arena_left[2] = {0, 0};

  int upload_tables() {
    Context *ctx = this;
    if (!tables_dirty) return OK;
    uploader_drain();   // it reads the tables that are about to move
    CU_TRY(cudaDeviceSynchronize());
    CU_TRY(pyr_dev.reserve(pyr_tab.size()));
    CU_TRY(cam_dev.reserve(cam_tab.size()));
    if (!pyr_tab.empty()) CU_TRY(cudaMemcpy(pyr_dev.d, pyr_tab.data(), pyr_tab.size() * sizeof(PyrDesc), cudaMemcpyHostToDevice));
    if (!cam_tab.empty()) CU_TRY(cudaMemcpy(cam_dev.d, cam_tab.data(), cam_tab.size() * sizeof(CamModel), cudaMemcpyHostToDevice));
    tables_dirty = false;
    return OK;
  }
  // device memory for one camera: a raw upload buffer and two pyramids (plus the halved frame with downsample_cameras)
  int alloc_camera(CamState &c, int in_w, int in_h, int pyr_levels, int win, bool downsample) {
    Context *ctx = this;
    // cv::Size(cols / 2.0, rows / 2.0): the doubles are truncated (VioManager.cpp:274,276)
    const int w = downsample ? in_w / 2 : in_w, h = downsample ? in_h / 2 : in_h;
    if (w < 1 || h < 1) return E_ARG;
    const int L = pyr_max_level(w, h, pyr_levels, win);
    if (L + 1 > MAX_LEVELS) return E_ARG;
    auto pitch_of = [](int ww) { return (ww + 127) / 128 * 128; };
    // the upload buffer keeps the rows as tight as 16-byte loads allow, so that a tightly packed host frame is one 1-D copy
    const int raw_pitch = (in_w + 15) / 16 * 16;
    size_t raw_bytes = ((size_t)raw_pitch * in_h + 255) / 256 * 256, pyr_bytes = 0;
    const int half_pitch = (w + 15) / 16 * 16;
    const size_t half_bytes = downsample ? ((size_t)half_pitch * h + 255) / 256 * 256 : 0;
    int lw = w, lh = h;
    for (int l = 0; l <= L; l++) {
      pyr_bytes += (size_t)pitch_of(lw) * lh;
      lw = (lw + 1) / 2;
      lh = (lh + 1) / 2;
    }
    // Upload buffers come from an arena in creation order: cameras created one after the other (the two cameras of a rig,
    // the rigs of a batch) get adjacent buffers, so frames that are adjacent in host memory go up as one copy.
    uint8_t *raw_base[2] = {nullptr, nullptr};
    for (size_t k = 0; k < raw_free.size(); k++)
      if (raw_free[k].bytes == raw_bytes) {   // a destroyed tracker's buffers of the same size
        raw_base[0] = raw_free[k].p[0];
        raw_base[1] = raw_free[k].p[1];
        raw_free.erase(raw_free.begin() + (long)k);
        break;
      }
    for (int par = 0; par < 2 && !raw_base[par]; par++) {
      if (raw_arena_left[par] < raw_bytes) {
        const size_t chunk = std::max<size_t>(raw_bytes, (size_t)64 << 20);
        CU_TRY(cudaMalloc((void **)&raw_arena[par], chunk + 256));
        CU_TRY(cudaMemset(raw_arena[par], 0, chunk + 256));
        allocs.push_back(raw_arena[par]);
        raw_arena_left[par] = chunk;
      }
      raw_base[par] = raw_arena[par];
      raw_arena[par] += raw_bytes;
      raw_arena_left[par] -= raw_bytes;
    }
    uint8_t *base = nullptr;
    CU_TRY(cudaMalloc((void **)&base, PYR_SLOTS * pyr_bytes + half_bytes + 256));
    CU_TRY(cudaMemset(base, 0, PYR_SLOTS * pyr_bytes + half_bytes + 256));
    allocs.push_back(base);
    c.dev_base = base;
    c.raw_ptr[0] = raw_base[0];
    c.raw_ptr[1] = raw_base[1];
    c.raw_bytes = raw_bytes;
    c.w = w;
    c.h = h;
    c.in_w = in_w;
    c.in_h = in_h;
    for (int par = 0; par < 2; par++) {
      c.raw_in[par] = (int)raw_tab.size();
      raw_tab.push_back(ImgView{raw_base[par], in_w, in_h, raw_pitch});
    }
    c.raw = -1;
    uint8_t *p = base;
    if (downsample) {
      c.raw = (int)raw_tab.size();
      raw_tab.push_back(ImgView{p, w, h, half_pitch});
      p += half_bytes;
    }
    for (int s = 0; s < PYR_SLOTS; s++) {
      PyrDesc d;
      memset(&d, 0, sizeof(d));
      d.nlv = L + 1;
      lw = w;
      lh = h;
      for (int l = 0; l <= L; l++) {
        d.lv[l] = ImgView{p, lw, lh, pitch_of(lw)};
        p += (size_t)pitch_of(lw) * lh;
        lw = (lw + 1) / 2;
        lh = (lh + 1) / 2;
      }
      c.pyr[s] = (int)pyr_tab.size();
      pyr_tab.push_back(d);
      c.mask[s].assign((size_t)w * h, 0);
      c.mask_nonzero[s] = false;
    }
    tables_dirty = true;
    return OK;
  }
  ~Context();
};

// Minimal persistent worker pool: run(n, fn) executes fn(0) on the caller and fn(1..n-1) on workers, then joins.  Workers
// spin briefly for the next job before they sleep (feed_batch calls arrive back to back when a pipeline is busy), and the
// caller spins for the join: both waits are far shorter than a futex round trip.
class Pool {
 public:
  explicit Pool(int workers) {
    for (int i = 0; i < workers; i++) threads.emplace_back([this] { loop(); });
  }
  ~Pool() {
    {
      std::lock_guard<std::mutex> lk(m);
      quit.store(true);
    }
    cv.notify_all();
    for (auto &t : threads) t.join();
  }
  void run(int n, const std::function<void(int)> &fn) {
    if (n <= 1 || threads.empty()) {
      for (int i = 0; i < n; i++) fn(i);
      return;
    }
    {
      std::lock_guard<std::mutex> lk(m);
      job = &fn;
      total = n;
      pending.store(n - 1);
      const unsigned long long g = generation.load() + 1;
      next.store((g << 32) | 1ull);   // claims carry the generation, so a straggler of an older run can never take one
      generation.store(g, std::memory_order_release);
    }
    if (sleepers.load() > 0) cv.notify_all();
    fn(0);
    drain(generation.load(), fn, n);   // help with whatever is left, then wait for the stragglers
    while (pending.load(std::memory_order_acquire) != 0) std::this_thread::yield();
    std::lock_guard<std::mutex> lk(m);
    job = nullptr;
  }
  // Asynchronous variant: start() hands fn(0..n-1) to the workers and returns; finish() helps with what is left and waits.
  // (One job at a time: run() and start() must not be called while a started job has not been finished.)
  void start(int n, std::function<void(int)> fn) {
    async_fn = std::move(fn);
    if (threads.empty()) {   // no workers: the work is done when it is waited for
      async_n = n;
      async_active = true;
      async_inline = true;
      return;
    }
    {
      std::lock_guard<std::mutex> lk(m);
      job = &async_fn;
      total = n;
      pending.store(n);
      const unsigned long long g = generation.load() + 1;
      next.store(g << 32);
      generation.store(g, std::memory_order_release);
    }
    if (sleepers.load() > 0) cv.notify_all();
    async_n = n;
    async_active = true;
    async_inline = false;
  }
  void finish() {
    if (!async_active) return;
    if (async_inline) {
      for (int i = 0; i < async_n; i++) async_fn(i);
    } else {
      drain(generation.load(), async_fn, async_n);
      while (pending.load(std::memory_order_acquire) != 0) std::this_thread::yield();
      std::lock_guard<std::mutex> lk(m);
      job = nullptr;
    }
    async_active = false;
  }
  bool busy() const { return async_active; }

 private:
  std::function<void(int)> async_fn;
  int async_n = 0;
  bool async_active = false, async_inline = false;
  void loop() {
    unsigned long long seen = 0;
    while (true) {
      const auto spin_until = std::chrono::steady_clock::now() + std::chrono::microseconds(300);
      while (generation.load(std::memory_order_acquire) == seen && !quit.load(std::memory_order_relaxed) &&
             std::chrono::steady_clock::now() < spin_until) {
#if defined(__x86_64__)
        __builtin_ia32_pause();
#endif
      }
      if (generation.load(std::memory_order_acquire) == seen) {
        std::unique_lock<std::mutex> lk(m);
        sleepers.fetch_add(1);
        cv.wait(lk, [&] { return quit.load() || generation.load() != seen; });
        sleepers.fetch_sub(1);
      }
      if (quit.load()) return;
      const std::function<void(int)> *fn;
      int n;
      {
        std::lock_guard<std::mutex> lk(m);   // pairs with run(): job/total are published under the mutex
        seen = generation.load();
        fn = job;
        n = total;
      }
      if (!fn) continue;
      drain(seen, *fn, n);
    }
  }
  void drain(unsigned long long gen, const std::function<void(int)> &fn, int n) {
    unsigned long long cur = next.load();
    while ((cur >> 32) == (gen & 0xffffffffull) && (int)(cur & 0xffffffffull) < n) {
      if (next.compare_exchange_weak(cur, cur + 1)) {
        fn((int)(cur & 0xffffffffull));
        pending.fetch_sub(1, std::memory_order_release);
        cur = next.load();
      }
    }
  }
  std::vector<std::thread> threads;
  std::mutex m;
  std::condition_variable cv;
  const std::function<void(int)> *job = nullptr;
  int total = 0;
  std::atomic<int> pending{0}, sleepers{0};
  std::atomic<unsigned long long> next{0}, generation{0};
  std::atomic<bool> quit{false};
};

// ---------------------------------------------------------------------------------------------------------------------
// One "unit" = what the reference tracks as a whole: a stereo pair (feed_stereo) or a single camera (feed_monocular).
// ---------------------------------------------------------------------------------------------------------------------
struct Side {                         // one camera of a unit during one feed
  CamState *cam = nullptr;
  int msg = -1;                       // index in the message
  int det_slot = 0;                   // pyramid slot detection runs on (last frame; current frame on the first feed)
  const uint8_t *det_mask = nullptr;  // mask of that image (null = all zeros)
  OccState occ;
  bool detect = false;                // FAST jobs were emitted for this side (left: final; right: speculative)
  GridGeom geom;
  std::vector<std::pair<int, int>> cells;              // cells FAST ran on
  std::vector<int> job_of_cell;                        // job index per cell (-1: cell skipped, Grider_GRID.h:117)
  std::vector<std::vector<KeyPointLite>> top;          // per cell: top-K in sort order (no mask test yet)
  std::vector<int> sp_of_cell;                         // first sub-pixel item of each cell
  std::vector<Pt> pts;                                 // pts_old after detection (input of the temporal LK)
  std::vector<size_t> ids;
  int lk_off = 0, lk_n = 0;                            // slice in the temporal LK arrays
  bool lk_run = false;                                 // >= 10 points (TrackKLT.cpp:848)
};

struct Unit {
  Rig *rig = nullptr;
  FeedRequest *req = nullptr;
  bool stereo = false, first = false;
  Side L, R;
  std::vector<KeyPointLite> selL;     // left corners that passed the mask test, in cell order (sub-pixel items sp0L..)
  int sp0L = 0, lk0 = 0;              // first sub-pixel item / first stereo-LK item of selL
  bool failed = false;
  // scratch that keeps its capacity from feed to feed
  std::vector<Pt> good_l, good_r, refined, new_r, ext;
  std::vector<size_t> ids_l, ids_r, obs_l_idx, obs_r_idx, new_r_ids;
  std::vector<int> keep;
  IdTable tab_a, tab_b;
};

// the units of one pass: pointers into the rigs' pools, iterated as references
struct UnitList {
  std::vector<Unit *> v;
  struct It {
    Unit *const *p;
    Unit &operator*() const { return **p; }
    It &operator++() { ++p; return *this; }
    bool operator!=(const It &o) const { return p != o.p; }
  };
  It begin() const { return It{v.data()}; }
  It end() const { return It{v.data() + v.size()}; }
  size_t size() const { return v.size(); }
  void push_back(Unit *u) { v.push_back(u); }
};

// ---------------------------------------------------------------------------------------------------------------------
// Device-side pipeline: per-tracker memory.  One device slab (state + work areas) and one mapped host slab (the state's
// mirror and the observation list) per tracker, laid out for the tracker's feature budget and image sizes.
// ---------------------------------------------------------------------------------------------------------------------
struct DvCamMem {
  float *pts_last = nullptr; unsigned long long *ids_last = nullptr; int *n_last = nullptr;        // device
  float *pts = nullptr; unsigned long long *ids = nullptr;
  uint8_t *close = nullptr; int *first = nullptr, *grid = nullptr, *cells = nullptr, *top_n = nullptr, *ctl = nullptr;
  uint32_t *top = nullptr;
  float *h_pts_last = nullptr; unsigned long long *h_ids_last = nullptr; int *h_n_last = nullptr;  // mapped: host pointers
  float *m_pts_last = nullptr; unsigned long long *m_ids_last = nullptr; int *m_n_last = nullptr;  // ... and their device aliases
  uint8_t *mask[PYR_SLOTS] = {nullptr, nullptr, nullptr};                                          // device copies of non-zero masks
};
struct DvUnitMem {
  int *scr[4] = {nullptr, nullptr, nullptr, nullptr}, *part = nullptr, *uctl = nullptr;
  float *tmp_pts = nullptr; unsigned long long *tmp_ids = nullptr;
  DvObs *h_obs = nullptr, *m_obs = nullptr, *d_obs = nullptr; int *h_n_obs = nullptr, *m_n_obs = nullptr;
};
struct DvRig {
  int nf = 0, P = 0, C = 0, K = 0, size_x = 0, size_y = 0, cell_cap = 0, obs_cap = 0;
  std::vector<int> cam_w, cam_h;
  void *slab = nullptr, *host_slab = nullptr;
  unsigned long long *currid = nullptr, *h_currid = nullptr, *m_currid = nullptr;   // device; mapped host pointer; its device alias
  std::vector<DvCamMem> cam;
  std::vector<DvUnitMem> unit;
  ~DvRig() {
    if (slab) cudaFree(slab);
    if (host_slab) cudaFreeHost(host_slab);
  }
};

inline double now_us() {
  return std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

// keep a caller mask for the detection of the next feed (TrackKLT.cpp:129,245 detect on img_mask_last); returns whether any
// pixel is non-zero.  An all-zero mask (the usual case) is only scanned, not copied.
inline bool mask_any_set(const HostImage &m, int w, int h) {
  if (!m.p) return false;
  bool nz = false;
  for (int y = 0; y < h && !nz; y++) {
    const uint8_t *r = m.p + (size_t)y * m.stride;
    uint64_t acc = 0;
    int x = 0;
    for (; x + 32 <= w; x += 32) {
      uint64_t a, b, c, d;
      memcpy(&a, r + x, 8);
      memcpy(&b, r + x + 8, 8);
      memcpy(&c, r + x + 16, 8);
      memcpy(&d, r + x + 24, 8);
      acc |= (a | b) | (c | d);
    }
    for (; x < w; x++) acc |= r[x];
    nz = acc != 0;
  }
  return nz;
}
inline bool copy_mask(const HostImage &m, std::vector<uint8_t> &dst, int w, int h) {
  HOSTPROF(copy_mask);
  if (!mask_any_set(m, w, h)) return false;
  for (int y = 0; y < h; y++) memcpy(dst.data() + (size_t)y * w, m.p + (size_t)y * m.stride, (size_t)w);
  return true;
}

// Everything one lane does for its share of a feed_batch call, cut at the three stream synchronisations into four stages so
// that one host thread can drive two lanes alternately: while it does the host bookkeeping of one lane, the GPU works on the
// rounds the other lane has already issued.
struct Pass {
  Lane *ctx = nullptr;
  FeedRequest **reqs = nullptr;
  int n = 0;
  UnitList units;
  int worst = OK;
  int rc = OK;          // E_CUDA once a stage failed; later stages are skipped
  int stage = 0;
  int npre = 0, njobs = 0, nsp = 0, nlk = 0, nrs = 0;
  unsigned wait_evts = 0;   // events of the context's submit ring that cover announced frames of this pass
  double t0 = 0, t1 = 0, t2 = 0, t3 = 0, tmark = 0;
  // make the lane's main stream wait for the new pyramids: its own upload stream and the submit rounds that hold its frames
  cudaError_t wait_new_frames();
  // occupancy passes + FAST jobs of every unit; launches FAST unless a first frame has to wait for its upload
  int issue_detection(bool ahead, bool &any_first, int &max_cw, int &max_ch);
  bool ahead_usable();
  void issue_ahead();
  bool detected = false;   // this pass found its whole detection finished by the previous pass's background part
  void background();       // runs after feed_batch has returned (pool workers): rounds 1-2 of the NEXT feed of these units

  void fail(FeedRequest &r, int code) {
    r.status = code;
    r.rig->last_error = code;
    if (code > worst) worst = code;
  }
  void seg(int k) {   // close host-profile segment k at the current time
    double t = now_us();
    ctx->seg_us[k] += t - tmark;
    tmark = t;
  }
  int stage0();
  int stage0_book();
  int stage0_issue();
  int device_feed();     // the whole feed as one chain of launches with the bookkeeping on the device (track_body.h): issue
  int device_collect();  // ... wait for it and take the results over
  int stage1();
  int stage2();
  int stage3();
  // run the next stage (synchronising the lane's stream first when the stage consumes GPU results)
  void step() {
    if (detected && stage == 1) {   // nothing to wait for, nothing to do: the detection is already there
      stage++;
      return;
    }
    if (rc == OK && n > 0) {
      cudaSetDevice(ctx->dev->device);
      g_prof = ctx->prof_on ? &ctx->prof : nullptr;
      if (stage > 0) {
        if (ctx->sync() != OK) rc = E_CUDA;
        else {
          seg(2 * stage - 1);
          if (stage == 3) ctx->prof_fold();   // every stream of the lane is idle now
        }
      }
      if (rc == OK) {
        int e = stage == 0 ? stage0() : stage == 1 ? stage1() : stage == 2 ? stage2() : stage3();
        if (e != OK) rc = e;
      }
      g_prof = nullptr;
    }
    stage++;
  }
  int result() const { return rc != OK ? rc : worst; }
};


// ---- functions defined in one translation unit and used by another ---------------------------------------------------------
void dv_destroy(DvRig *d);
bool device_pipeline_enabled();
int dv_prepare(Context *ctx, Rig &rig);
int issue_preprocess(Context *ctx, cudaStream_t st, const PreBuffers &B, const std::vector<PreItem> &items, unsigned long long *stats,
                     int group = 0, cudaStream_t copy_st = nullptr, cudaEvent_t copied = nullptr);
bool ahead_enabled();
bool background_enabled();
void cancel_ahead(Lane &ln);
int validate_request(Context *ctx, FeedRequest &rq, bool announcing);

}  // namespace uvio
