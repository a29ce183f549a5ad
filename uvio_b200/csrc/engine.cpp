// Per-device context and the batched frame pipeline of the GPU TrackKLT.  See engine.h for the division of labour.
//
// One frame of a batch of rigs costs three device round trips, whatever the batch size:
//   round 1  upload + equalise + pyramid of the new frames; FAST on every cell that may need new features
//            (left cells, and a superset of the right cells that is cheap to prove sufficient - see phase_a)
//   round 2  cornerSubPix of the per-cell top-K corners; left->right LK of the new left corners (chained on the device)
//   round 3  temporal LK of every tracked point, undistortion, fundamental-matrix RANSAC
// Between rounds the host runs the reference's order-dependent bookkeeping (hostlogic.h).
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

thread_local KernelProf *g_prof = nullptr;

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
static HostProfTable g_hostprof;
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
static bool jobs_mapped() {
  static const bool v = [] {
    const char *e = getenv("UVIO_KLT_JOBS");
    return e && strcmp(e, "mapped") == 0;
  }();
  return v;
}
static bool pdl_on() {   // UVIO_KLT_PDL=0 turns the programmatic dependent launches off (A/B measurements)
  static const bool v = [] {
    const char *e = getenv("UVIO_KLT_PDL");
    return !(e && e[0] == '0');
  }();
  return v;
}
static bool jobs_copied() {
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
static std::atomic<bool> g_contexts_gone{false};   // set when process exit has torn the device contexts down

// after the stream was synchronised: every (kernel mark, next mark) interval is one kernel's duration
static void fold_prof(KernelProf &prof, double *ms_sum, unsigned long long *launches) {
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
      (e = cudaStreamCreateWithPriority(&c->pf_copy_stream, cudaStreamNonBlocking, prio_lo)) != cudaSuccess) {
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

struct DvRig;
static void dv_destroy(DvRig *d);   // defined with DvRig below

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

static const uint8_t *mask_ptr(CamState &c, int slot) { return c.mask_nonzero[slot] ? c.mask[slot].data() : nullptr; }

// emit FAST jobs for the cells of one side
static void emit_fast_jobs(Lane *ctx, Side &s, int threshold, int &njobs, size_t &nwords) {
  HOSTPROF(emit_fast_jobs);
  const int cols = s.cam->w, rows = s.cam->h;
  s.job_of_cell.assign(s.cells.size(), -1);
  for (size_t r = 0; r < s.cells.size(); r++) {
    int x = s.cells[r].first * s.geom.size_x, y = s.cells[r].second * s.geom.size_y;
    if (x + s.geom.size_x > cols || y + s.geom.size_y > rows) continue;   // Grider_GRID.h:117
    FastJob j;
    j.img = s.cam->pyr[s.det_slot];
    j.x0 = x; j.y0 = y; j.w = s.geom.size_x; j.h = s.geom.size_y;
    j.out_off = (int)nwords;
    j.out_cap = ((j.w + 1) / 2) * ((j.h + 1) / 2);
    j.threshold = threshold;
    ctx->fast_jobs.h[njobs] = j;
    s.job_of_cell[r] = njobs++;
    nwords += (size_t)j.out_cap;
  }
}

// Phase A: occupancy passes and the list of cells to run FAST on.
// ahead: run at the end of a feed for the NEXT one (the frame that is current now will be "last" then).
static void phase_a(Unit &u, bool ahead = false) {
  HOSTPROF(phase_a);
  const Params &P = u.rig->prm;
  Side &L = u.L;
  const int thr = std::min(20, (int)(0.5 * P.num_features));
  {
    CamState &c = *L.cam;
    L.det_slot = (u.first || ahead) ? c.cur : c.last();
    L.det_mask = mask_ptr(c, L.det_slot);
    { HOSTPROF(a_occ_init); occ_init(L.occ, c.w, c.h, P.min_px_dist, P.grid_x, P.grid_y); }
    HOSTPROF(a_left_rest);
    if (!u.first) occ_feed(L.occ, c.pts_last.data(), c.ids_last.data(), c.pts_last.size(), L.det_mask, nullptr);
    const int need = P.num_features - (int)L.occ.pts.size();
    // mono returns when need < thr (TrackKLT.cpp:470); stereo detects when need > thr (:609)
    L.detect = u.stereo ? (need > thr) : !(need < thr);
    L.cells.clear();
    if (L.detect) {
      valid_cells(L.occ, L.det_mask, P.num_features, L.cells);
      L.geom = grider_geom(c.w, c.h, P.num_features, P.grid_x, P.grid_y);
    }
  }
  if (u.stereo) {
    Side &R = u.R;
    CamState &c = *R.cam;
    R.det_slot = (u.first || ahead) ? c.cur : c.last();
    R.det_mask = mask_ptr(c, R.det_slot);
    occ_init(R.occ, c.w, c.h, P.min_px_dist, P.grid_x, P.grid_y);
    // The right pass (:722-780) runs after the left detection appended new stereo points to pts1, but it is a plain
    // in-order loop: its first |pts_last| steps only see the old points, and the new ids are fresh, so they cannot
    // change the stereo-id exemption of an old point. Run those steps now; phase C continues the same loop.
    IdTable &stereo_ids = u.tab_a;
    { HOSTPROF(a_idtable); stereo_ids.fill(L.occ.ids.data(), L.occ.ids.size()); }
    HOSTPROF(a_right_feed);
    if (!u.first) occ_feed(R.occ, c.pts_last.data(), c.ids_last.data(), c.pts_last.size(), R.det_mask, &stereo_ids);
    // Cells: new points only raise counts and lower the need, so the cells valid now are a superset of the cells the
    // reference will extract from; FAST + top-K + sub-pixel run on the superset, phase C keeps the right subset.
    const int need = P.num_features - (int)R.occ.pts.size();
    R.detect = need > thr;
    R.cells.clear();
    if (R.detect) {
      valid_cells(R.occ, R.det_mask, P.num_features, R.cells);
      R.geom = grider_geom(c.w, c.h, P.num_features, P.grid_x, P.grid_y);
    }
  }
}

// Phase B: per-cell top-K (reference sort order), left mask test, sub-pixel / stereo-LK items.
static void phase_b(Lane *ctx, Unit &u, int &nsp, int &nlk, std::vector<uint32_t> &tmp, std::vector<uint32_t> &grid) {
  HOSTPROF(phase_b);
  for (int side = 0; side < (u.stereo ? 2 : 1); side++) {
    Side &s = side ? u.R : u.L;
    if (s.top.size() < s.cells.size()) s.top.resize(s.cells.size());   // entries beyond cells.size() are stale scratch
    for (size_t r = 0; r < s.cells.size(); r++) s.top[r].clear();
    if (!s.detect) continue;
    HOSTPROF(b_topk);
    for (size_t r = 0; r < s.cells.size(); r++) {
      int j = s.job_of_cell[r];
      if (j < 0) continue;
      const FastJob &job = ctx->fast_jobs.h[j];
#ifdef UVIO_HOSTPROF
      { static long long cells = 0, corners = 0; cells++; corners += ctx->fast_cnt.h[j]; if ((cells & 255) == 0) fprintf(stderr, "[hostprof] cells %lld corners %lld\n", cells, corners); }
#endif
      cell_top_k(ctx->fast_out.h + job.out_off, ctx->fast_cnt.h[j], job.w, job.h, job.x0, job.y0, s.geom.per_cell, tmp, grid, s.top[r]);
    }
  }
  // left: mask test against mask_updated = mask | painted boxes (Grider_GRID.h:142-148), then sub-pixel
  Side &L = u.L;
  u.selL.clear();
  u.sp0L = nsp;
  if (L.detect) {
    HOSTPROF(b_masktest);
    const int cols = L.cam->w, rows = L.cam->h;
    for (size_t r = 0; r < L.cells.size(); r++)
      for (auto &k : L.top[r]) {
        int x = (int)k.x, y = (int)k.y;
        if (x < 0 || x > cols || y < 0 || y > rows) continue;
        size_t at = (size_t)y * cols + x;
        if ((L.det_mask && L.det_mask[at] > 127) || occ_boxed(L.occ, at)) continue;
        u.selL.push_back(k);
      }
  }
  for (auto &k : u.selL) ctx->sp_items.h[nsp++] = SubpixItem{k.x, k.y, L.cam->pyr[L.det_slot], 0};
  u.lk0 = nlk;
  if (u.stereo) {
    // left -> right LK of every selected left corner, chained on the sub-pixel output (item i <-> sub-pixel item i)
    for (size_t i = 0; i < u.selL.size(); i++)
      ctx->lk_items.h[nlk++] = LkItem{0, 0, 0, 0, L.cam->pyr[L.det_slot], u.R.cam->pyr[u.R.det_slot]};
  }
}

static void phase_b_right(Lane *ctx, Unit &u, int &nsp) {
  HOSTPROF(phase_b_right);
  Side &R = u.R;
  R.sp_of_cell.assign(R.cells.size(), -1);
  if (!R.detect) return;
  for (size_t r = 0; r < R.cells.size(); r++) {
    R.sp_of_cell[r] = nsp;
    for (auto &k : R.top[r]) ctx->sp_items.h[nsp++] = SubpixItem{k.x, k.y, R.cam->pyr[R.det_slot], 0};
  }
}

// Phase C: finish the detection with the refined corners (min-distance reject, ids, right side), then either store the
// first-frame result or emit the temporal LK / RANSAC work.
static int phase_c_detect(Lane *ctx, Unit &u) {
  HOSTPROF(phase_c_detect);
  const Params &P = u.rig->prm;
  Side &L = u.L;
  const int thr = std::min(20, (int)(0.5 * P.num_features));
  std::vector<Pt> &refined = u.refined;
  refined.resize(u.selL.size());
  for (size_t i = 0; i < u.selL.size(); i++) refined[i] = Pt{ctx->sp_out.h[2 * (u.sp0L + i)], ctx->sp_out.h[2 * (u.sp0L + i) + 1]};
  std::vector<int> &keep = u.keep;
  close_reject(L.occ, refined.data(), refined.size(), keep);       // TrackKLT.cpp:497-512 / :635-650
  std::vector<Pt> &new_r = u.new_r;
  std::vector<size_t> &new_r_ids = u.new_r_ids;
  new_r.clear();
  new_r_ids.clear();
  if (!u.stereo) {
    for (int i : keep) {                                            // :519-527
      L.occ.pts.push_back(refined[i]);
      L.occ.ids.push_back(++u.rig->currid);
    }
  } else {
    Side &R = u.R;
    const int cols0 = L.cam->w, rows0 = L.cam->h, cols1 = R.cam->w, rows1 = R.cam->h;
    for (int i : keep) {                                            // :675-708
      const Pt p0 = refined[i];
      const LkOut o = ctx->lk_out.h[u.lk0 + i];
      const bool oob_left = (int)p0.x < 0 || (int)p0.x >= cols0 || (int)p0.y < 0 || (int)p0.y >= rows0;
      const bool oob_right = (int)o.x < 0 || (int)o.x >= cols1 || (int)o.y < 0 || (int)o.y >= rows1;
      if (!oob_left && !oob_right && ctx->lk_status.h[u.lk0 + i] == 1) {
        size_t id = ++u.rig->currid;
        L.occ.pts.push_back(p0);
        L.occ.ids.push_back(id);
        new_r.push_back(Pt{o.x, o.y});
        new_r_ids.push_back(id);
      } else if (!oob_left) {
        L.occ.pts.push_back(p0);
        L.occ.ids.push_back(++u.rig->currid);
      }
    }
    // right: continue the occupancy loop over the appended stereo points (:722-780), then the right top-up (:784-826)
    IdTable &stereo_ids = u.tab_a;
    stereo_ids.fill(L.occ.ids.data(), L.occ.ids.size());
    occ_feed(R.occ, new_r.data(), new_r_ids.data(), new_r.size(), R.det_mask, &stereo_ids);
    const int need = P.num_features - (int)R.occ.pts.size();
    if (need > thr) {
      std::vector<std::pair<int, int>> cells;
      valid_cells(R.occ, R.det_mask, P.num_features, cells);
      std::vector<Pt> &ext = u.ext;
      ext.clear();
      const uint8_t *mask0 = (cols0 == cols1 && rows0 == rows1) ? L.det_mask : nullptr;   // :721 clones mask0 (sic)
      for (auto &cell : cells) {
        size_t r = 0;
        while (r < R.cells.size() && R.cells[r] != cell) r++;
        if (r == R.cells.size() || !R.detect) return E_ARG;   // cannot happen: R.cells is a superset (phase_a)
        for (size_t k = 0; k < R.top[r].size(); k++) {
          const KeyPointLite &kp = R.top[r][k];
          int x = (int)kp.x, y = (int)kp.y;
          if (x < 0 || x > cols1 || y < 0 || y > rows1) continue;
          size_t at = (size_t)y * cols1 + x;
          if ((mask0 && mask0[at] > 127) || occ_boxed(R.occ, at)) continue;
          int sp = R.sp_of_cell[r] + (int)k;
          ext.push_back(Pt{ctx->sp_out.h[2 * sp], ctx->sp_out.h[2 * sp + 1]});
        }
      }
      close_reject(R.occ, ext.data(), ext.size(), keep);            // :811-825
      for (int i : keep) {
        R.occ.pts.push_back(ext[i]);
        R.occ.ids.push_back(++u.rig->currid);
      }
    }
    R.pts.swap(R.occ.pts);
    R.ids.swap(R.occ.ids);
  }
  L.pts.swap(L.occ.pts);
  L.ids.swap(L.occ.ids);
  return OK;
}

static void store_last(Side &s, std::vector<Pt> &pts, std::vector<size_t> &ids) {
  s.cam->pts_last.swap(pts);
  s.cam->ids_last.swap(ids);
  s.cam->have_last = true;
}

static void emit_temporal(Lane *ctx, Unit &u, int &nlk, int &nrs) {
  HOSTPROF(emit_temporal);
  for (int side = 0; side < (u.stereo ? 2 : 1); side++) {
    Side &s = side ? u.R : u.L;
    CamState &c = *s.cam;
    s.lk_off = nlk;
    s.lk_n = 0;
    s.lk_run = s.pts.size() >= 10;                                  // TrackKLT.cpp:848-852
    if (!s.lk_run) continue;
    s.lk_n = (int)s.pts.size();
    for (auto &p : s.pts) ctx->lk_items.h[nlk++] = LkItem{p.x, p.y, p.x, p.y, c.pyr[c.last()], c.pyr[c.cur]};
    RansacJob j;
    j.off = s.lk_off;
    j.n = s.lk_n;
    j.cam0 = j.cam1 = c.cam_model_idx;
    j.thr = 2.0 / c.max_focal;                                      // :870-873 (id0 == id1 for the temporal match)
    j.run = 1;
    ctx->rs_jobs.h[nrs++] = j;
  }
}

// Phase D: merge / filter the matches and produce the feature-database observations (TrackKLT.cpp:143-189, :285-381).
static void phase_d(Lane *ctx, Unit &u) {
  HOSTPROF(phase_d);
  Rig &rig = *u.rig;
  Side &L = u.L, &R = u.R;
  auto new_pt = [&](Side &s, size_t i) { const LkOut o = ctx->lk_out.h[s.lk_off + i]; return Pt{o.x, o.y}; };
  auto ok = [&](Side &s, size_t i) { return s.lk_run && ctx->rs_mask.h[s.lk_off + i]; };
  auto add_obs = [&](Side &s, size_t i, size_t id, Pt p) {
    rig.obs.push_back(Obs{id, s.cam->cam_id, p.x, p.y, ctx->p1n.h[2 * (s.lk_off + i)], ctx->p1n.h[2 * (s.lk_off + i) + 1]});
  };
  std::vector<Pt> &good_l = u.good_l, &good_r = u.good_r;   // scratch kept in the unit; swapped into pts_last at the end
  std::vector<size_t> &ids_l = u.ids_l, &ids_r = u.ids_r;
  good_l.clear(); good_r.clear(); ids_l.clear(); ids_r.clear();
  good_l.reserve(L.pts.size()); ids_l.reserve(L.pts.size());
  good_r.reserve(R.pts.size()); ids_r.reserve(R.pts.size());
  rig.obs.reserve(L.pts.size() + R.pts.size());
  if (!u.stereo) {
    CamState &c = *L.cam;
    if (L.pts.empty()) {                                            // :143-152 (mask_ll empty)
      store_last(L, good_l, ids_l);
      return;
    }
    const uint8_t *mask = mask_ptr(c, c.cur);                       // :166 tests the current message mask
    for (size_t i = 0; i < L.pts.size(); i++) {                     // :159-173
      if (!L.lk_run) break;
      Pt p = new_pt(L, i);
      if (p.x < 0 || p.y < 0 || (int)p.x >= c.w || (int)p.y >= c.h) continue;
      if (mask && mask[(size_t)(int)p.y * c.w + (int)p.x] > 127) continue;
      if (ok(L, i)) {
        good_l.push_back(p);
        ids_l.push_back(L.ids[i]);
        add_obs(L, i, L.ids[i], p);
      }
    }
    store_last(L, good_l, ids_l);
    return;
  }
  if (L.pts.empty() && R.pts.empty()) {                             // :285-299
    store_last(L, good_l, ids_l);
    store_last(R, good_r, ids_r);
    return;
  }
  const int cols_l = L.cam->w, rows_l = L.cam->h, cols_r = R.cam->w, rows_r = R.cam->h;
  IdTable &first_right = u.tab_a;                                   // id -> first index in ids_right_old (:314-320)
  first_right.fill(R.ids.data(), R.ids.size());
  std::vector<size_t> &obs_l_idx = u.obs_l_idx, &obs_r_idx = u.obs_r_idx;
  obs_l_idx.clear(); obs_r_idx.clear();
  obs_l_idx.reserve(L.pts.size()); obs_r_idx.reserve(R.pts.size());
  if (L.lk_run)
    for (size_t i = 0; i < L.pts.size(); i++) {                     // :306-338
      Pt p = new_pt(L, i);
      if (p.x < 0 || p.y < 0 || (int)p.x > cols_l || (int)p.y > rows_l) continue;   // '>' on the left (:308-309)
      const int ir = first_right.find(L.ids[i]);
      const bool found_right = ir >= 0;
      if (ok(L, i) && found_right && ok(R, (size_t)ir)) {
        Pt q = new_pt(R, (size_t)ir);
        if (q.x < 0 || q.y < 0 || (int)q.x >= cols_r || (int)q.y >= rows_r) continue;
        good_l.push_back(p);
        ids_l.push_back(L.ids[i]);
        obs_l_idx.push_back(i);
        good_r.push_back(q);
        ids_r.push_back(R.ids[ir]);
        obs_r_idx.push_back((size_t)ir);
      } else if (ok(L, i)) {
        good_l.push_back(p);
        ids_l.push_back(L.ids[i]);
        obs_l_idx.push_back(i);
      }
    }
  if (R.lk_run) {
    IdTable &added = u.tab_b;
    added.reset(ids_r.size() + R.pts.size());
    for (size_t k = 0; k < ids_r.size(); k++) added.insert(ids_r[k], (int)k);
    for (size_t i = 0; i < R.pts.size(); i++) {                     // :341-354
      Pt q = new_pt(R, i);
      if (q.x < 0 || q.y < 0 || (int)q.x >= cols_r || (int)q.y >= rows_r) continue;
      if (ok(R, i) && added.find(R.ids[i]) < 0) {
        good_r.push_back(q);
        ids_r.push_back(R.ids[i]);
        obs_r_idx.push_back(i);
        added.insert(R.ids[i], (int)i);
      }
    }
  }
  for (size_t k = 0; k < good_l.size(); k++) add_obs(L, obs_l_idx[k], ids_l[k], good_l[k]);   // :357-366 left, then right
  for (size_t k = 0; k < good_r.size(); k++) add_obs(R, obs_r_idx[k], ids_r[k], good_r[k]);
  store_last(L, good_l, ids_l);
  store_last(R, good_r, ids_r);
}

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

static void dv_destroy(DvRig *d) { delete d; }

static bool device_pipeline_enabled() {   // UVIO_KLT_PIPELINE=host selects the round-1 path (bookkeeping on host threads)
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
static int dv_prepare(Context *ctx, Rig &rig) {
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

static double now_us() {
  return std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

// keep a caller mask for the detection of the next feed (TrackKLT.cpp:129,245 detect on img_mask_last); returns whether any
// pixel is non-zero.  An all-zero mask (the usual case) is only scanned, not copied.
static bool mask_any_set(const HostImage &m, int w, int h) {
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
static bool copy_mask(const HostImage &m, std::vector<uint8_t> &dst, int w, int h) {
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

// H2D of the frames (merged into as few copies as their layout allows), the optional halving, equalisation and pyramids,
// all on stream st.  Frames already in device memory are consumed where they lie.
// copy_st / copied: when given, the copies go to a stream of their own and st waits for them through the event, so that the
// upload of one batch overlaps the kernels of the batch before it (the frames alternate between two upload buffers).
static int issue_preprocess(Context *ctx, cudaStream_t st, const PreBuffers &B, const std::vector<PreItem> &items, unsigned long long *stats,
                            int group = 0, cudaStream_t copy_st = nullptr, cudaEvent_t copied = nullptr) {
  const bool split = copy_st != nullptr && copy_st != st;
  if (!split) copy_st = st;   // (with a copy stream the groups still apply: each group's kernels wait for its own copies only)
  HOSTPROF(s0_memcpy);
  const int nimg = (int)items.size();
  if (nimg == 0) return OK;
  int any_clahe = 0, max_band = 0, ndown = 0, max_dw = 1, max_dh = 1, n_down_jobs = 0, max_w = 0, max_h = 0, max_nlv = 0;
  std::vector<char> on_dev((size_t)nimg, 0);
  for (auto &it : items) {
    any_clahe |= it.method == 2;
    n_down_jobs += it.downsample;
  }
  CU_TRY(B.pre->reserve((size_t)nimg));
  if (n_down_jobs) CU_TRY(B.down->reserve((size_t)n_down_jobs));
  if (any_clahe) CU_TRY(B.clahe->reserve((size_t)nimg * CLAHE_TILES * CLAHE_TILES * 256));
  // Copies whose source and destination both continue the previous one are merged (a batch of tightly packed frames is one copy).
  uint8_t *pend_dst = nullptr;
  const uint8_t *pend_src = nullptr;
  size_t pend_bytes = 0;
  // A long copy is cut into pieces: the rounds of the feed that runs meanwhile read their job lists and write their results
  // across the same PCIe link, and one 20 MB transfer ahead of them costs each round ~100 us (measured: tools/pcie_probe.py);
  // between pieces the link takes their requests.
  static const size_t piece = [] {
    const char *e = getenv("UVIO_KLT_UPLOAD_PIECE");
    long v = e ? atol(e) : (1l << 20);
    return (size_t)(v < 4096 ? 4096 : v);
  }();
  auto flush = [&]() -> cudaError_t {
    cudaError_t e = cudaSuccess;
    for (size_t off = 0; off < pend_bytes && e == cudaSuccess; off += piece)
      e = cudaMemcpyAsync(pend_dst + off, pend_src + off, std::min(piece, pend_bytes - off), cudaMemcpyHostToDevice, copy_st);
    pend_bytes = 0;
    return e;
  };
  // Pass 1: the job lists (one small copy each, ahead of the frames).
  { HOSTPROF(pre_pass1);
  for (int j = 0; j < nimg; j++) {
    const PreItem &it = items[j];
    CamState &c = *it.cam;
    const HostImage &im = it.im;
    B.pre->h[j] = PreJob{ctx->raw_tab[it.downsample ? c.raw : c.raw_in[it.par]], c.pyr[it.slot], it.method};
    max_band = std::max(max_band, ((c.h + 7) / 8 + 1) * c.w);
    max_w = std::max(max_w, c.w);
    max_h = std::max(max_h, c.h);
    max_nlv = std::max(max_nlv, ctx->pyr_tab[c.pyr[it.slot]].nlv);
    cudaPointerAttributes attr;
    const bool on_device = cudaPointerGetAttributes(&attr, im.p) == cudaSuccess &&
                           (attr.type == cudaMemoryTypeDevice || attr.type == cudaMemoryTypeManaged);
    if (!on_device) cudaGetLastError();   // an unregistered host pointer is not an error
    on_dev[j] = on_device;
    if (it.downsample) {
      const ImgView full = on_device ? ImgView{im.p, im.w, im.h, (int)im.stride} : ctx->raw_tab[c.raw_in[it.par]];
      B.down->h[ndown++] = DownJob{full, ctx->raw_tab[c.raw]};
      max_dw = std::max(max_dw, c.w);
      max_dh = std::max(max_dh, c.h);
    } else if (on_device) {
      B.pre->h[j].src = ImgView{im.p, im.w, im.h, (int)im.stride};
    }
  }
  }
  CU_TRY(B.pre->upload(copy_st, (size_t)nimg));
  if (ndown > 0) CU_TRY(B.down->upload(copy_st, (size_t)ndown));
  // Pass 2: group by group, the frames go up and are (halved and) pre-processed.  A feed issues one group; submit_batch cuts
  // a batch into several so that the pyramids of the first frames are built while the last ones are still on the link, and
  // so that no launch takes every SM from the rounds of the feed that is running meanwhile.
  if (group <= 0 || group > nimg) group = nimg;
  int down_at = 0;
  HOSTPROF(pre_pass2);
  for (int g0 = 0; g0 < nimg; g0 += group) {
    const int g1 = std::min(nimg, g0 + group);
    int gdown = 0;
    for (int j = g0; j < g1; j++) {
      const PreItem &it = items[j];
      const HostImage &im = it.im;
      const ImgView &raw = ctx->raw_tab[it.cam->raw_in[it.par]];
      gdown += it.downsample;
      if (on_dev[j]) continue;
      stats[2] += (unsigned long long)im.w * im.h;
      if ((size_t)raw.pitch == im.stride) {
        const size_t bytes = (size_t)raw.pitch * (im.h - 1) + im.w, span = (size_t)raw.pitch * im.h;
        if (pend_bytes && pend_src + pend_bytes == im.p && pend_dst + pend_bytes == raw.p) {
          pend_bytes += span;
        } else {
          CU_TRY(flush());
          pend_dst = (uint8_t *)raw.p;
          pend_src = im.p;
          pend_bytes = span;
        }
        if (span != bytes && (size_t)im.w != im.stride) {   // never read past the caller's last row
          pend_bytes -= span - bytes;
          CU_TRY(flush());
        }
      } else {
        CU_TRY(flush());
        CU_TRY(cudaMemcpy2DAsync((void *)raw.p, (size_t)raw.pitch, im.p, im.stride, (size_t)im.w, (size_t)im.h, cudaMemcpyDefault, copy_st));
      }
    }
    CU_TRY(flush());
    if (split) {
      CU_TRY(cudaEventRecord(copied, copy_st));
      CU_TRY(cudaStreamWaitEvent(st, copied, 0));
    }
    if (gdown > 0) {   // downsample_cameras: halve the frames where they landed, in front of the pre-processing
      CU_TRY(launch_downsample(st, B.down->d + down_at, gdown, max_dw, max_dh));
      down_at += gdown;
      stats[1] += 1;
    }
    HOSTPROF(s0_launch_preprocess);
    CU_TRY(launch_preprocess(st, B.pre->d + g0, g1 - g0, ctx->pyr_dev.d, any_clahe ? B.clahe->d + (size_t)g0 * CLAHE_TILES * CLAHE_TILES * 256 : nullptr,
                             any_clahe, max_band, max_w, max_h, max_nlv));
    stats[1] += (unsigned)(1 + 2 * any_clahe);
  }
  return OK;
}

cudaError_t Pass::wait_new_frames() {
  cudaError_t e = cudaStreamWaitEvent(ctx->stream, ctx->up_done, 0);
  for (int i = 0; i < Context::PF_RING && e == cudaSuccess; i++)
    if ((wait_evts >> i) & 1u) e = cudaStreamWaitEvent(ctx->stream, ctx->dev->pf_evt[i], 0);
  return e;
}

int Pass::issue_detection(bool ahead, bool &any_first, int &max_cw, int &max_ch) {
  Context *dev = ctx->dev;
  size_t ncells = 0, nwords = 0;
  max_cw = max_ch = 1;
  any_first = false;
  for (auto &u : units) {
    if (u.failed) continue;
    phase_a(u, ahead);
    any_first |= u.first;
    for (int side = 0; side < (u.stereo ? 2 : 1); side++) {
      Side &s = side ? u.R : u.L;
      if (!s.detect) continue;
      ncells += s.cells.size();
      nwords += s.cells.size() * (size_t)(((s.geom.size_x + 1) / 2) * ((s.geom.size_y + 1) / 2));
      max_cw = std::max(max_cw, s.geom.size_x);
      max_ch = std::max(max_ch, s.geom.size_y);
    }
  }
  CU_TRY(ctx->fast_jobs.reserve(ncells));
  CU_TRY(ctx->fast_cnt.reserve(ncells));
  CU_TRY(ctx->fast_out.reserve(nwords));
  CU_TRY(ctx->fast_out_dev.reserve(nwords));
  CU_TRY(ctx->fast_cnt_dev.reserve(ncells, true));
  CU_TRY(ctx->fast_ticket.reserve(ncells, true));
  njobs = 0;
  nwords = 0;
  for (auto &u : units) {
    if (u.failed) continue;
    for (int side = 0; side < (u.stereo ? 2 : 1); side++) {
      Side &s = side ? u.R : u.L;
      if (s.detect) emit_fast_jobs(ctx, s, u.rig->prm.fast_threshold, njobs, nwords);
    }
  }
  CU_TRY(ctx->fast_jobs.upload(ctx->stream, (size_t)njobs));
  if (!any_first) {
    HOSTPROF(s0_fast_launch);
    CU_TRY(launch_fast(ctx->stream, ctx->fast_jobs.d, njobs, max_cw, max_ch, dev->pyr_dev.d, ctx->fast_out_dev.d, ctx->fast_cnt_dev.d,
                       ctx->fast_ticket.d, ctx->fast_out.d, ctx->fast_cnt.d));
  }
  return OK;
}

static bool ahead_enabled() {   // UVIO_KLT_AHEAD=0: every feed issues its own detection (A/B measurements)
  static const bool v = [] {
    const char *e = getenv("UVIO_KLT_AHEAD");
    return !(e && e[0] == '0');
  }();
  return v;
}

// A background detection that will not be used: the id counters it advanced go back to where they were.  Every tracker in
// the list is unfed since (a feed of any of them through another lane cancels first: feed_batch).
static void cancel_ahead(Lane &ln) {
  Lane::Ahead &a = ln.ahead;
  if (a.valid && a.level == 2) {
    for (size_t i = a.unit.size(); i-- > 0;) {   // in reverse: a tracker with several units ends at its earliest value
      a.unit[i]->rig->currid = a.currid_before[i];
      a.unit[i]->rig->bg_lane = -1;
    }
  }
  a.valid = false;
  a.level = 0;
}

// The detection issued when the previous pass of this lane ended serves this pass if the units are the same and nothing
// it read has changed: same trackers (by serial: an address may be reused), no set_num_features / change_feat_id since,
// no feed of any of them through another lane since, same stereo / mono shape, and no unit starting over (a tracker that
// lost every point detects on the NEW frame).
bool Pass::ahead_usable() {
  const Lane::Ahead &a = ctx->ahead;
  if (!a.valid || a.unit.size() != units.size()) return false;
  for (size_t i = 0; i < units.v.size(); i++) {
    const Unit *u = units.v[i];
    if (u != a.unit[i] || u->rig->serial != a.serial[i] || u->rig->epoch != a.epoch[i] || u->rig->feeds != a.feeds[i] + 1 ||
        (char)u->stereo != a.stereo[i] || u->first || u->failed || u->L.cam != a.cam_l[i] || (u->stereo && u->R.cam != a.cam_r[i]))
      return false;
  }
  return true;
}

static bool background_enabled() {   // UVIO_KLT_BACKGROUND=1 turns it on (off until it has been through the GPU suite)
  static const bool v = [] {
    const char *e = getenv("UVIO_KLT_BACKGROUND");
    return e && e[0] == '1';
  }();
  return v;
}

// End of a pass: issue the next pass's detection (see Lane::Ahead).  All or nothing per lane.
void Pass::issue_ahead() {
  Lane::Ahead &a = ctx->ahead;
  a.valid = false;
  a.level = 0;
  // (when FAST itself is being timed, every feed issues its own detection inside the timed window)
  if (!ahead_enabled() || (ctx->prof_on && ((ctx->prof.mask >> K_FAST) & 1u)) || rc != OK || units.size() == 0) return;
  for (auto &u : units) {
    if (u.failed || u.req->status != OK) return;
    const bool empty = u.stereo ? (u.L.cam->pts_last.empty() && u.R.cam->pts_last.empty()) : u.L.cam->pts_last.empty();
    if (empty) return;                              // its next feed starts over on the new frame (TrackKLT.cpp:110, :221)
  }
  a.serial.clear(); a.epoch.clear(); a.feeds.clear(); a.unit.clear(); a.stereo.clear(); a.cam_l.clear(); a.cam_r.clear();
  for (auto &u : units) {
    u.first = false;
    a.cam_l.push_back(u.L.cam);
    a.cam_r.push_back(u.stereo ? u.R.cam : nullptr);
    a.serial.push_back(u.rig->serial);
    a.epoch.push_back(u.rig->epoch);
    a.feeds.push_back(u.rig->feeds);
    a.unit.push_back(&u);
    a.stereo.push_back((char)u.stereo);
  }
  bool any_first = false;
  int cw, ch;
  if (issue_detection(true, any_first, cw, ch) != OK || any_first) return;
  a.njobs = njobs;
  a.valid = true;
  a.level = 1;
}

// After feed_batch has returned: the rest of the NEXT feed's detection for the units of this pass -- wait for the FAST that
// issue_ahead launched, per-cell top-K, sub-pixel + stereo LK (round 2), then ids and the right-image pass -- so that the
// next feed starts at its temporal LK.  Nothing here touches what the caller may read meanwhile (observations, pts_last);
// the trackers' id counters are advanced and put back by cancel_ahead if the result is not used.
void Pass::background() {
  Lane::Ahead &a = ctx->ahead;
  if (!a.valid || a.level != 1 || !background_enabled()) return;
  if (ctx->prof_on && ((ctx->prof.mask >> K_SUBPIX) & 1u)) return;   // sub-pixel timed: keep it inside the timed window
  cudaSetDevice(ctx->dev->device);
  g_prof = ctx->prof_on ? &ctx->prof : nullptr;
  auto give_up = [&] {
    a.valid = false;
    a.level = 0;
    g_prof = nullptr;
  };
  if (ctx->sync() != OK) return give_up();
  njobs = a.njobs;
  tmark = now_us();
  if (stage1() != OK || ctx->sync() != OK) return give_up();
  a.currid_before.clear();
  for (auto &u : units) a.currid_before.push_back(u.rig->currid);
  bool ok = true;
  for (auto &u : units) ok = ok && phase_c_detect(ctx, u) == OK;
  a.level = 2;            // from here on cancel_ahead puts the id counters back
  a.nsp = nsp;
  a.nlk = nlk;
  for (auto &u : units) u.rig->bg_lane = ctx->index;
  if (!ok) cancel_ahead(*ctx);
  g_prof = nullptr;
}

int Pass::stage0() {
  int e = stage0_book();
  if (e != OK) return e;
  return stage0_issue();
}

// host bookkeeping of the new messages: frame slots, masks, pre-processing jobs, units (shared by the host and the device path)
int Pass::stage0_book() {
  Context *dev = ctx->dev;
  t0 = tmark = now_us();
  ctx->prof_begin();
  g_prof = ctx->prof_on ? &ctx->prof : nullptr;
  // ---- host bookkeeping of the new messages: frame slots, masks, pre-processing jobs, units ---------------------------------
  int n_mask_jobs = 0;
  size_t mask_in_bytes = 0, mask_out_bytes = 0;
  std::vector<PreItem> &items = ctx->pre_items;
  items.clear();
  wait_evts = 0;
  std::vector<HostImage> &zero_masks = ctx->zero_masks;   // masks of this pass already found to be all zero
  zero_masks.clear();
  unsigned long long need_seq = 0;   // last announced batch this pass consumes: its event must have been recorded
  for (int i = 0; i < n; i++) {
    FeedRequest &rq = *reqs[i];
    Rig &rig = *rq.rig;
    const Params &P = rig.prm;
    rig.feeds++;
    for (int k = 0; k < rq.n; k++) {
      CamState &c = *rig.find_cam(rq.sensor_ids[k]);
      // frame number n_fed + 1 of this camera: announced earlier (submit_batch; already on its way or done) or uploaded now
      const bool announced = c.n_sub > c.n_fed;
      if (!announced) c.n_sub++;
      c.n_fed++;
      c.cur = (int)(c.n_fed % PYR_SLOTS);
      if (announced) {
        wait_evts |= 1u << c.sub_evt[c.cur];
        need_seq = std::max(need_seq, c.sub_seq[c.cur]);
      }
      else items.push_back(PreItem{&c, rq.images[k], c.cur, P.histogram_method, P.downsample_cameras != 0, (int)(c.n_sub & 1u)});
      // the same mask memory given to several cameras of this call (a rig's shared mask, a batch's common one) is scanned once
      const HostImage &mk = rq.masks[k];
      bool known_zero = false;
      for (const HostImage &z : zero_masks) known_zero |= z.p == mk.p && z.stride == mk.stride && z.w == mk.w && z.h == mk.h;
      if (known_zero) c.mask_nonzero[c.cur] = false;
      else if (!P.downsample_cameras) c.mask_nonzero[c.cur] = copy_mask(mk, c.mask[c.cur], c.w, c.h);
      else c.mask_nonzero[c.cur] = mask_any_set(mk, c.in_w, c.in_h);
      if (!c.mask_nonzero[c.cur] && mk.p && !known_zero && zero_masks.size() < 16) zero_masks.push_back(mk);
      if (P.downsample_cameras && c.mask_nonzero[c.cur]) {   // halved on the device below
        n_mask_jobs++;
        mask_in_bytes += (size_t)((c.in_w + 15) / 16 * 16) * c.in_h;
        mask_out_bytes += (size_t)((c.w + 3) / 4 * 4) * c.h;
      }
    }
    // units (TrackKLT.cpp:80-93)
    auto make_side = [&](Side &s, int k) { s.cam = rig.find_cam(rq.sensor_ids[k]); s.msg = k; };
    if (rq.n == 2 && P.use_stereo) {
      Unit &u = *rig.unit_at(0);
      u.req = &rq; u.stereo = true;
      make_side(u.L, 0);
      make_side(u.R, 1);
      u.first = u.L.cam->no_points() && u.R.cam->no_points();                                         // :221
      if (!u.first && (!u.L.cam->have_last || !u.R.cam->have_last)) { fail(rq, E_BADMSG); continue; }
      units.push_back(&u);
    } else {
      for (int k = 0; k < rq.n; k++) {
        Unit &u = *rig.unit_at((size_t)k);
        u.req = &rq; u.stereo = false;
        make_side(u.L, k);
        u.first = u.L.cam->no_points();                                                                // :110
        units.push_back(&u);
      }
    }
  }
  if (need_seq && dev->wait_issued(need_seq) != OK) { ctx->cuda_err = dev->cuda_err; return E_CUDA; }
  if (n_mask_jobs > 0) {
    // downsample_cameras with masks that mask something: cv::pyrDown of the mask (VioManager.cpp:275-277).  The host
    // bookkeeping below reads the halved mask (first frames detect on the frame being fed), so this small job runs on the
    // idle main stream and is waited for; the frames themselves go up on the upload stream meanwhile.
    CU_TRY(ctx->mask_jobs.reserve((size_t)n_mask_jobs));
    CU_TRY(ctx->mask_in.reserve(mask_in_bytes));
    CU_TRY(ctx->mask_out.reserve(mask_out_bytes));
    size_t in_off = 0, out_off = 0;
    int nj = 0, max_w = 1, max_h = 1;
    for (int i = 0; i < n; i++) {
      FeedRequest &rq = *reqs[i];
      if (!rq.rig->prm.downsample_cameras) continue;
      for (int k = 0; k < rq.n; k++) {
        CamState &c = *rq.rig->find_cam(rq.sensor_ids[k]);
        if (!c.mask_nonzero[c.cur]) continue;
        const int ip = (c.in_w + 15) / 16 * 16, op = (c.w + 3) / 4 * 4;
        CU_TRY(cudaMemcpy2DAsync(ctx->mask_in.d + in_off, (size_t)ip, rq.masks[k].p, rq.masks[k].stride, (size_t)c.in_w, (size_t)c.in_h,
                                 cudaMemcpyDefault, ctx->stream));
        ctx->mask_jobs.h[nj++] = DownJob{ImgView{ctx->mask_in.d + in_off, c.in_w, c.in_h, ip}, ImgView{ctx->mask_out.d + out_off, c.w, c.h, op}};
        in_off += (size_t)ip * c.in_h;
        out_off += (size_t)op * c.h;
        max_w = std::max(max_w, c.w);
        max_h = std::max(max_h, c.h);
        ctx->stats[2] += (unsigned long long)c.in_w * c.in_h;
        ctx->stats[3] += (unsigned long long)c.w * c.h;
      }
    }
    CU_TRY(ctx->mask_jobs.upload(ctx->stream, (size_t)nj));
    CU_TRY(launch_downsample(ctx->stream, ctx->mask_jobs.d, nj, max_w, max_h));
    ctx->stats[1] += 1;
    if (ctx->sync() != OK) return E_CUDA;
    out_off = 0;
    for (int i = 0; i < n; i++) {
      FeedRequest &rq = *reqs[i];
      if (!rq.rig->prm.downsample_cameras) continue;
      for (int k = 0; k < rq.n; k++) {
        CamState &c = *rq.rig->find_cam(rq.sensor_ids[k]);
        if (!c.mask_nonzero[c.cur]) continue;
        const int op = (c.w + 3) / 4 * 4;
        for (int y = 0; y < c.h; y++) memcpy(c.mask[c.cur].data() + (size_t)y * c.w, ctx->mask_out.h + out_off + (size_t)y * op, (size_t)c.w);
        out_off += (size_t)op * c.h;
      }
    }
  }
  if (!ctx->sp_mask_ready) {
    CU_TRY(ctx->sp_mask.reserve(SP_N));
    CU_TRY(upload_subpix_mask(ctx->sp_mask.d, ctx->stream));
    ctx->sp_mask_ready = true;
  }
  return OK;
}

int Pass::stage0_issue() {
  Context *dev = ctx->dev;
  std::vector<PreItem> &items = ctx->pre_items;
  // ---- round 1a: FAST.  Detection runs on the previous frame's pyramid, which is already in HBM, so it is issued ahead of
  // the upload (its job list would otherwise queue behind the images on the copy engine) -- or was issued when the previous
  // feed of these trackers ended.  First frames detect on the frame that is being uploaded and wait for it.
  int max_cw = 1, max_ch = 1;
  bool any_first = false;
  if (ahead_usable()) {
    njobs = ctx->ahead.njobs;
    if (ctx->ahead.level == 2) {
      detected = true;
      nsp = ctx->ahead.nsp;
      nlk = ctx->ahead.nlk;
      for (auto &u : units) u.rig->bg_lane = -1;
    }
    ctx->ahead.valid = false;
    ctx->ahead.level = 0;
  } else {
    cancel_ahead(*ctx);
    int e = issue_detection(false, any_first, max_cw, max_ch);
    if (e != OK) return e;
  }
  // ---- round 1b: upload + equalise + pyramid on the upload stream (frames announced by submit_batch are on their way) -------
  npre = (int)items.size();
  {
    const PreBuffers B{&ctx->pre_jobs, &ctx->down_jobs, &ctx->clahe_lut};
    int e = issue_preprocess(dev, ctx->up_stream, B, items, ctx->stats);
    if (e != OK) { ctx->cuda_err = dev->cuda_err; return e; }
    CU_TRY(cudaEventRecord(ctx->up_done, ctx->up_stream));
  }
  if (any_first) {
    CU_TRY(wait_new_frames());
    CU_TRY(launch_fast(ctx->stream, ctx->fast_jobs.d, njobs, max_cw, max_ch, dev->pyr_dev.d, ctx->fast_out_dev.d, ctx->fast_cnt_dev.d,
                       ctx->fast_ticket.d, ctx->fast_out.d, ctx->fast_cnt.d));
  }
  seg(0);
  return OK;
}

int Pass::stage1() {
  Context *dev = ctx->dev;
  ctx->stats[1] += njobs > 0;
  for (int j = 0; j < njobs; j++) ctx->stats[5] += (unsigned long long)ctx->fast_jobs.h[j].w * ctx->fast_jobs.h[j].h;
  for (int j = 0; j < njobs; j++) ctx->stats[3] += 4ull * ctx->fast_cnt.h[j] + 4ull;
  t1 = now_us();
  // ---- round 2: sub-pixel + stereo LK --------------------------------------------------------------------------------
  size_t sp_bound = 0;
  for (auto &u : units)
    for (int side = 0; side < (u.stereo ? 2 : 1); side++) {
      Side &s = side ? u.R : u.L;
      if (s.detect) sp_bound += s.cells.size() * (size_t)s.geom.per_cell;
    }
  CU_TRY(ctx->sp_items.reserve(sp_bound));
  CU_TRY(ctx->sp_out.reserve(2 * sp_bound));
  CU_TRY(ctx->sp_out_dev.reserve(2 * sp_bound));
  CU_TRY(ctx->lk_items.reserve(sp_bound));
  CU_TRY(ctx->lk_out.reserve(sp_bound));
  CU_TRY(ctx->lk_status.reserve(sp_bound));
  CU_TRY(ctx->lk_out_dev.reserve(sp_bound));
  CU_TRY(ctx->lk_status_dev.reserve(sp_bound));
  nsp = 0;
  nlk = 0;
  {
    // sub-pixel items: all left selections first (so that stereo-LK item i chains on sub-pixel item i), then the right
    std::vector<uint32_t> tmp;
    std::vector<uint32_t> grid;
    for (auto &u : units) phase_b(ctx, u, nsp, nlk, tmp, grid);
    // a unit's lk0 must equal its sp0L for the chaining: true when only stereo units select points; mono units add
    // sub-pixel items without LK items, so stereo units are given explicit coordinates below instead
    for (auto &u : units)
      if (u.stereo) phase_b_right(ctx, u, nsp);
  }
  HOSTPROF(s1_launches);
  CU_TRY(ctx->sp_items.upload(ctx->stream, (size_t)nsp));
  CU_TRY(launch_subpix(ctx->stream, ctx->sp_items.d, nsp, dev->pyr_dev.d, ctx->sp_mask.d, ctx->sp_out.d, ctx->sp_out_dev.d));
  // stereo LK items read their start point from the sub-pixel output: build the index map
  bool chained = true;
  for (auto &u : units) chained &= (!u.stereo || u.lk0 == u.sp0L);
  if (nlk > 0) {
    if (!chained) {
      // mixed batch: wait for the refined corners and fill the LK items on the host
      if (ctx->sync() != OK) return E_CUDA;
      for (auto &u : units)
        if (u.stereo)
          for (size_t i = 0; i < u.selL.size(); i++) {
            LkItem &it = ctx->lk_items.h[u.lk0 + i];
            it.px = it.nx = ctx->sp_out.h[2 * (u.sp0L + i)];
            it.py = it.ny = ctx->sp_out.h[2 * (u.sp0L + i) + 1];
          }
    }
    CU_TRY(ctx->lk_items.upload(ctx->stream, (size_t)nlk));
    // chained and with its items in mapped memory, the launch may overlap the sub-pixel kernel (it waits for it on the device)
    CU_TRY(launch_lk(ctx->stream, ctx->lk_items.d, nlk, dev->pyr_dev.d, chained ? ctx->sp_out_dev.d : nullptr, 30, 0.01,
                     ctx->lk_out.d, ctx->lk_status.d, ctx->lk_out_dev.d, ctx->lk_status_dev.d, chained && ctx->lk_items.mapped && pdl_on()));
  }
  seg(2);
  return OK;
}

int Pass::stage2() {
  Context *dev = ctx->dev;
  ctx->stats[1] += (nsp > 0) + (nlk > 0);
  ctx->stats[3] += 8ull * nsp + 9ull * nlk;
  // ---- host: finish detection; round 3: temporal LK + RANSAC -----------------------------------------------------------
  size_t lk_bound = 0;
  for (auto &u : units) {
    int e = detected ? OK : phase_c_detect(ctx, u);
    if (e != OK) { fail(*u.req, e); u.failed = true; continue; }
    if (u.first) {                                                   // TrackKLT.cpp:110-123 / :221-240: store, no DB update
      store_last(u.L, u.L.pts, u.L.ids);
      if (u.stereo) store_last(u.R, u.R.pts, u.R.ids);
      continue;
    }
    lk_bound += u.L.pts.size() + (u.stereo ? u.R.pts.size() : 0);
  }
  t2 = now_us();
  CU_TRY(ctx->lk_items.reserve(lk_bound));
  CU_TRY(ctx->lk_out.reserve(lk_bound));
  CU_TRY(ctx->lk_status.reserve(lk_bound));
  CU_TRY(ctx->lk_out_dev.reserve(lk_bound));
  CU_TRY(ctx->lk_status_dev.reserve(lk_bound));
  CU_TRY(ctx->rs_jobs.reserve(2 * units.size()));
  CU_TRY(ctx->p0n_dev.reserve(2 * lk_bound));
  CU_TRY(ctx->p1n_dev.reserve(2 * lk_bound));
  CU_TRY(ctx->p1n.reserve(2 * lk_bound));
  CU_TRY(ctx->rs_mask.reserve(lk_bound));
  CU_TRY(ctx->rs_mask_dev.reserve(lk_bound));
  nlk = 0;
  nrs = 0;
  int max_n = 1;
  for (auto &u : units) {
    if (u.first || u.failed) continue;
    emit_temporal(ctx, u, nlk, nrs);
    max_n = std::max(max_n, std::max(u.L.lk_n, u.R.lk_n));
  }
  HOSTPROF(s2_launches);
  CU_TRY(wait_new_frames());   // the temporal LK reads the new pyramids
  CU_TRY(ctx->lk_items.upload(ctx->stream, (size_t)nlk));
  CU_TRY(ctx->rs_jobs.upload(ctx->stream, (size_t)nrs));
  CU_TRY(launch_lk(ctx->stream, ctx->lk_items.d, nlk, dev->pyr_dev.d, nullptr, 30, 0.01, ctx->lk_out.d, ctx->lk_status.d,
                   ctx->lk_out_dev.d, ctx->lk_status_dev.d));
  CU_TRY(launch_ransac(ctx->stream, ctx->rs_jobs.d, nrs, max_n, ctx->lk_items.d, ctx->lk_out_dev.d, ctx->lk_status_dev.d,
                       dev->cam_dev.d, ctx->p0n_dev.d, ctx->p1n_dev.d, ctx->p1n.d, ctx->rs_mask.d, ctx->rs_mask_dev.d,
                       nlk > 0 && ctx->rs_jobs.mapped && ctx->lk_items.mapped && pdl_on()));
  seg(4);
  return OK;
}

int Pass::stage3() {
  ctx->stats[1] += (nlk > 0) + (nrs > 0);
  ctx->stats[3] += 9ull * nlk + (nrs > 0 ? 9ull * nlk : 0ull);
  ctx->stats[0] += (unsigned long long)n;
  t3 = now_us();
  for (auto &u : units)
    if (!u.first && !u.failed) phase_d(ctx, u);
  for (auto &u : units) {   // a device-side copy of this state, if any, is stale now
    u.rig->dv_dirty = true;
    u.rig->dv_mirror_stale = false;
    for (auto &c : u.rig->cams) c.dv_n_last = -1;
  }
  const double t4 = now_us();
  issue_ahead();
  seg(6);
  ctx->seg_us[7] += 1;
  ctx->prof_end();
  for (int i = 0; i < n; i++) {
    float *s = reqs[i]->rig->stage_us;
    s[0] = (float)(t1 - t0);   // upload + equalise + pyramid (+ FAST of round 1)
    s[1] = (float)(t2 - t1);   // detection (top-K, sub-pixel, id assignment)
    s[2] = (float)(t3 - t2);   // temporal KLT + RANSAC
    s[3] = 0.f;                // stereo KLT is part of the detection round here
    s[4] = (float)(t4 - t3);   // merge + observations
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

// The checks of TrackKLT::feed_new_camera on one message; new cameras get their device memory here.
static int validate_request(Context *ctx, FeedRequest &rq, bool announcing) {
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
