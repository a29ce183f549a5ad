// Host-callable launch wrappers of the sm_100a kernels (implemented in kernels.cu).  Plain C++ types only.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "fast_body.h"
#include "ransac_body.h"
#include "simt.h"

namespace uvio {

constexpr int MAX_LEVELS = 8;          // pyramid levels per image (level 0 included)
constexpr int HIST_BANDS = 16;         // row bands (thread blocks) per image in the histogram / LUT passes
constexpr int CLAHE_TILES = 8;         // TrackKLT.cpp:62 (8x8 tiles)

struct PyrDesc {                       // one image pyramid resident in HBM
  ImgView lv[MAX_LEVELS];
  int nlv;                             // levels built (max_level + 1)
};

struct PreJob {                        // one image to pre-process this frame
  ImgView src;                         // the frame in device-visible memory: the upload buffer, or the caller's own device memory
  int pyr;                             // destination pyramid
  int method;                          // 0 none, 1 equalizeHist, 2 CLAHE   (TrackBase::HistogramMethod)
};

struct DownJob {                       // one image or mask to halve before the tracker sees it (VioManager.cpp:268-278)
  ImgView src;                         // full-size frame (upload buffer or the caller's device memory) or mask
  ImgView dst;                         // floor(w/2) x floor(h/2); device memory, or mapped host memory for masks
};

struct SubpixItem { float x, y; int pyr; int pad; };
struct LkItem { float px, py, nx, ny; int prev_pyr, next_pyr; };
struct LkOut { float x, y; };

struct RansacJob {
  int off, n;                          // slice of the LK item / output arrays (the points of one camera of one rig)
  int cam0, cam1;                      // camera models used to undistort the old / new points
  double thr;                          // 2.0 / max focal length (TrackKLT.cpp:870-873)
  int run;                             // 0: fewer than 10 points -> mask stays zero (TrackKLT.cpp:848-852), only undistort
};

// Optional per-kernel timing: when g_prof is set, every launch wrapper records a CUDA event before each kernel it launches
// (and the engine one after the last), on the launching stream.
enum KernelId { K_HIST = 0, K_LUT_APPLY, K_CLAHE_TILES, K_CLAHE_APPLY, K_PYRDOWN, K_FAST, K_SUBPIX, K_LK, K_RANSAC, K_PREPROCESS, K_DOWNSAMPLE, K_DV_LOGIC, K_DV_TOPK, K_COUNT };
struct KernelProf {
  static constexpr int MAX_MARKS = 48;
  cudaEvent_t ev[MAX_MARKS];
  int id[MAX_MARKS];
  int n = 0;
  bool ready = false;
  unsigned mask = ~0u;   // kernels to time (bit = KernelId)
  bool open = false;     // the last mark recorded started a kernel interval
};
extern thread_local KernelProf *g_prof;   // defined in engine.cpp; set by the lane thread that is launching
inline void prof_mark(cudaStream_t st, int kernel_id) {
  KernelProf *p = g_prof;
  if (!p || p->n >= KernelProf::MAX_MARKS) return;
  const bool wanted = kernel_id >= 0 && ((p->mask >> kernel_id) & 1u);
  if (!wanted && !p->open) return;   // neither starts a timed kernel nor ends one
  cudaEventRecord(p->ev[p->n], st);
  p->id[p->n++] = wanted ? kernel_id : -1;
  p->open = wanted;
}

// ---- image-wide passes ------------------------------------------------------------------------------------------------
// equalise (or copy) raw frames into level 0 of their pyramids and build the remaining levels: one cluster launch
cudaError_t launch_preprocess(cudaStream_t st, const PreJob *jobs_dev, int njobs, const PyrDesc *pyr_tab_dev, uint8_t *clahe_lut_dev /*njobs*64*256, CLAHE only*/, int any_clahe,
                              int max_band_bytes /* max over the jobs of (ceil(h/8)+1) * w: the rows one block stages by TMA */,
                              int max_w = 0, int max_h = 0, int max_nlv = 0 /* largest frame and deepest pyramid of the jobs: the levels then go in launches of their own */);

// cv::pyrDown(src, dst, Size(cols / 2.0, rows / 2.0)) of every job: the downsample_cameras step in front of the tracker
// pyramid launches launch_preprocess makes for frames of at most max_w x max_h with max_nlv levels (level 1, and level 2 of
// large frames, as launches of their own; the remaining levels in one): for the launch statistics
inline int pyramid_launches(int max_w, int max_h, int max_nlv) {
  if (max_nlv <= 1 || max_w <= 0 || max_h <= 0) return 0;
  int w = max_w, h = max_h, l = 1, n = 0;
  for (; l < max_nlv; l++) {
    w = (w + 1) / 2;
    h = (h + 1) / 2;
    if (l > 1 && ((w + 3) / 4) * ((h + 3) / 4) <= 4096) break;
    n++;
  }
  return n + (l < max_nlv ? 1 : 0);
}

cudaError_t launch_downsample(cudaStream_t st, const DownJob *jobs_dev, int njobs, int max_dst_w, int max_dst_h);

// ---- detection ----------------------------------------------------------------------------------------------------------
// FAST + NMS on grid cells; compacts each job's corner list into out_host (mapped pinned) and writes its count.
cudaError_t launch_fast(cudaStream_t st, const FastJob *jobs, int njobs, int max_cell_w, int max_cell_h, const PyrDesc *pyr_tab_dev,
                        uint32_t *out_dev, unsigned *count_dev,
                        unsigned *ticket_dev, uint32_t *out_host, unsigned *count_host);

cudaError_t launch_subpix(cudaStream_t st, const SubpixItem *items, int n, const PyrDesc *pyr_tab_dev, const float *mask_dev,
                          float *out_xy /*2n*/, float *out_xy_dev /*2n, may be null*/,
                          const int *n_dev = nullptr /* device-built compact list: its length; item.pad = output slot */);

// ---- tracking -----------------------------------------------------------------------------------------------------------
// If init_from_dev != null, item i takes (px,py,nx,ny) from init_from_dev[2*i..] (used to chain subpix -> stereo LK).
cudaError_t launch_lk(cudaStream_t st, const LkItem *items, int n, const PyrDesc *pyr_tab_dev, const float *init_from_dev,
                      int max_count, double eps, LkOut *out_host, uint8_t *status_host, LkOut *out_dev, uint8_t *status_dev,
                      bool overlap = false /* programmatic dependent launch: may start while the kernel before it in the stream runs */,
                      const int *out_idx_dev = nullptr /* result i goes to out[out_idx[i]] instead of out[i] */,
                      const int *n_dev = nullptr /* device-built compact list: its length (n is then its capacity) */);

// undistort old/new points of every job and run the fundamental-matrix RANSAC; mask = klt_status && inlier
cudaError_t launch_ransac(cudaStream_t st, const RansacJob *jobs, int njobs, int max_n, const LkItem *items, const LkOut *lk_dev,
                          const uint8_t *status_dev, const CamModel *cams_dev, float *p0n_dev, float *p1n_dev,
                          float *p1n_host, uint8_t *mask_host, uint8_t *mask_dev, bool overlap = false);

// ---- device-side bookkeeping (track_body.h): occupancy -> FAST -> top-K -> select -> [sub-pixel, stereo LK] -> assign ->
// [temporal LK, RANSAC] -> merge, all on one stream without a host round trip
struct DvUnit;
struct DvHead;
struct DvGlobals;
struct DvFeedBits;
cudaError_t launch_dv_occupancy(cudaStream_t st, const DvUnit *units_dev, const DvHead *heads_dev, int n_heads, const DvGlobals &X, int max_p,
                                const DvFeedBits &bits /* by value in the kernel parameters: what changed since the last feed */);
cudaError_t launch_dv_fast(cudaStream_t st, const DvGlobals &X, const PyrDesc *pyr_tab_dev, int n_slots, int max_cell_w, int max_cell_h);
cudaError_t launch_dv_topk(cudaStream_t st, const DvUnit *units_dev, int n_units, int max_c, const DvGlobals &X);
cudaError_t launch_dv_select(cudaStream_t st, const DvUnit *units_dev, const DvHead *heads_dev, int n_heads, const DvGlobals &X, int max_p);
cudaError_t launch_dv_assign(cudaStream_t st, const DvUnit *units_dev, const DvHead *heads_dev, int n_heads, const DvGlobals &X, int max_p);
cudaError_t launch_dv_merge(cudaStream_t st, const DvUnit *units_dev, const DvHead *heads_dev, int n_heads, const DvGlobals &X, int max_p);

cudaError_t upload_subpix_mask(float *mask_dev, cudaStream_t st);
// LK work counters since the last reset: [0] point-levels (templates staged), [1] iterations, [2] points
cudaError_t read_lk_work(unsigned long long out[4], int reset);

}  // namespace uvio
