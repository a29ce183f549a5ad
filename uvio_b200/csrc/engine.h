// Host side of the GPU TrackKLT: per-rig tracker state and the reference's sequential, order-dependent control flow
// (ov_core/src/track/TrackKLT.cpp, Grider_GRID.h), plus the per-device context that batches the GPU rounds of many rigs.
//
// What runs where: pixel / feature arithmetic is on the GPU; decisions whose outcome depends on
// iteration order (occupancy grids, per-cell std::sort tie order, id assignment, stereo merge) stay on the host and
// mirror the reference statement by statement.
#pragma once
#include <stddef.h>
#include <stdint.h>

#include <mutex>
#include <unordered_map>
#include <utility>
#include <vector>

#include "kernels.h"

namespace uvio {

struct Params {                 // TrackKLT ctor arguments (TrackKLT.h:54-57) + the reference's hard-coded constants
  int num_features = 200;
  int num_aruco = 0;
  int use_stereo = 1;
  int histogram_method = 1;     // TrackBase::HistogramMethod {NONE, HISTOGRAM, CLAHE} (TrackBase.h:78)
  int fast_threshold = 20;
  int grid_x = 5, grid_y = 5;
  int min_px_dist = 10;
  int pyr_levels = 5;           // TrackKLT.h:143
  int win_size = 15;            // TrackKLT.h:144 (only 15 is implemented)
  int lk_max_iters = 30;        // TrackKLT.cpp:670,857
  double lk_eps = 0.01;
  int downsample_cameras = 0;   // VioManagerOptions.h:409: halve every image and mask in front of the tracker (VioManager.cpp:268-278)
};

struct CamCalib {
  size_t cam_id;
  CamModel model;
};

struct Obs {                    // one FeatureDatabase::update_feature call (FeatureDatabase.cpp:59)
  size_t id;
  size_t cam_id;
  float u, v, un, vn;
};

struct Pt { float x, y; };

struct HostImage {              // caller-owned view
  const uint8_t *p;
  size_t stride;
  int w, h;
};

constexpr int PYR_SLOTS = 3;

struct CamState {               // per camera id (TrackBase.h:180-190: img_last, img_mask_last, pts_last, ids_last)
  size_t cam_id = 0;
  int cam_model_idx = -1;       // index into the context's CamModel table
  double max_focal = 0;
  int w = 0, h = 0;             // size of the images this camera tracks on (fixed after the first frame)
  int in_w = 0, in_h = 0;       // size of the images it is fed: (w, h), or twice that with downsample_cameras
  int raw_in[2] = {-1, -1};     // upload buffers of the fed frames: frame number k of this camera lands in raw_in[k & 1], so
                                // that the upload of an announced frame overlaps the pre-processing of the one before it
  int pyr[PYR_SLOTS] = {-1, -1, -1};   // pyramid slots in the context's table; frame number k of this camera lives in slot k % 3
  int cur = 0;                  // slot holding the frame being fed ("curr")
  int last() const { return (cur + PYR_SLOTS - 1) % PYR_SLOTS; }   // slot of the frame before it ("last")
  // frames announced ahead of their feed (submit_batch) are uploaded and pre-processed into the third slot while the feed of
  // the frame before them is still reading the other two
  unsigned n_sub = 0, n_fed = 0;                 // frames uploaded (by submit or by feed itself) / frames fed
  const uint8_t *sub_src[PYR_SLOTS] = {nullptr, nullptr, nullptr};   // image pointer a submitted frame was announced with
  int sub_evt[PYR_SLOTS] = {-1, -1, -1};         // context event that marks its pre-processing done
  unsigned long long sub_seq[PYR_SLOTS] = {0, 0, 0};   // number of the announced batch it went up with
  void *dev_base = nullptr;     // the camera's own device allocation (pyramids, halved frame); freed with its tracker
  uint8_t *raw_ptr[2] = {nullptr, nullptr};   // its upload buffers (arena slots, recycled through the context's free list)
  size_t raw_bytes = 0;
  int raw = -1;                 // downsample_cameras: the halved frame the pre-processing reads (else it reads raw_in[])
  bool have_last = false;
  int dv_n_last = -1;           // device-side pipeline: points after the last feed (-1: pts_last below is current)
  bool no_points() const { return dv_n_last >= 0 ? dv_n_last == 0 : pts_last.empty(); }
  std::vector<Pt> pts_last;
  std::vector<size_t> ids_last;
  std::vector<uint8_t> mask[PYR_SLOTS]; // host copies of the masks, indexed like pyr[]
  bool mask_nonzero[PYR_SLOTS] = {false, false, false};
};

enum Status { OK = 0, E_BADMSG = 1, E_CAMID = 2, E_CUDA = 3, E_ARG = 4, E_SIZE = 5 };

class Context;
struct Unit;                    // per-feed scratch of one stereo pair / mono camera (engine.cpp)

class Rig {                     // one tracker = one TrackKLT instance
 public:
  Rig(Context *ctx, const Params &p, const CamCalib *cams, int n_cams);
  ~Rig();
  Context *ctx;
  Params prm;
  std::vector<CamState> cams;                       // in the order given at creation
  std::unordered_map<size_t, int> cam_index;        // camera id -> index in cams
  size_t currid;                                    // TrackBase.cpp:34
  unsigned long long serial = 0;                    // unique per Rig object ever created (handles may reuse an address)
  int bg_lane = -1;                                 // lane that holds a finished background detection of this tracker
  unsigned long long feeds = 0;                     // feeds this tracker has been through (whatever lane ran them)
  unsigned epoch = 0;                               // bumped by set_num_features / change_feat_id: state a detection that was
                                                    // issued ahead of its feed may have read
  std::vector<Obs> obs;                             // observations produced by the last feed, in update_feature order
  float stage_us[5] = {0, 0, 0, 0, 0};              // pyramid, detection, temporal klt, stereo klt, db update
  int last_error = 0;
  int cuda_error = 0;
  std::mutex mtx;                                   // serialises feed / getters on this handle
  CamState *find_cam(size_t id);
  struct DvRig *dv = nullptr;                       // device-side pipeline: the tracker's state and work areas in HBM (engine.cpp)
  bool dv_dirty = true;                             // pts_last / ids_last / currid changed on the host since the device copy was made
  bool dv_wants_mirror = false;                     // somebody has read pts_last / ids_last: feeds publish them to mapped memory from now on
  bool dv_mirror_published = false;                 // the last feed did
  bool dv_mirror_stale = false;                     // the last feed's pts_last / ids_last are still in mapped memory only (dv_sync_mirror)
  std::vector<std::pair<const Obs *, int>> dv_obs;  // the last feed's observations where the GPU wrote them (mapped memory), in order
  std::vector<Unit *> unit_pool;                    // scratch reused from feed to feed (keeps its vectors' capacity)
  Unit *unit_at(size_t i);
};

struct FeedRequest {            // one feed_new_camera call (utils/sensor_data.h:55-80 CameraData)
  Rig *rig;
  double timestamp;
  int n;
  const size_t *sensor_ids;
  const HostImage *images;
  const HostImage *masks;       // mask[i].p may be null = all zeros
  int status;                   // out
};

// Feeds one frame to each of the given rigs (all on one device) with batched GPU rounds. Returns the worst status.
int feed_batch(Context *ctx, FeedRequest *reqs, int n);
// Announces the next message of each rig ahead of its feed: the frames are uploaded and pre-processed asynchronously and the
// call returns at once.  The rig's next feed must be given the same images (same pointers), valid until that feed returns.
int submit_batch(Context *ctx, FeedRequest *reqs, int n);

Context *context_for_device(int device, int *err);
// device-side pipeline: brings CamState::pts_last / ids_last up to date with what the last feed left in mapped memory
void dv_sync_mirror(Rig &rig);
int context_cuda_error(Context *ctx);
// waits (under the context lock) until no background part of a feed is running: call before mutating tracker state it reads
void context_quiesce(Context *ctx);
void context_stats(Context *ctx, unsigned long long out[12]);
int context_timer(Context *ctx, int stop, float *ms);
void context_host_profile(Context *ctx, int reset, double us[8]);
int context_profile(Context *ctx, int enable, double ms[K_COUNT], unsigned long long launches[K_COUNT]);

// stage entry points (see engine.cpp); slot 0 = frame fed last, 1 = the one before
int stage_get_level(Rig *rig, size_t cam_id, int slot, int level, uint8_t *out, size_t out_stride, int *w, int *h, int *nlv);
int stage_fast(Rig *rig, size_t cam_id, int slot, int x0, int y0, int cw, int ch, int threshold, int *xy, float *resp, int cap,
               int *n_out);
int stage_subpix(Rig *rig, size_t cam_id, int slot, const float *pts, int n, float *out);
int stage_lk(Rig *rig, size_t cam0, int slot0, size_t cam1, int slot1, const float *pts0, const float *pts1_init, int n, float *out,
             uint8_t *status, uint8_t *mask, float *p1n);
int stage_ransac(Rig *rig, size_t cam0, size_t cam1, const float *pts0, const float *pts1, int n, uint8_t *mask, float *p0n,
                 float *p1n);

}  // namespace uvio
