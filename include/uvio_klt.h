/* uvio_klt.h — C ABI of the B200 (sm_100a) TrackKLT.
 *
 * Drop-in boundary for the reference's visual front-end hot path.  The reference has no FFI: its "plugin API" is the
 * C++ virtual class ov_core::TrackBase (ov_core/src/track/TrackBase.h:72-196) held as
 * std::shared_ptr<TrackBase> trackFEATS (ov_msckf/src/core/VioManager.h:189) and constructed as a TrackKLT at
 * ov_msckf/src/core/VioManager.cpp:133-135.  Every entry point below names the reference member it stands in for; the
 * header-only shim  uvio_b200/cpp/TrackKLTGpu.h  turns them back into a TrackBase subclass.
 *
 * Conventions: plain pointers and sizes only; no C++/OpenCV/torch types; nothing throws; every function returns an
 * UVIO_* status (0 = ok).  Host pointers passed to feed() are not retained after the call returns.  There is no CPU
 * fallback: create() fails with UVIO_E_CUDA when no CUDA device is present.
 */
#ifndef UVIO_KLT_H_
#define UVIO_KLT_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define UVIO_OK 0
#define UVIO_E_BADMSG 1 /* malformed message: the reference calls std::exit(EXIT_FAILURE) (TrackKLT.cpp:37-43, :90-93) */
#define UVIO_E_CAMID 2  /* unknown camera id: the reference's camera_calib.at() throws (TrackKLT.cpp:53-54) */
#define UVIO_E_CUDA 3   /* CUDA runtime error (sticky; see uvio_klt_cuda_error) or no CUDA device */
#define UVIO_E_ARG 4    /* invalid argument / unsupported parameter value */
#define UVIO_E_SIZE 5   /* image size differs from the first frame of that camera */

/* TrackBase::HistogramMethod (TrackBase.h:78) */
#define UVIO_HIST_NONE 0
#define UVIO_HIST_HISTOGRAM 1
#define UVIO_HIST_CLAHE 2
/* camera models: cam/CamRadtan.h, cam/CamEqui.h */
#define UVIO_CAM_RADTAN 0
#define UVIO_CAM_EQUI 1

typedef struct uvio_klt uvio_klt; /* one tracker = one ov_core::TrackKLT instance (one rig) */

/* Arguments of TrackKLT::TrackKLT (TrackKLT.h:54-57) plus the constants the reference hard-codes. */
typedef struct UvioKltParams {
  int num_features;     /* numfeats: features per camera */
  int num_aruco;        /* numaruco: first feature id is 4*num_aruco + 2 (TrackBase.cpp:34) */
  int use_stereo;       /* stereo */
  int histogram_method; /* histmethod: UVIO_HIST_* */
  int fast_threshold;   /* fast_threshold */
  int grid_x, grid_y;   /* gridx, gridy */
  int min_px_dist;      /* minpxdist */
  int pyr_levels;       /* TrackKLT.h:143 hard-codes 5 */
  int win_size;         /* TrackKLT.h:144 hard-codes 15; only 15 is supported */
  int lk_max_iters;     /* TrackKLT.cpp:670,857: 30; only 30 is supported (UVIO_E_ARG otherwise) */
  double lk_eps;        /* TrackKLT.cpp:670,857: 0.01; only 0.01 is supported */
  /* VioManagerOptions.h:409 downsample_cameras.  Non-zero: every fed image and mask is halved on the device with
   * cv::pyrDown(src, dst, Size(cols / 2.0, rows / 2.0)) before the tracker sees it, which is what
   * VioManager::track_image_and_update does in front of trackFEATS->feed_new_camera (ov_msckf/src/core/VioManager.cpp:268-278);
   * the caller then feeds the full-size message and passes the already halved intrinsics (VioManagerOptions.h:251-260). */
  int downsample_cameras;
} UvioKltParams;

/* One entry of the camera_calib map given to the TrackBase ctor (TrackBase.h:96, cam/CamBase.h). */
typedef struct UvioCamCalib {
  size_t cam_id;
  int model;            /* UVIO_CAM_* */
  double intrinsics[4]; /* fx fy cx cy */
  double distortion[4]; /* k1 k2 p1 p2 (radtan) | k1 k2 k3 k4 (equidistant) */
} UvioCamCalib;

/* One FeatureDatabase::update_feature(id, timestamp, cam_id, u, v, u_n, v_n) call (feat/FeatureDatabase.cpp:59). */
typedef struct UvioObs {
  size_t id;
  size_t cam_id;
  float u, v, un, vn;
} UvioObs;

/* One ov_core::CameraData message (utils/sensor_data.h:55-80). mask[i] may be NULL (= all zeros, nothing masked). */
typedef struct UvioFrame {
  double timestamp;
  int n_img;
  const size_t *sensor_ids;
  const uint8_t *const *img; /* CV_8UC1 */
  const size_t *img_stride;  /* bytes per row (cv::Mat::step) */
  const uint8_t *const *mask;
  const size_t *mask_stride;
  const int *width;
  const int *height;
  int status; /* out: UVIO_* for this message (feed_batch) */
} UvioFrame;

void uvio_klt_default_params(UvioKltParams *p);

/* TrackKLT::TrackKLT (TrackKLT.h:54-57) / TrackBase::TrackBase (TrackBase.cpp:30-41) */
int uvio_klt_create(const UvioKltParams *params, const UvioCamCalib *cams, int n_cams, int cuda_device, uvio_klt **out);
/* TrackBase::~TrackBase (TrackBase.h:100) */
void uvio_klt_destroy(uvio_klt *h);

/* TrackKLT::feed_new_camera (TrackKLT.cpp:34-94). Synchronous: on return the observations of this frame are available
 * through uvio_klt_get_new_observations, as the caller of the reference expects of the FeatureDatabase
 * (VioManager.cpp:281 then :369-375).
 * Two documented differences from the reference, both inside cv::findFundamentalMat (TrackKLT.cpp:873):
 *  - every correspondence of a camera identical (the same frame delivered twice, two blank frames): OpenCV asserts
 *    (matrix.cpp:764) and the reference, which does not catch it, terminates; this tracker keeps the tracked points (the
 *    zero-motion model fits all of them) and returns UVIO_OK;
 *  - a camera holding 10..13 points: OpenCV silently switches to LMedS, whose winner is decided by rounding noise (cv2's own
 *    mask changes under a one-ulp change of one input in 60-80 % of such cases); the mask returned here is a valid LMedS
 *    outcome (tests/test_lmeds.py) but not necessarily the one a given OpenCV build draws.  DESIGN.md section 5. */
int uvio_klt_feed(uvio_klt *h, const UvioFrame *frame);
/* The same for n independent trackers living on one device, with the GPU work of all of them batched (one launch set
 * per stage for the whole batch).  The reference's equivalent is n TrackKLT instances fed from n threads. */
int uvio_klt_feed_batch(uvio_klt *const *h, UvioFrame *frames, int n);

/* Optional: announce the NEXT message of a tracker ahead of its feed -- what a driver does in the image-arrival callback
 * (ov_msckf/src/ros/ROS1Visualizer.cpp:571-580 builds the CameraData there) while the estimator is still busy with the
 * previous frame.  The images are uploaded, (halved,) equalised and their pyramids built asynchronously on a stream of
 * their own; the call returns as soon as the work is queued.  Contract: the next uvio_klt_feed / uvio_klt_feed_batch
 * of that tracker must be given the same message (same image pointers), and the image memory must stay valid and
 * unchanged until that feed returns; masks are read by the feed, not here.  At most two messages per camera may be
 * announced and not yet fed (UVIO_E_ARG otherwise).  Results are identical to a feed that was not announced. */
int uvio_klt_submit(uvio_klt *h, const UvioFrame *frame);
int uvio_klt_submit_batch(uvio_klt *const *h, UvioFrame *frames, int n);

/* The update_feature calls the last feed made, in the reference's order (all left, then all right; TrackKLT.cpp:357-366,
 * mono :176-179). *n receives the count; at most cap entries are written. */
int uvio_klt_get_new_observations(uvio_klt *h, UvioObs *out, int cap, int *n);
/* TrackBase::get_last_obs / get_last_ids for one camera (TrackBase.h:134-146). uv = n x 2 floats. */
int uvio_klt_get_last(uvio_klt *h, size_t cam_id, size_t *ids, float *uv, int cap, int *n);
/* TrackBase::set_num_features (TrackBase.h:152) */
int uvio_klt_set_num_features(uvio_klt *h, int n);
/* TrackBase::change_feat_id (TrackBase.cpp:230-248): the ids_last rewrite (the database side stays in the host shim) */
int uvio_klt_change_feat_id(uvio_klt *h, size_t id_old, size_t id_new);
/* the reference's [TIME-KLT] stage stamps (TrackKLT.cpp:385-392): pyramid, detection, temporal klt, stereo klt, db update */
int uvio_klt_get_timing(uvio_klt *h, float us[5]);
int uvio_klt_last_error(uvio_klt *h);
int uvio_klt_cuda_error(uvio_klt *h); /* last cudaError_t seen on this tracker's device context */
const char *uvio_klt_version(void);

/* ---- measurement helpers (no reference counterpart; the reference stamps wall-clock rT1..rT7, TrackBase.h:195) ------
 * Counters of the device context since its creation: [0] frames fed, [1] kernels launched, [2] image bytes uploaded,
 * [3] bytes the GPU wrote to host memory (corner lists, refined / tracked points, status, masks), [4] feed_batch calls,
 * [5] pixels scanned by FAST, [6] LK point-levels (templates staged), [7] LK iterations, [8] LK iterations whose mismatch
 * sums needed the ordered fp32 chains (the others were exact integer reductions), [9..11] reserved (0). */
int uvio_klt_device_stats(int cuda_device, unsigned long long out[12]);
/* CUDA-event stopwatch recorded on the stream all kernels of that device's trackers are launched on. */
int uvio_klt_timer_start(int cuda_device);
int uvio_klt_timer_stop(int cuda_device, float *ms);
/* Per-kernel CUDA-event timing on that stream. enable: 1 = reset and start (all kernels), 1 | (mask << 1) = reset and start
 * for the kernels whose UVIO_K_* bit is set in mask only (two event records per timed launch, none for the others),
 * 0 = stop, -1 = leave as is. ms / launches (16 entries each, may be NULL) receive the sums accumulated so far, indexed by
 * UVIO_K_*. */
#define UVIO_K_HIST 0
#define UVIO_K_LUT_APPLY 1
#define UVIO_K_CLAHE_TILES 2
#define UVIO_K_CLAHE_APPLY 3
#define UVIO_K_PYRDOWN 4
#define UVIO_K_FAST 5
#define UVIO_K_SUBPIX 6
#define UVIO_K_LK 7
#define UVIO_K_RANSAC 8
#define UVIO_K_PREPROCESS 9 /* equalizeHist + pyramid, one cluster launch (ids 0, 1, 4 are the retired separate passes) */
int uvio_klt_profile(int cuda_device, int enable, double ms[16], unsigned long long launches[16]);
/* Host wall-clock of the feed, summed over the lanes of that device's context, in microseconds: [0] upload + round-1 launch
 * issue, [1] wait for round 1, [2] host selection + round-2 issue, [3] wait for round 2, [4] host detection + round-3 issue,
 * [5] wait for round 3, [6] host merge + observations, [7] lane passes counted. reset != 0 clears the sums afterwards. */
int uvio_klt_host_profile(int cuda_device, int reset, double us[8]);
/* uvio_klt_get_new_observations for n trackers at once: tracker i writes at out + i*cap_each, counts[i] = its count */
int uvio_klt_get_new_observations_batch(uvio_klt *const *h, int n, UvioObs *out, int cap_each, int *counts);

/* ---- stage entry points: one kernel on the pyramids the tracker already holds (parity tests, micro-benchmarks) ------
 * slot 0 = the frame fed last (img_pyramid_last after the feed returns), slot 1 = the frame before it. */
/* level `level` of cv::buildOpticalFlowPyramid's output (after equalisation, TrackKLT.cpp:58-71); out may be NULL */
int uvio_klt_stage_get_level(uvio_klt *h, size_t cam_id, int slot, int level, uint8_t *out, size_t out_stride, int *w, int *hgt,
                             int *n_levels);
/* cv::FAST(img(Rect(x0,y0,cw,ch)), kps, threshold, true) (Grider_GRID.h:125): row-major keypoints, cell-relative */
int uvio_klt_stage_fast(uvio_klt *h, size_t cam_id, int slot, int x0, int y0, int cw, int ch, int threshold, int *xy, float *resp,
                        int cap, int *n);
/* cv::cornerSubPix(img, pts, (5,5), (-1,-1), {COUNT+EPS,20,0.001}) (Grider_GRID.h:163-174) */
int uvio_klt_stage_subpix(uvio_klt *h, size_t cam_id, int slot, const float *pts, int n, float *out);
/* cv::calcOpticalFlowPyrLK(..., OPTFLOW_USE_INITIAL_FLOW) (TrackKLT.cpp:671,858); if mask != NULL also the undistort +
 * cv::findFundamentalMat of TrackKLT::perform_matching (:860-879): mask = status && inlier, p1n = undistorted result */
int uvio_klt_stage_lk(uvio_klt *h, size_t cam0, int slot0, size_t cam1, int slot1, const float *pts0, const float *pts1_init, int n,
                      float *out, uint8_t *status, uint8_t *mask, float *p1n);
/* CamBase::undistort_cv on both point sets + cv::findFundamentalMat(FM_RANSAC, 2/max_focal, 0.999) (TrackKLT.cpp:860-873) */
int uvio_klt_stage_ransac(uvio_klt *h, size_t cam0, size_t cam1, const float *pts0, const float *pts1, int n, uint8_t *mask,
                          float *p0n, float *p1n);

/* ---- allocation-free entry point --------------------------------------------------------------------------------------
 * No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (code under CUDA-graph capture, tight serving loops, harnesses that forbid it).
 *
 * CLAHE exactly as the tracker applies it (cv::createCLAHE(10.0, Size(8,8))->apply; the arithmetic of k_clahe_tiles /
 * k_clahe_apply, bit-identical to level 0 of uvio_klt_stage_get_level for a UVIO_HIST_CLAHE tracker) of n_images CV_8UC1
 * images that already live in DEVICE memory: image i is read at src_dev + i*src_image_stride (rows src_pitch bytes apart)
 * and written to dst_dev + i*dst_image_stride (rows dst_pitch apart); src and dst must not overlap.  The two kernels are
 * queued on cuda_stream (a cudaStream_t, NULL = the default stream) of the CURRENT device and the call returns without
 * synchronising.  It calls no cudaMalloc / cudaFree / allocating routine: the per-tile LUT scratch (16 KiB per image)
 * is a file-scope __device__ array of UVIO_CLAHE_NOALLOC_MAX_IMAGES * 16 KiB = 512 KiB that every GPU loading the library
 * holds until exit.  Limits: n_images <= UVIO_CLAHE_NOALLOC_MAX_IMAGES, 1 <= width, height <= UVIO_CLAHE_NOALLOC_MAX_DIM;
 * anything beyond is refused with UVIO_E_ARG and a message in err (err_cap bytes, may be NULL).  The first call on a
 * device resolves the array's device address (cudaGetSymbolAddress): make that call outside graph capture.  All calls on
 * a GPU share the one array, which keeps the last call's LUTs (every call rewrites all it reads, nothing needs clearing):
 * two calls whose kernels may overlap on the GPU -- different streams without an ordering event -- race; order them. */
#define UVIO_CLAHE_NOALLOC_MAX_IMAGES 32
#define UVIO_CLAHE_NOALLOC_MAX_DIM 8192
int uvio_klt_clahe_noalloc(const uint8_t *src_dev, size_t src_pitch, size_t src_image_stride, uint8_t *dst_dev, size_t dst_pitch,
                           size_t dst_image_stride, int width, int height, int n_images, void *cuda_stream, char *err,
                           size_t err_cap);

#ifdef __cplusplus
}
#endif
#endif /* UVIO_KLT_H_ */
