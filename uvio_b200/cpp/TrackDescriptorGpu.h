// TrackDescriptorGpu -- the reference-facing side of the descriptor tracker: a TrackBase subclass that forwards to the C ABI
// (include/uvio_desc.h) and replays the returned observations into the tracker-owned FeatureDatabase in the reference's order
// (TrackDescriptor.cpp:311-326: left and right of each feature in turn), so every consumer of ov_core::TrackBase keeps working:
//
//     // ov_msckf/src/core/VioManager.cpp:137-140
//     trackFEATS = std::shared_ptr<TrackBase>(new TrackDescriptorGpu(state->_cam_intrinsics_cameras, init_max_features,
//                      state->_options.max_aruco_features, params.use_stereo, params.histogram_method,
//                      params.fast_threshold, params.grid_x, params.grid_y, params.min_px_dist, params.knn_ratio));
//
// Same rules as TrackKLTGpu.h: only the reference's virtual members are overridden; the plain members the estimator calls
// through the base pointer (set_num_features, get_last_obs / get_last_ids, change_feat_id; TrackBase.h:134-152) work because
// the base's own state is kept in step with the handle around every feed.
#pragma once
#include <cstdio>
#include <cstdlib>
#include <mutex>
#include <stdexcept>
#include <unordered_map>
#include <vector>

#include "../../include/uvio_desc.h"

#ifdef UVIO_USE_REAL_OV_CORE
#include "cam/CamBase.h"
#include "cam/CamEqui.h"
#include "feat/FeatureDatabase.h"
#include "track/TrackBase.h"
#else
#include "ov_core_mirror.h"
#endif

namespace ov_core {

class TrackDescriptorGpu : public TrackBase {
public:
#ifdef UVIO_USE_REAL_OV_CORE
  typedef std::unordered_map<size_t, std::shared_ptr<CamBase>> CamMap;
  typedef cv::KeyPoint KeyPointT;
#else
  typedef std::unordered_map<size_t, std::shared_ptr<CamCalibMirror>> CamMap;
  typedef KeyPointMirror KeyPointT;
#endif

  // same argument list as TrackDescriptor::TrackDescriptor (TrackDescriptor.h:55-59); cuda_device is the only addition
  explicit TrackDescriptorGpu(CamMap cameras, int numfeats, int numaruco, bool stereo, HistogramMethod histmethod, int fast_threshold,
                              int gridx, int gridy, int minpxdist, double knnratio, int cuda_device = 0)
      : TrackBase(cameras, numfeats, numaruco, stereo, histmethod) {
    UvioDescParams p;
    uvio_desc_default_params(&p);
    p.num_features = numfeats;
    p.num_aruco = numaruco;
    p.use_stereo = stereo ? 1 : 0;
    p.histogram_method = (int)histmethod;          // NONE, HISTOGRAM, CLAHE = 0, 1, 2 on both sides
    p.fast_threshold = fast_threshold;
    p.grid_x = gridx;
    p.grid_y = gridy;
    p.min_px_dist = minpxdist;
    p.knn_ratio = knnratio;
    std::vector<UvioCamCalib> cc;
    for (auto const &kv : cameras) {
      UvioCamCalib c;
      c.cam_id = kv.first;
#ifdef UVIO_USE_REAL_OV_CORE
      c.model = std::dynamic_pointer_cast<CamEqui>(kv.second) ? UVIO_CAM_EQUI : UVIO_CAM_RADTAN;
      Eigen::MatrixXd v = kv.second->get_value();
      for (int i = 0; i < 4; i++) { c.intrinsics[i] = v(i); c.distortion[i] = v(4 + i); }
#else
      c.model = kv.second->is_fisheye ? UVIO_CAM_EQUI : UVIO_CAM_RADTAN;
      for (int i = 0; i < 4; i++) { c.intrinsics[i] = kv.second->values[i]; c.distortion[i] = kv.second->values[4 + i]; }
#endif
      cc.push_back(c);
    }
    int rc = uvio_desc_create(&p, cc.data(), (int)cc.size(), cuda_device, &handle);
    if (rc != UVIO_OK) throw std::runtime_error("TrackDescriptorGpu: uvio_desc_create failed (no CUDA device? there is no CPU fallback)");
    pushed_num_features = numfeats;
  }
  ~TrackDescriptorGpu() override { uvio_desc_destroy(handle); }

  // TrackDescriptor::feed_new_camera (TrackDescriptor.cpp:33-60)
  void feed_new_camera(const CameraData &message) override {
    const size_t n = message.images.size();
    if (message.sensor_ids.empty() || message.sensor_ids.size() != n || n != message.masks.size()) {
      std::fprintf(stderr, "[ERROR]: MESSAGE DATA SIZES DO NOT MATCH OR EMPTY!!!\n");   // TrackDescriptor.cpp:36-42
      std::exit(EXIT_FAILURE);
    }
    std::vector<size_t> ids(n), istep(n), mstep(n);
    std::vector<const uint8_t *> img(n), msk(n);
    std::vector<int> w(n), h(n);
    for (size_t i = 0; i < n; i++) {
      ids[i] = (size_t)message.sensor_ids[i];
      img[i] = message.images[i].data;
      istep[i] = (size_t)message.images[i].step;
      msk[i] = message.masks[i].data;
      mstep[i] = (size_t)message.masks[i].step;
      w[i] = message.images[i].cols;
      h[i] = message.images[i].rows;
    }
    UvioFrame f;
    f.timestamp = message.timestamp;
    f.n_img = (int)n;
    f.sensor_ids = ids.data();
    f.img = img.data();
    f.img_stride = istep.data();
    f.mask = msk.data();
    f.mask_stride = mstep.data();
    f.width = w.data();
    f.height = h.data();
    f.status = 0;
    sync_to_handle();
    int rc = uvio_desc_feed(handle, &f);
    if (rc == UVIO_E_BADMSG) {                                                        // TrackDescriptor.cpp:57-58
      std::fprintf(stderr, "[ERROR]: invalid number of images passed %zu, we only support mono or stereo tracking\n", n);
      std::exit(EXIT_FAILURE);
    }
    if (rc == UVIO_E_CAMID) throw std::out_of_range("TrackDescriptorGpu: unknown camera id");   // .at() in the reference (:67, :182-183)
    if (rc != UVIO_OK) throw std::runtime_error("TrackDescriptorGpu: uvio_desc_feed failed");
    replay_observations(message.timestamp);
    sync_from_handle();
  }

  // get_last_obs / get_last_ids / set_num_features / change_feat_id are NOT virtual in the reference (TrackBase.h:134-152) and
  // the estimator calls them through std::shared_ptr<TrackBase> (VioManagerHelper.cpp:124, :209-210), so they cannot be
  // overridden: the shim keeps the BASE's own state in step with the handle instead (sync_to_handle / sync_from_handle).
  uvio_desc *native_handle() { return handle; }

protected:
  // base -> handle: what the non-virtual setters changed on the base since the last feed
  void sync_to_handle() {
    if (num_features != pushed_num_features) {                       // TrackBase::set_num_features (TrackBase.h:152)
      uvio_desc_set_num_features(handle, num_features);
      pushed_num_features = num_features;
    }
    // TrackBase::change_feat_id rewrote entries of ids_last in place (TrackBase.cpp:241-247): forward every (old -> new)
    std::lock_guard<std::mutex> lckv(mtx_last_vars);
    for (auto const &kv : ids_synced) {
      auto it = ids_last.find(kv.first);
      if (it == ids_last.end() || it->second.size() != kv.second.size()) continue;
      for (size_t i = 0; i < kv.second.size(); i++)
        if (it->second[i] != kv.second[i]) uvio_desc_change_feat_id(handle, kv.second[i], it->second[i]);
    }
  }
  // handle -> base: pts_last / ids_last as TrackDescriptor leaves them (TrackDescriptor.cpp:333-345), under the same mutex
  void sync_from_handle() {
    std::vector<size_t> ids;
    std::vector<float> uv;
    std::lock_guard<std::mutex> lckv(mtx_last_vars);
    for (auto const &kv : camera_calib) {
      last_of(kv.first, ids, uv);
      std::vector<KeyPointT> &kps = pts_last[kv.first];
      kps.assign(ids.size(), KeyPointT());
      for (size_t i = 0; i < ids.size(); i++) { kps[i].pt.x = uv[2 * i]; kps[i].pt.y = uv[2 * i + 1]; }
      ids_last[kv.first] = ids;
      ids_synced[kv.first] = ids;
    }
  }
  void replay_observations(double timestamp) {
    int n = 0;
    uvio_desc_get_new_observations(handle, nullptr, 0, &n);
    obs.resize((size_t)n);
    if (n) uvio_desc_get_new_observations(handle, obs.data(), n, &n);
    for (int i = 0; i < n; i++)
      database->update_feature(obs[i].id, timestamp, obs[i].cam_id, obs[i].u, obs[i].v, obs[i].un, obs[i].vn);
  }
  void last_of(size_t cam, std::vector<size_t> &ids, std::vector<float> &uv) {
    int n = 0;
    uvio_desc_get_last(handle, cam, nullptr, nullptr, 0, &n);
    ids.resize((size_t)n);
    uv.resize((size_t)2 * n);
    if (n) uvio_desc_get_last(handle, cam, ids.data(), uv.data(), n, &n);
  }
  uvio_desc *handle = nullptr;
  std::vector<UvioObs> obs;
  int pushed_num_features = 0;                                        // the feature budget the handle was last given
  std::unordered_map<size_t, std::vector<size_t>> ids_synced;         // ids_last as the handle had them after the last feed
};

}  // namespace ov_core
