// Dependency-free mirror of the reference types on either side of the tracker, for builds without OpenCV / Eigen / Boost
// (this container has none of them).  Names, members and method signatures follow the reference so that TrackKLTGpu.h
// compiles unchanged against either this header or the real ov_core headers (define UVIO_USE_REAL_OV_CORE).
//
//   ov_core::CameraData        ov_core/src/utils/sensor_data.h:55-80     images are {ptr, step, cols, rows} views here
//   ov_core::Feature           ov_core/src/feat/Feature.h:40-97          Eigen::VectorXf -> std::array<float, 2>
//   ov_core::FeatureDatabase   ov_core/src/feat/FeatureDatabase.{h,cpp}  update_feature :59-85, queries :87-209, cleanup :211-263
//   ov_core::CamBase (calib)   ov_core/src/cam/CamBase.h                 only what the tracker needs: model + 8 values
//   ov_core::TrackBase         ov_core/src/track/TrackBase.h:72-196
#pragma once
#include <algorithm>
#include <array>
#include <atomic>
#include <cstddef>
#include <cstdint>
#include <map>
#include <memory>
#include <mutex>
#include <unordered_map>
#include <vector>

namespace ov_core {

struct MatView {                 // stands in for a CV_8UC1 cv::Mat
  const uint8_t *data = nullptr;
  size_t step = 0;
  int cols = 0, rows = 0;
  bool empty() const { return data == nullptr; }
};

struct CameraData {              // sensor_data.h:55-80
  double timestamp = -1;
  std::vector<int> sensor_ids;
  std::vector<MatView> images;
  std::vector<MatView> masks;
  bool operator<(const CameraData &other) const {
    if (timestamp == other.timestamp) {
      int id = *std::min_element(sensor_ids.begin(), sensor_ids.end());
      int id_other = *std::min_element(other.sensor_ids.begin(), other.sensor_ids.end());
      return id < id_other;
    }
    return timestamp < other.timestamp;
  }
};

struct CamCalibMirror {          // what TrackBase needs from CamBase: model + camera_values (CamBase.h:60-75)
  bool is_fisheye = false;       // CamEqui vs CamRadtan
  double values[8] = {0};        // fx fy cx cy d0 d1 d2 d3
  int w = 0, h = 0;
};

class Feature {                  // Feature.h:40-97
public:
  size_t featid = 0;
  bool to_delete = false;
  std::unordered_map<size_t, std::vector<std::array<float, 2>>> uvs;
  std::unordered_map<size_t, std::vector<std::array<float, 2>>> uvs_norm;
  std::unordered_map<size_t, std::vector<double>> timestamps;
};

class FeatureDatabase {          // FeatureDatabase.h / .cpp
public:
  // FeatureDatabase.cpp:59-85
  void update_feature(size_t id, double timestamp, size_t cam_id, float u, float v, float u_n, float v_n) {
    std::lock_guard<std::mutex> lck(mtx);
    auto it = features_idlookup.find(id);
    if (it != features_idlookup.end()) {
      Feature &feat = *it->second;
      feat.uvs[cam_id].push_back({u, v});
      feat.uvs_norm[cam_id].push_back({u_n, v_n});
      feat.timestamps[cam_id].push_back(timestamp);
      return;
    }
    auto feat = std::make_shared<Feature>();
    feat->featid = id;
    feat->uvs[cam_id].push_back({u, v});
    feat->uvs_norm[cam_id].push_back({u_n, v_n});
    feat->timestamps[cam_id].push_back(timestamp);
    features_idlookup[id] = feat;
  }
  // FeatureDatabase.cpp:36-57
  std::shared_ptr<Feature> get_feature(size_t id, bool remove = false) {
    std::lock_guard<std::mutex> lck(mtx);
    auto it = features_idlookup.find(id);
    if (it == features_idlookup.end()) return nullptr;
    std::shared_ptr<Feature> temp = it->second;
    if (remove) features_idlookup.erase(it);
    return temp;
  }
  // FeatureDatabase.cpp:87-126
  std::vector<std::shared_ptr<Feature>> features_not_containing_newer(double timestamp, bool remove = false, bool skip_deleted = false) {
    std::vector<std::shared_ptr<Feature>> feats_old;
    std::lock_guard<std::mutex> lck(mtx);
    for (auto it = features_idlookup.begin(); it != features_idlookup.end();) {
      if (skip_deleted && it->second->to_delete) { ++it; continue; }
      bool has_newer_measurement = false;
      for (auto const &pair : it->second->timestamps) {
        has_newer_measurement = (!pair.second.empty() && pair.second.back() > timestamp);
        if (has_newer_measurement) break;
      }
      if (!has_newer_measurement) {
        feats_old.push_back(it->second);
        if (remove) { it = features_idlookup.erase(it); continue; }
      }
      ++it;
    }
    return feats_old;
  }
  // FeatureDatabase.cpp:169-209
  std::vector<std::shared_ptr<Feature>> features_containing(double timestamp, bool remove = false, bool skip_deleted = false) {
    std::vector<std::shared_ptr<Feature>> feats_has_timestamp;
    std::lock_guard<std::mutex> lck(mtx);
    for (auto it = features_idlookup.begin(); it != features_idlookup.end();) {
      if (skip_deleted && it->second->to_delete) { ++it; continue; }
      bool has_timestamp = false;
      for (auto const &pair : it->second->timestamps) {
        has_timestamp = (std::find(pair.second.begin(), pair.second.end(), timestamp) != pair.second.end());
        if (has_timestamp) break;
      }
      if (has_timestamp) {
        feats_has_timestamp.push_back(it->second);
        if (remove) { it = features_idlookup.erase(it); continue; }
      }
      ++it;
    }
    return feats_has_timestamp;
  }
  // FeatureDatabase.cpp:211-224
  void cleanup() {
    std::lock_guard<std::mutex> lck(mtx);
    for (auto it = features_idlookup.begin(); it != features_idlookup.end();) {
      if (it->second->to_delete) it = features_idlookup.erase(it);
      else ++it;
    }
  }
  size_t size() {
    std::lock_guard<std::mutex> lck(mtx);
    return features_idlookup.size();
  }
  // FeatureDatabase.h:146-149: returned BY VALUE (which makes TrackBase::change_feat_id half a no-op)
  std::unordered_map<size_t, std::shared_ptr<Feature>> get_internal_data() {
    std::lock_guard<std::mutex> lck(mtx);
    return features_idlookup;
  }

protected:
  std::mutex mtx;
  std::unordered_map<size_t, std::shared_ptr<Feature>> features_idlookup;
};

struct KeyPointMirror {          // cv::KeyPoint: consumers read .pt only (VioManagerHelper.cpp:209-235)
  struct { float x, y; } pt;
  float size = 7.f, angle = -1.f, response = 0.f;
  int octave = 0, class_id = -1;
};

class TrackBase {                // TrackBase.h:72-196 (display_* left out: drawing only)
public:
  enum HistogramMethod { NONE, HISTOGRAM, CLAHE };
  TrackBase(std::unordered_map<size_t, std::shared_ptr<CamCalibMirror>> cameras, int numfeats, int numaruco, bool stereo,
            HistogramMethod histmethod)
      : camera_calib(std::move(cameras)), database(new FeatureDatabase()), num_features(numfeats), use_stereo(stereo),
        histogram_method(histmethod) {
    currid = 4 * (size_t)numaruco + 1;   // TrackBase.cpp:34
  }
  virtual ~TrackBase() {}
  virtual void feed_new_camera(const CameraData &message) = 0;
  std::shared_ptr<FeatureDatabase> get_feature_database() { return database; }
  virtual std::unordered_map<size_t, std::vector<KeyPointMirror>> get_last_obs() = 0;
  virtual std::unordered_map<size_t, std::vector<size_t>> get_last_ids() = 0;
  int get_num_features() { return num_features; }
  virtual void set_num_features(int _num_features) { num_features = _num_features; }
  virtual void change_feat_id(size_t id_old, size_t id_new) = 0;

protected:
  std::unordered_map<size_t, std::shared_ptr<CamCalibMirror>> camera_calib;
  std::shared_ptr<FeatureDatabase> database;
  int num_features;
  bool use_stereo;
  HistogramMethod histogram_method;
  std::atomic<size_t> currid;
};

}  // namespace ov_core
