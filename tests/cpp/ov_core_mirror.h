// TEST SCAFFOLDING (not part of the product): dependency-free mirror of the reference types on either side of the tracker,
// for builds without OpenCV / Eigen / Boost (this container has none of them).  Names, members, method signatures and --
// what matters for a drop-in -- the virtual / non-virtual split of TrackBase follow the reference, so that TrackKLTGpu.h
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
#include <string>
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
  // Feature.cpp:60-82: drop every measurement with a timestamp <= the given one
  void clean_older_measurements(double timestamp) {
    for (auto &pair : timestamps) {
      auto &ts = pair.second;
      auto &a = uvs[pair.first];
      auto &b = uvs_norm[pair.first];
      size_t k = 0;
      for (size_t i = 0; i < ts.size(); i++)
        if (!(ts[i] <= timestamp)) { ts[k] = ts[i]; a[k] = a[i]; b[k] = b[i]; k++; }
      ts.resize(k); a.resize(k); b.resize(k);
    }
  }
  // Feature.cpp:36-58 / :84-106: keep only (valid) / drop (invalid) the measurements at the listed timestamps
  void clean_old_measurements(const std::vector<double> &valid_times) {
    for (auto &pair : timestamps) {
      auto &ts = pair.second;
      auto &a = uvs[pair.first];
      auto &b = uvs_norm[pair.first];
      size_t k = 0;
      for (size_t i = 0; i < ts.size(); i++)
        if (std::find(valid_times.begin(), valid_times.end(), ts[i]) != valid_times.end()) { ts[k] = ts[i]; a[k] = a[i]; b[k] = b[i]; k++; }
      ts.resize(k); a.resize(k); b.resize(k);
    }
  }
  void clean_invalid_measurements(const std::vector<double> &invalid_times) {
    for (auto &pair : timestamps) {
      auto &ts = pair.second;
      auto &a = uvs[pair.first];
      auto &b = uvs_norm[pair.first];
      size_t k = 0;
      for (size_t i = 0; i < ts.size(); i++)
        if (std::find(invalid_times.begin(), invalid_times.end(), ts[i]) == invalid_times.end()) { ts[k] = ts[i]; a[k] = a[i]; b[k] = b[i]; k++; }
      ts.resize(k); a.resize(k); b.resize(k);
    }
  }
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
        has_newer_measurement = (!pair.second.empty() && pair.second.back() >= timestamp);   // FeatureDatabase.cpp:104 (>=)
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
  // FeatureDatabase.cpp:226-243
  void cleanup_measurements(double timestamp) {
    std::lock_guard<std::mutex> lck(mtx);
    for (auto it = features_idlookup.begin(); it != features_idlookup.end();) {
      it->second->clean_older_measurements(timestamp);
      int ct_meas = 0;
      for (const auto &pair : it->second->timestamps) ct_meas += (int)pair.second.size();
      if (ct_meas < 1) it = features_idlookup.erase(it);
      else ++it;
    }
  }
  // FeatureDatabase.cpp:128-167
  std::vector<std::shared_ptr<Feature>> features_containing_older(double timestamp, bool remove = false, bool skip_deleted = false) {
    std::vector<std::shared_ptr<Feature>> feats_old;
    std::lock_guard<std::mutex> lck(mtx);
    for (auto it = features_idlookup.begin(); it != features_idlookup.end();) {
      if (skip_deleted && it->second->to_delete) { ++it; continue; }
      bool found = false;
      for (auto const &pair : it->second->timestamps) {
        found = (!pair.second.empty() && pair.second.at(0) < timestamp);
        if (found) break;
      }
      if (found) {
        feats_old.push_back(it->second);
        if (remove) { it = features_idlookup.erase(it); continue; }
      }
      ++it;
    }
    return feats_old;
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

class TrackBase {                // TrackBase.h:72-196, with the reference's virtual / non-virtual split kept exactly:
public:                          // virtual: ~TrackBase (:91), feed_new_camera (:97), display_active (:106), display_history (:116);
                                 // plain members: get_feature_database (:123), change_feat_id (:134), get_last_obs (:137),
                                 // get_last_ids (:143), get_num_features (:149), set_num_features (:152)
  enum HistogramMethod { NONE, HISTOGRAM, CLAHE };
  TrackBase(std::unordered_map<size_t, std::shared_ptr<CamCalibMirror>> cameras, int numfeats, int numaruco, bool stereo,
            HistogramMethod histmethod)
      : camera_calib(std::move(cameras)), database(new FeatureDatabase()), num_features(numfeats), use_stereo(stereo),
        histogram_method(histmethod) {
    currid = 4 * (size_t)numaruco + 1;   // TrackBase.cpp:34
    if (mtx_feeds.empty() || mtx_feeds.size() != camera_calib.size()) {   // TrackBase.cpp:37-40
      std::vector<std::mutex> list(camera_calib.size());
      mtx_feeds.swap(list);
    }
  }
  virtual ~TrackBase() {}
  virtual void feed_new_camera(const CameraData &message) = 0;
  virtual void display_active(MatView &, int, int, int, int, int, int, std::string = "") {}                               // drawing only
  virtual void display_history(MatView &, int, int, int, int, int, int, std::vector<size_t> = {}, std::string = "") {}   // drawing only
  std::shared_ptr<FeatureDatabase> get_feature_database() { return database; }
  // TrackBase.cpp:230-248, statement for statement (get_internal_data() returns the map BY VALUE, so the erase / insert act
  // on temporaries and only feat->featid changes in the database; ids_last is rewritten without taking mtx_last_vars)
  void change_feat_id(size_t id_old, size_t id_new) {
    if (database->get_internal_data().find(id_old) != database->get_internal_data().end()) {
      std::shared_ptr<Feature> feat = database->get_internal_data().at(id_old);
      database->get_internal_data().erase(id_old);
      feat->featid = id_new;
      database->get_internal_data().insert({id_new, feat});
    }
    for (auto &cam_ids_pair : ids_last)
      for (size_t i = 0; i < cam_ids_pair.second.size(); i++)
        if (cam_ids_pair.second.at(i) == id_old) ids_last.at(cam_ids_pair.first).at(i) = id_new;
  }
  std::unordered_map<size_t, std::vector<KeyPointMirror>> get_last_obs() {
    std::lock_guard<std::mutex> lckv(mtx_last_vars);
    return pts_last;
  }
  std::unordered_map<size_t, std::vector<size_t>> get_last_ids() {
    std::lock_guard<std::mutex> lckv(mtx_last_vars);
    return ids_last;
  }
  int get_num_features() { return num_features; }
  void set_num_features(int _num_features) { num_features = _num_features; }

protected:
  std::unordered_map<size_t, std::shared_ptr<CamCalibMirror>> camera_calib;
  std::shared_ptr<FeatureDatabase> database;
  std::map<size_t, bool> camera_fisheye;
  int num_features;
  bool use_stereo;
  HistogramMethod histogram_method;
  std::vector<std::mutex> mtx_feeds;
  std::mutex mtx_last_vars;
  std::map<size_t, MatView> img_last;
  std::map<size_t, MatView> img_mask_last;
  std::unordered_map<size_t, std::vector<KeyPointMirror>> pts_last;
  std::unordered_map<size_t, std::vector<size_t>> ids_last;
  std::atomic<size_t> currid;
};

}  // namespace ov_core
