// Drives the GPU tracker exactly the way ov_core/src/test_tracking.cpp:288-381 drives TrackKLT: through TrackBase, with
// CameraData messages, reading the FeatureDatabase afterwards.  Reads raw stereo frames from a file written by the
// pytest wrapper and prints one line per update_feature call so the wrapper can compare them with the golden stream.
//   usage: test_trackbase_shim frames.bin W H N [base]
// With "base" the run follows the estimator's use of the NON-virtual TrackBase members through std::shared_ptr<TrackBase>:
// the tracker starts with 25 features per camera (init_max_features / num_cameras, VioManager.cpp:131-135), set_num_features
// raises the budget after a few frames (VioManagerHelper.cpp:124), get_last_obs() / get_last_ids() are read every frame
// (VioManagerHelper.cpp:209-210, :235-237) and one id is renamed with change_feat_id (TrackBase.cpp:230-248).
#include <cstdio>
#include <cstdlib>
#include <memory>
#include <string>
#include <vector>

#include "../../uvio_b200/cpp/TrackKLTGpu.h"

using namespace ov_core;

int main(int argc, char **argv) {
  if (argc < 5) return 2;
  const int W = std::atoi(argv[2]), H = std::atoi(argv[3]), N = std::atoi(argv[4]);
  FILE *f = std::fopen(argv[1], "rb");
  if (!f) return 2;
  std::vector<uint8_t> buf((size_t)N * 2 * W * H), zero((size_t)W * H, 0);
  if (std::fread(buf.data(), 1, buf.size(), f) != buf.size()) return 2;
  std::fclose(f);
  // euroc_mav calibration (config/euroc_mav/kalibr_imucam_chain.yaml:11-14,24-27)
  TrackKLTGpu::CamMap cams;
  const double c0[8] = {458.654, 457.296, 367.215, 248.375, -0.28340811, 0.07395907, 0.00019359, 1.76187114e-05};
  const double c1[8] = {457.587, 456.134, 379.999, 255.238, -0.28368365, 0.07451284, -0.00010473, -3.55590700e-05};
  for (int c = 0; c < 2; c++) {
    auto cam = std::make_shared<CamCalibMirror>();
    for (int i = 0; i < 8; i++) cam->values[i] = c ? c1[i] : c0[i];
    cams[(size_t)c] = cam;
  }
  const bool base_mode = argc > 5 && std::string(argv[5]) == "base";
  std::shared_ptr<TrackBase> trackFEATS(new TrackKLTGpu(cams, base_mode ? 25 : 200, 0, true, TrackBase::HISTOGRAM, 20, 5, 5, 10));
  if (base_mode) {
    for (int k = 0; k < N; k++) {
      CameraData message;
      message.timestamp = 0.05 * k;
      for (int c = 0; c < 2; c++) {
        message.sensor_ids.push_back(c);
        message.images.push_back(MatView{buf.data() + ((size_t)k * 2 + c) * W * H, (size_t)W, W, H});
        message.masks.push_back(MatView{zero.data(), (size_t)W, W, H});
      }
      if (k == 6) trackFEATS->set_num_features(150);                  // base member, not virtual
      if (k == 9) {                                                   // rename the first left id, as a loop closure would
        auto ids = trackFEATS->get_last_ids();
        if (!ids.at(0).empty()) trackFEATS->change_feat_id(ids.at(0).front(), ids.at(0).front() + 100000);
      }
      trackFEATS->feed_new_camera(message);
      std::unordered_map<size_t, std::vector<KeyPointMirror>> last_obs = trackFEATS->get_last_obs();
      std::unordered_map<size_t, std::vector<size_t>> last_ids = trackFEATS->get_last_ids();
      std::printf("nf %d %d\n", k, trackFEATS->get_num_features());
      for (size_t c = 0; c < 2; c++) {
        const std::vector<size_t> &ids = last_ids.at(c);              // .at(): throws if the base's maps were never filled
        const std::vector<KeyPointMirror> &kp = last_obs.at(c);
        if (ids.size() != kp.size()) return 3;
        for (size_t i = 0; i < ids.size(); i++) std::printf("last %d %zu %zu %.9g %.9g\n", k, c, ids[i], kp[i].pt.x, kp[i].pt.y);
      }
    }
    return 0;
  }
  size_t prev = 0;
  for (int k = 0; k < N; k++) {
    CameraData message;
    message.timestamp = 0.05 * k;
    for (int c = 0; c < 2; c++) {
      message.sensor_ids.push_back(c);
      message.images.push_back(MatView{buf.data() + ((size_t)k * 2 + c) * W * H, (size_t)W, W, H});
      message.masks.push_back(MatView{zero.data(), (size_t)W, W, H});
    }
    trackFEATS->feed_new_camera(message);
    // what VioManager does next (VioManager.cpp:369-375): query the database for this timestamp
    auto feats = trackFEATS->get_feature_database()->features_containing(message.timestamp);
    size_t nobs = 0;
    for (auto &ft : feats)
      for (auto &kv : ft->timestamps) nobs += kv.second.back() == message.timestamp;
    auto ids = trackFEATS->get_last_ids();
    std::printf("frame %d db_features %zu feats_at_t %zu obs_at_t %zu last0 %zu last1 %zu\n", k, trackFEATS->get_feature_database()->size(),
                feats.size(), nobs, ids[0].size(), ids[1].size());
    prev = nobs;
  }
  // dump every feature's left-camera track for the wrapper
  auto all = trackFEATS->get_feature_database()->get_internal_data();
  for (auto &kv : all)
    for (auto &cam : kv.second->uvs)
      for (size_t i = 0; i < cam.second.size(); i++)
        std::printf("obs %zu %zu %.9g %.9g %.9g\n", kv.first, cam.first, kv.second->timestamps[cam.first][i], cam.second[i][0], cam.second[i][1]);
  (void)prev;
  return 0;
}
