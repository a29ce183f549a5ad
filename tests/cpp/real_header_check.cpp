// Syntax check of the UVIO_USE_REAL_OV_CORE branch of uvio_b200/cpp/TrackKLTGpu.h against the REFERENCE'S OWN headers
// (track/TrackBase.h, cam/CamBase.h, cam/CamEqui.h, feat/FeatureDatabase.h, utils/sensor_data.h under
// /root/reference/ov_core/src) with stub OpenCV / Eigen / Boost headers (tests/cpp/stubs) that declare only the names used.
// It proves what the mirror cannot: the class derives from the real ov_core::TrackBase, overrides only members that are
// virtual there (an `override` of a plain member is a hard error), reads cv::Mat / CamBase the way those headers declare
// them, and the estimator's calls through std::shared_ptr<TrackBase> bind (VioManager.cpp:133-135, VioManagerHelper.cpp:124,
// :209-210, :235-237).  g++ -fsyntax-only; nothing is linked or run.
#define UVIO_USE_REAL_OV_CORE
#include "uvio_b200/cpp/TrackDescriptorGpu.h"
#include "uvio_b200/cpp/TrackKLTGpu.h"

int main() {
  std::unordered_map<size_t, std::shared_ptr<ov_core::CamBase>> cams;
  std::shared_ptr<ov_core::TrackBase> trackFEATS(
      new ov_core::TrackKLTGpu(cams, 25, 0, true, ov_core::TrackBase::HistogramMethod::HISTOGRAM, 20, 5, 5, 10));
  trackFEATS->set_num_features(150);
  std::unordered_map<size_t, std::vector<cv::KeyPoint>> last_obs = trackFEATS->get_last_obs();
  std::unordered_map<size_t, std::vector<size_t>> last_ids = trackFEATS->get_last_ids();
  trackFEATS->change_feat_id(1, 2);
  ov_core::CameraData message;
  trackFEATS->feed_new_camera(message);
  std::shared_ptr<ov_core::FeatureDatabase> db = trackFEATS->get_feature_database();
  // the descriptor tracker behind the same base pointer (VioManager.cpp:137-140)
  std::shared_ptr<ov_core::TrackBase> trackDESC(
      new ov_core::TrackDescriptorGpu(cams, 25, 0, true, ov_core::TrackBase::HistogramMethod::HISTOGRAM, 20, 5, 5, 10, 0.70));
  trackDESC->set_num_features(150);
  trackDESC->feed_new_camera(message);
  trackDESC->change_feat_id(1, 2);
  return (int)last_obs.size() + (int)last_ids.size() + (db ? 1 : 0);
}
