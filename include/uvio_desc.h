/* C ABI of the B200 descriptor tracker: stands in for ov_core::TrackDescriptor (ov_core/src/track/TrackDescriptor.h:38-173,
 * TrackDescriptor.cpp:33-575) behind the same TrackBase interface as the KLT tracker
 * (include/uvio_klt.h: UvioFrame = CameraData, UvioObs = one FeatureDatabase::update_feature call, UvioCamCalib, UVIO_* codes).
 *
 * Per frame and camera: equalise (TrackDescriptor.cpp:72-82) -> Grider_FAST::perform_griding (Grider_FAST.h:74-173) ->
 * cv::ORB::compute on those corners (:368, :424) -> BruteForce-Hamming knnMatch both ways + ratio + symmetry tests +
 * cv::findFundamentalMat (robust_match, :487-540) left->right and last->new -> ids and the database update (:252-345).
 * Everything runs as one chain of kernels per (batched) feed with a single synchronisation at its end.
 *
 * Differences from the reference, all on paths where the reference's own result is undefined or unordered:
 *  - robust_match reads mask_rsc[i] even when findFundamentalMat found no model and left the mask empty (:536-539);
 *    here no match survives in that case.
 *  - a non-stereo tracker with several cameras is fed by cv::parallel_for_ (:50-55), its cameras drawing ids from one
 *    counter concurrently (the id order changes from run to run); here the cameras take their ids one after the other in
 *    message order, the same on every run.  Up to four cameras per message.
 *  - 10..14 symmetric matches send cv::findFundamentalMat to LMedS: see include/uvio_klt.h on that regime.
 */
#ifndef UVIO_DESC_H_
#define UVIO_DESC_H_

#include "uvio_klt.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct uvio_desc uvio_desc; /* one tracker = one ov_core::TrackDescriptor instance */

/* Arguments of TrackDescriptor::TrackDescriptor (TrackDescriptor.h:55-59). */
typedef struct UvioDescParams {
  int num_features;     /* numfeats */
  int num_aruco;        /* numaruco: first feature id is 4*num_aruco + 2 (TrackBase.cpp:34) */
  int use_stereo;       /* stereo */
  int histogram_method; /* histmethod: UVIO_HIST_* */
  int fast_threshold;   /* fast_threshold */
  int grid_x, grid_y;   /* gridx, gridy */
  int min_px_dist;      /* minpxdist */
  double knn_ratio;     /* knnratio (config: 0.70) */
} UvioDescParams;

void uvio_desc_default_params(UvioDescParams *p);

/* TrackDescriptor::TrackDescriptor.  UVIO_E_CUDA when the device cannot be used: there is no CPU path. */
int uvio_desc_create(const UvioDescParams *params, const UvioCamCalib *cams, int n_cams, int cuda_device, uvio_desc **out);
void uvio_desc_destroy(uvio_desc *h);

/* TrackDescriptor::feed_new_camera (TrackDescriptor.cpp:33-60).  Synchronous: on return the observations of this frame can be
 * read.  UVIO_E_BADMSG where the reference exits (:36-42, :57-58), UVIO_E_CAMID where .at() throws, UVIO_E_SIZE when a
 * camera's image size changes.  Image pointers may be host or device memory; nothing is retained after return. */
int uvio_desc_feed(uvio_desc *h, const UvioFrame *frame);
/* One frame for each of n different trackers of one device, batched into the same launches; frames[i].status per message. */
int uvio_desc_feed_batch(uvio_desc *const *h, UvioFrame *frames, int n);

/* The FeatureDatabase::update_feature calls of the last feed, in the reference's order (:311-326, :152-155). */
int uvio_desc_get_new_observations(uvio_desc *h, UvioObs *out, int cap, int *n);
/* TrackBase::get_last_obs / get_last_ids (TrackBase.h:137-146) of one camera. */
int uvio_desc_get_last(uvio_desc *h, size_t cam_id, size_t *ids, float *uv, int cap, int *n);
/* desc_last[cam_id] (TrackDescriptor.h:170): n rows of 32 bytes, as cv::ORB writes them. */
int uvio_desc_get_last_descriptors(uvio_desc *h, size_t cam_id, uint8_t *desc, int cap, int *n);
/* TrackBase::set_num_features (TrackBase.h:152). */
int uvio_desc_set_num_features(uvio_desc *h, int n);
/* TrackBase::change_feat_id (TrackBase.cpp:230-248), the part that sticks: every id_old in ids_last becomes id_new. */
int uvio_desc_change_feat_id(uvio_desc *h, size_t id_old, size_t id_new);
int uvio_desc_cuda_error(uvio_desc *h);
/* CUDA-event time of the last feed (or batch) this tracker was part of, upload to results, in microseconds. */
int uvio_desc_get_timing(uvio_desc *h, float *us);

#ifdef __cplusplus
}
#endif
#endif /* UVIO_DESC_H_ */
