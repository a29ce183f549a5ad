/* uvio_tri.h — C ABI of the batched feature triangulation.
 *
 * Stands in for ov_core::FeatureInitializer (ov_core/src/feat/FeatureInitializer.cpp): single_triangulation (:30-112),
 * single_triangulation_1d (:114-195), single_gaussnewton (:197-360) and compute_error (:362-421), which the updaters run
 * feature by feature on what the FeatureDatabase queries return (ov_msckf/src/update/UpdaterMSCKF.cpp, UpdaterSLAM.cpp) and
 * whose linear part VioManager::retriangulate_active_tracks repeats for the active tracks
 * (ov_msckf/src/core/VioManagerHelper.cpp:190-330).  Here every feature of a query result is solved by its own GPU thread
 * (fp64, the reference's summation order), straight from the device-resident result of uvio_featdb_query.
 *
 * Plain pointers and sizes only; UVIO_* status codes of uvio_klt.h.  No CPU fallback.
 */
#ifndef UVIO_TRI_H_
#define UVIO_TRI_H_

#include "uvio_featdb.h"

#ifdef __cplusplus
extern "C" {
#endif

/* One entry of clonesCAM[cam_id][timestamp] (FeatureInitializer.h:51-77 ClonePose): camera pose of a clone */
typedef struct UvioClonePose {
  size_t cam_id;
  double timestamp;
  double R_GtoC[9]; /* row-major rotation global -> camera */
  double p_CinG[3]; /* camera position in the global frame */
} UvioClonePose;

/* ov_core::FeatureInitializerOptions (FeatureInitializerOptions.h:35-69), same names and defaults */
typedef struct UvioTriOptions {
  int triangulate_1d;   /* false */
  int refine_features;  /* true */
  int max_runs;         /* 5 */
  double init_lamda;    /* 1e-3 */
  double max_lamda;     /* 1e10 */
  double min_dx;        /* 1e-6 */
  double min_dcost;     /* 1e-6 */
  double lam_mult;      /* 10 */
  double min_dist;      /* 0.10 */
  double max_dist;      /* 60 */
  double max_baseline;  /* 40 */
  double max_cond_number; /* 10000 */
} UvioTriOptions;

typedef struct UvioTriResult {
  int ok_triangulation; /* return value of single_triangulation(_1d); -1: a measurement has no clone pose (.at() throws in the reference) */
  int ok_refine;        /* return value of single_gaussnewton (0 when it was not run) */
  int anchor_cam_id;    /* Feature::anchor_cam_id */
  int n_meas;
  double anchor_clone_timestamp; /* Feature::anchor_clone_timestamp */
  double p_FinA[3];     /* Feature::p_FinA (as left by the last step that ran) */
  double p_FinG[3];     /* Feature::p_FinG */
} UvioTriResult;

void uvio_tri_default_options(UvioTriOptions *opt);

/* Triangulates (and, with refine_features, refines) every feature of the store's last query result, in result order.
 * out must hold n_features (of that query) entries.  Cameras with the same number of measurements tie for the anchor: the
 * reference takes whichever std::unordered_map iterates first (unspecified); here the higher camera id wins, which is what
 * libstdc++ yields for a feature first seen in camera 0. */
int uvio_tri_triangulate_featdb(uvio_featdb *db, const UvioClonePose *clones, int n_clones, const UvioTriOptions *opt, UvioTriResult *out);

/* The same for features given as host arrays in the layout of uvio_featdb_fetch: row0[n_features + 1], row_cam[rows],
 * off[rows + 1], t[observations], uvn[2 * observations]. */
int uvio_tri_triangulate(int cuda_device, int n_features, const int *row0, const int *row_cam, const int *off, const double *t,
                         const float *uvn, const UvioClonePose *clones, int n_clones, const UvioTriOptions *opt, UvioTriResult *out);

#ifdef __cplusplus
}
#endif
#endif /* UVIO_TRI_H_ */
