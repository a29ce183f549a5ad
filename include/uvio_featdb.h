/* uvio_featdb.h — C ABI of the device-side feature-track store.
 *
 * Stands in for ov_core::FeatureDatabase (ov_core/src/feat/FeatureDatabase.h, .cpp:36-263; ov_core/src/feat/Feature.h:40-97),
 * the object every tracker fills through update_feature and the estimator queries once per frame
 * (ov_msckf/src/core/VioManager.cpp:369-375, :462-466, :579-591; VioManagerHelper.cpp:61, :123).  The reference keeps an
 * unordered_map<size_t, shared_ptr<Feature>> of per-camera std::vectors under one mutex; here the observations of a tracker
 * live in HBM as a structure-of-arrays log in arrival order plus a feature table (insertion order) and an id hash, and the
 * queries run as one thread block per store and return COMPACT per-feature observation lists (CSR), which is all the
 * estimator reads from a Feature (featid, to_delete, uvs / uvs_norm / timestamps per camera).
 *
 * Result layout of a query (device-side until fetched): n_features selected features in insertion order (the reference's
 * order is that of std::unordered_map iteration, i.e. unspecified); feature k owns rows row0[k] .. row0[k+1]-1, one row per
 * camera that has measurements (ascending camera id); row r owns observations off[r] .. off[r+1]-1 in time order:
 * timestamps t[], pixel coordinates uv[2*], normalised coordinates uvn[2*] -- Feature::timestamps / uvs / uvs_norm.
 *
 * Plain pointers and sizes only, nothing throws, every function returns an UVIO_* status of uvio_klt.h.  No CPU fallback.
 */
#ifndef UVIO_FEATDB_H_
#define UVIO_FEATDB_H_

#include "uvio_desc.h"
#include "uvio_klt.h"

#ifdef __cplusplus
extern "C" {
#endif

#define UVIO_FEATDB_MAX_CAMS 4 /* camera ids 0..3 per store (the reference's rigs are mono or stereo) */

/* query kinds */
#define UVIO_FDB_NOT_CONTAINING_NEWER 0 /* FeatureDatabase::features_not_containing_newer (FeatureDatabase.cpp:87-126) */
#define UVIO_FDB_CONTAINING_OLDER 1     /* FeatureDatabase::features_containing_older    (FeatureDatabase.cpp:128-167) */
#define UVIO_FDB_CONTAINING 2           /* FeatureDatabase::features_containing          (FeatureDatabase.cpp:169-209) */
#define UVIO_FDB_GET_FEATURE 3          /* FeatureDatabase::get_feature(id, remove)      (FeatureDatabase.cpp:36-57); the id goes in `key` */
#define UVIO_FDB_ALL 4                  /* FeatureDatabase::get_internal_data            (FeatureDatabase.h:146-149) */

typedef struct uvio_featdb uvio_featdb;

/* FeatureDatabase::FeatureDatabase.  max_observations / max_features bound what the store holds between two cleanups; an
 * update that does not fit returns UVIO_E_SIZE and leaves the store unchanged. */
int uvio_featdb_create(int cuda_device, int max_observations, int max_features, uvio_featdb **out);
void uvio_featdb_destroy(uvio_featdb *db);

/* n FeatureDatabase::update_feature(id, timestamp, cam_id, u, v, u_n, v_n) calls (FeatureDatabase.cpp:59-85), in order: the
 * observation list of one feed (uvio_klt_get_new_observations). */
int uvio_featdb_update(uvio_featdb *db, double timestamp, const UvioObs *obs, int n);
/* the same, taking the list straight from the tracker's last feed (what TrackKLTGpu::replay_observations does on the host) */
int uvio_featdb_update_from_tracker(uvio_featdb *db, uvio_klt *tracker, double timestamp);

/* One of the queries above.  remove / skip_deleted as in the reference.  The result stays on the device; the counts come
 * back here (any may be NULL). */
int uvio_featdb_query(uvio_featdb *db, int kind, double timestamp, size_t key, int remove, int skip_deleted, int *n_features,
                      int *n_rows, int *n_observations);
/* Copies the last query's result to the host (any pointer may be NULL): featid[n_features] (Feature::featid), key[n_features]
 * (the map key: differs from featid after change_feat_id, TrackBase.cpp:230-248), to_delete[n_features], row0[n_features + 1],
 * row_cam[n_rows], off[n_rows + 1], t[n_observations], uv[2 * n_observations], uvn[2 * n_observations]. */
int uvio_featdb_fetch(uvio_featdb *db, size_t *featid, size_t *key, uint8_t *to_delete, int *row0, int *row_cam, int *off, double *t,
                      float *uv, float *uvn);
/* Device pointers of the same arrays, for consumers that stay on the GPU (the batched triangulation): order as in fetch. */
int uvio_featdb_result_device(uvio_featdb *db, const void *ptrs[9]);

/* feat->to_delete = true on features a query returned (what the updaters do, e.g. UpdaterMSCKF.cpp; consumed by cleanup) */
int uvio_featdb_mark_to_delete(uvio_featdb *db, const size_t *keys, int n);
/* FeatureDatabase::cleanup (FeatureDatabase.cpp:211-224): drops every feature marked to_delete */
int uvio_featdb_cleanup(uvio_featdb *db);
/* FeatureDatabase::cleanup_measurements (FeatureDatabase.cpp:226-243): drops measurements with time <= timestamp, then
 * features left without any */
int uvio_featdb_cleanup_measurements(uvio_featdb *db, double timestamp);
/* FeatureDatabase::cleanup_measurements_exact (FeatureDatabase.cpp:245-263): drops measurements at exactly timestamp */
int uvio_featdb_cleanup_measurements_exact(uvio_featdb *db, double timestamp);
/* the database half of TrackBase::change_feat_id as it really behaves (TrackBase.cpp:233-238 act on a copy of the map, so
 * only feat->featid changes and the feature stays under its old key) */
int uvio_featdb_change_feat_id(uvio_featdb *db, size_t id_old, size_t id_new);
/* FeatureDatabase::size (FeatureDatabase.h:138-141) and the number of stored measurements */
int uvio_featdb_size(uvio_featdb *db, int *n_features, int *n_observations);

/* ---- the same operations on n stores of one device in ONE launch (one block per store): what a host that runs many
 * trackers side by side calls once per step instead of n times.  Results per store as in the single calls; the return value is
 * the worst status, status[i] (may be NULL) the one of store i. ---- */
/* counts[i] observations of store i, concatenated in obs; timestamps[i] its message time */
int uvio_featdb_update_batch(uvio_featdb *const *dbs, int n, const double *timestamps, const UvioObs *obs, const int *counts, int *status);
/* the observation lists come straight from the trackers' last (batched) feed */
int uvio_featdb_update_from_trackers(uvio_featdb *const *dbs, uvio_klt *const *trackers, int n, const double *timestamps, int *status);
/* the same for descriptor trackers (include/uvio_desc.h) */
int uvio_featdb_update_from_desc_trackers(uvio_featdb *const *dbs, uvio_desc *const *trackers, int n, const double *timestamps, int *status);
/* one query kind for all; keys may be NULL unless kind = UVIO_FDB_GET_FEATURE; n_features / n_rows / n_observations[n] may be NULL */
int uvio_featdb_query_batch(uvio_featdb *const *dbs, int n, int kind, const double *timestamps, const size_t *keys, int remove, int skip_deleted,
                            int *n_features, int *n_rows, int *n_observations, int *status);
/* mode 0 = cleanup, 1 = cleanup_measurements(timestamps[i]), 2 = cleanup_measurements_exact(timestamps[i]) */
int uvio_featdb_cleanup_batch(uvio_featdb *const *dbs, int n, int mode, const double *timestamps, int *status);

#ifdef __cplusplus
}
#endif
#endif /* UVIO_FEATDB_H_ */
