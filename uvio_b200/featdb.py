"""ctypes binding of include/uvio_featdb.h: the device-side feature-track store (ov_core::FeatureDatabase's queries on the GPU).
Method names and argument meaning follow ov_core/src/feat/FeatureDatabase.h so that tests read like the reference's callers
(ov_msckf/src/core/VioManager.cpp:369-375, :579-591)."""
from __future__ import annotations

import ctypes as C
from typing import Dict, List

import numpy as np

from . import track

NOT_CONTAINING_NEWER, CONTAINING_OLDER, CONTAINING, GET_FEATURE, ALL = range(5)
_bound = None


def _lib():
    global _bound
    L = track.lib()
    if _bound is not L:
        vp, sz, ci, d = C.c_void_p, C.c_size_t, C.c_int, C.c_double
        L.uvio_featdb_create.argtypes = [ci, ci, ci, C.POINTER(vp)]
        L.uvio_featdb_destroy.argtypes = [vp]
        L.uvio_featdb_destroy.restype = None
        L.uvio_featdb_update.argtypes = [vp, d, vp, ci]
        L.uvio_featdb_update_from_tracker.argtypes = [vp, vp, d]
        L.uvio_featdb_query.argtypes = [vp, ci, d, sz, ci, ci, C.POINTER(ci), C.POINTER(ci), C.POINTER(ci)]
        L.uvio_featdb_fetch.argtypes = [vp] * 10
        L.uvio_featdb_result_device.argtypes = [vp, C.POINTER(vp)]
        L.uvio_featdb_mark_to_delete.argtypes = [vp, vp, ci]
        L.uvio_featdb_cleanup.argtypes = [vp]
        L.uvio_featdb_cleanup_measurements.argtypes = [vp, d]
        L.uvio_featdb_cleanup_measurements_exact.argtypes = [vp, d]
        L.uvio_featdb_change_feat_id.argtypes = [vp, sz, sz]
        L.uvio_featdb_size.argtypes = [vp, C.POINTER(ci), C.POINTER(ci)]
        L.uvio_featdb_update_batch.argtypes = [C.POINTER(vp), ci, vp, vp, vp, vp]
        L.uvio_featdb_update_from_trackers.argtypes = [C.POINTER(vp), C.POINTER(vp), ci, vp, vp]
        L.uvio_featdb_update_from_desc_trackers.argtypes = [C.POINTER(vp), C.POINTER(vp), ci, vp, vp]
        L.uvio_featdb_query_batch.argtypes = [C.POINTER(vp), ci, ci, vp, vp, ci, ci, vp, vp, vp, vp]
        L.uvio_featdb_cleanup_batch.argtypes = [C.POINTER(vp), ci, ci, vp, vp]
        _bound = L
    return L


class FeatureView:
    """What the estimator reads of an ov_core::Feature (feat/Feature.h:40-97)."""

    def __init__(self, featid: int, key: int, to_delete: bool):
        self.featid, self.key, self.to_delete = featid, key, to_delete
        self.uvs: Dict[int, np.ndarray] = {}
        self.uvs_norm: Dict[int, np.ndarray] = {}
        self.timestamps: Dict[int, np.ndarray] = {}


class DeviceFeatureDatabase:
    def __init__(self, max_observations: int = 1 << 16, max_features: int = 1 << 13, device: int = 0):
        self._h = C.c_void_p()
        rc = _lib().uvio_featdb_create(int(device), int(max_observations), int(max_features), C.byref(self._h))
        if rc != track.UVIO_OK:
            self._h = None
            raise track.UvioError(rc, "uvio_featdb_create")

    def close(self):
        if getattr(self, "_h", None):
            _lib().uvio_featdb_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc, what):
        if rc != track.UVIO_OK:
            raise track.UvioError(rc, what)

    def update(self, timestamp: float, obs: np.ndarray):
        """n update_feature calls: obs is an array of track.OBS_DTYPE (uvio_klt_get_new_observations)"""
        obs = np.ascontiguousarray(obs, track.OBS_DTYPE)
        self._check(_lib().uvio_featdb_update(self._h, float(timestamp), obs.ctypes.data, len(obs)), "uvio_featdb_update")

    def update_feature(self, fid, timestamp, cam_id, u, v, u_n, v_n):
        o = np.zeros(1, track.OBS_DTYPE)
        o[0] = (fid, cam_id, u, v, u_n, v_n)
        self.update(timestamp, o)

    def update_from_tracker(self, trk, timestamp: float):
        self._check(_lib().uvio_featdb_update_from_tracker(self._h, trk._h, float(timestamp)), "uvio_featdb_update_from_tracker")

    def _query(self, kind, timestamp=0.0, key=0, remove=False, skip_deleted=False) -> List[FeatureView]:
        nf, nr, no = C.c_int(), C.c_int(), C.c_int()
        self._check(_lib().uvio_featdb_query(self._h, kind, float(timestamp), int(key), int(remove), int(skip_deleted), C.byref(nf),
                                             C.byref(nr), C.byref(no)), "uvio_featdb_query")
        return self._fetch(nf.value, nr.value, no.value)

    def _fetch(self, nf, nr, no) -> List[FeatureView]:
        featid, key_a = np.zeros(max(nf, 1), np.uint64), np.zeros(max(nf, 1), np.uint64)
        dele = np.zeros(max(nf, 1), np.uint8)
        row0, row_cam, off = np.zeros(nf + 1, np.int32), np.zeros(max(nr, 1), np.int32), np.zeros(nr + 1, np.int32)
        t, uv, uvn = np.zeros(max(no, 1), np.float64), np.zeros((max(no, 1), 2), np.float32), np.zeros((max(no, 1), 2), np.float32)
        self._check(_lib().uvio_featdb_fetch(self._h, featid.ctypes.data, key_a.ctypes.data, dele.ctypes.data, row0.ctypes.data,
                                             row_cam.ctypes.data, off.ctypes.data, t.ctypes.data, uv.ctypes.data, uvn.ctypes.data), "uvio_featdb_fetch")
        out = []
        for k in range(nf):
            f = FeatureView(int(featid[k]), int(key_a[k]), bool(dele[k]))
            for r in range(row0[k], row0[k + 1]):
                c = int(row_cam[r])
                f.timestamps[c], f.uvs[c], f.uvs_norm[c] = t[off[r]:off[r + 1]], uv[off[r]:off[r + 1]], uvn[off[r]:off[r + 1]]
            out.append(f)
        return out

    def features_not_containing_newer(self, timestamp, remove=False, skip_deleted=False):
        return self._query(NOT_CONTAINING_NEWER, timestamp, 0, remove, skip_deleted)

    def features_containing_older(self, timestamp, remove=False, skip_deleted=False):
        return self._query(CONTAINING_OLDER, timestamp, 0, remove, skip_deleted)

    def features_containing(self, timestamp, remove=False, skip_deleted=False):
        return self._query(CONTAINING, timestamp, 0, remove, skip_deleted)

    def get_feature(self, fid, remove=False):
        r = self._query(GET_FEATURE, 0.0, fid, remove, False)
        return r[0] if r else None

    def get_internal_data(self):
        return {f.key: f for f in self._query(ALL)}

    def mark_to_delete(self, keys):
        a = np.ascontiguousarray(list(keys), np.uint64)
        self._check(_lib().uvio_featdb_mark_to_delete(self._h, a.ctypes.data, len(a)), "uvio_featdb_mark_to_delete")

    def cleanup(self):
        self._check(_lib().uvio_featdb_cleanup(self._h), "uvio_featdb_cleanup")

    def cleanup_measurements(self, timestamp):
        self._check(_lib().uvio_featdb_cleanup_measurements(self._h, float(timestamp)), "uvio_featdb_cleanup_measurements")

    def cleanup_measurements_exact(self, timestamp):
        self._check(_lib().uvio_featdb_cleanup_measurements_exact(self._h, float(timestamp)), "uvio_featdb_cleanup_measurements_exact")

    def change_feat_id(self, id_old, id_new):
        self._check(_lib().uvio_featdb_change_feat_id(self._h, int(id_old), int(id_new)), "uvio_featdb_change_feat_id")

    def size(self):
        nf, no = C.c_int(), C.c_int()
        _lib().uvio_featdb_size(self._h, C.byref(nf), C.byref(no))
        return nf.value

    def num_measurements(self):
        nf, no = C.c_int(), C.c_int()
        _lib().uvio_featdb_size(self._h, C.byref(nf), C.byref(no))
        return no.value


# ---- the same operations on many stores of one device in one launch (uvio_featdb_*_batch) ------------------------------------------
def _handles(dbs):
    return (C.c_void_p * len(dbs))(*[d._h for d in dbs])


def _check_batch(rc, status, what):
    if rc != track.UVIO_OK:
        bad = [i for i, s in enumerate(status) if s != track.UVIO_OK]
        raise track.UvioError(rc, f"{what} (stores {bad[:8]})")


def update_batch(dbs, timestamps, obs_lists):
    n = len(dbs)
    ts = np.ascontiguousarray(timestamps, np.float64)
    counts = np.array([len(o) for o in obs_lists], np.int32)
    allobs = np.ascontiguousarray(np.concatenate([np.asarray(o, track.OBS_DTYPE) for o in obs_lists]) if n else np.zeros(0, track.OBS_DTYPE))
    status = np.zeros(n, np.int32)
    rc = _lib().uvio_featdb_update_batch(_handles(dbs), n, ts.ctypes.data, allobs.ctypes.data, counts.ctypes.data, status.ctypes.data)
    _check_batch(rc, status, "uvio_featdb_update_batch")


def update_from_trackers(dbs, trackers, timestamps):
    n = len(dbs)
    ts = np.ascontiguousarray(timestamps, np.float64)
    status = np.zeros(n, np.int32)
    hs = (C.c_void_p * n)(*[t._h for t in trackers])
    fn = _lib().uvio_featdb_update_from_desc_trackers if type(trackers[0]).__name__ == "TrackDescriptor" else _lib().uvio_featdb_update_from_trackers
    rc = fn(_handles(dbs), hs, n, ts.ctypes.data, status.ctypes.data)
    _check_batch(rc, status, "uvio_featdb_update_from_trackers")


def query_batch(dbs, kind, timestamps, keys=None, remove=False, skip_deleted=False, fetch=True):
    """One query kind on every store in one launch; returns the per-store results (fetch=False: only the (features, rows,
    observations) counts -- the results stay on the device for uvio_tri_triangulate_featdb)."""
    n = len(dbs)
    ts = np.ascontiguousarray(timestamps, np.float64)
    ka = None if keys is None else np.ascontiguousarray(keys, np.uint64)
    nf, nr, no, status = (np.zeros(n, np.int32) for _ in range(4))
    rc = _lib().uvio_featdb_query_batch(_handles(dbs), n, int(kind), ts.ctypes.data, None if ka is None else ka.ctypes.data, int(remove), int(skip_deleted),
                                        nf.ctypes.data, nr.ctypes.data, no.ctypes.data, status.ctypes.data)
    _check_batch(rc, status, "uvio_featdb_query_batch")
    if not fetch:
        return list(zip(nf.tolist(), nr.tolist(), no.tolist()))
    return [d._fetch(int(a), int(b), int(c)) for d, a, b, c in zip(dbs, nf, nr, no)]


def cleanup_batch(dbs, mode=0, timestamps=None):
    n = len(dbs)
    ts = None if timestamps is None else np.ascontiguousarray(timestamps, np.float64)
    status = np.zeros(n, np.int32)
    rc = _lib().uvio_featdb_cleanup_batch(_handles(dbs), n, int(mode), None if ts is None else ts.ctypes.data, status.ctypes.data)
    _check_batch(rc, status, "uvio_featdb_cleanup_batch")
