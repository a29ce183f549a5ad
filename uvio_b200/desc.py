"""ctypes binding of include/uvio_desc.h plus a Python mirror of ov_core::TrackDescriptor (track/TrackDescriptor.h:38-173),
used by the tests.  Same library and conventions as uvio_b200.track (CameraData, FeatureDatabase, error codes); there is
no CPU fallback: creating a tracker fails without a GPU."""
from __future__ import annotations

import ctypes as C
from typing import Dict, Sequence

import numpy as np

from . import track
from .track import OBS_DTYPE, UVIO_E_BADMSG, UVIO_OK, CameraData, FeatureDatabase, UvioCamCalib, UvioError, UvioFrame, _pack_frame


class UvioDescParams(C.Structure):
    _fields_ = [("num_features", C.c_int), ("num_aruco", C.c_int), ("use_stereo", C.c_int), ("histogram_method", C.c_int),
                ("fast_threshold", C.c_int), ("grid_x", C.c_int), ("grid_y", C.c_int), ("min_px_dist", C.c_int),
                ("knn_ratio", C.c_double)]


_bound = None


def lib():
    global _bound
    L = track.lib()
    if _bound is not L:
        vp, sz, ci = C.c_void_p, C.c_size_t, C.c_int
        L.uvio_desc_default_params.argtypes = [C.POINTER(UvioDescParams)]
        L.uvio_desc_default_params.restype = None
        L.uvio_desc_create.argtypes = [C.POINTER(UvioDescParams), C.POINTER(UvioCamCalib), ci, ci, C.POINTER(vp)]
        L.uvio_desc_destroy.argtypes = [vp]
        L.uvio_desc_destroy.restype = None
        L.uvio_desc_feed.argtypes = [vp, C.POINTER(UvioFrame)]
        L.uvio_desc_feed_batch.argtypes = [C.POINTER(vp), C.POINTER(UvioFrame), ci]
        L.uvio_desc_get_new_observations.argtypes = [vp, vp, ci, C.POINTER(ci)]
        L.uvio_desc_get_last.argtypes = [vp, sz, vp, vp, ci, C.POINTER(ci)]
        L.uvio_desc_get_last_descriptors.argtypes = [vp, sz, vp, ci, C.POINTER(ci)]
        L.uvio_desc_set_num_features.argtypes = [vp, ci]
        L.uvio_desc_change_feat_id.argtypes = [vp, sz, sz]
        L.uvio_desc_cuda_error.argtypes = [vp]
        L.uvio_desc_get_timing.argtypes = [vp, C.POINTER(C.c_float)]
        _bound = L
    return L


class TrackDescriptor:
    """GPU ov_core::TrackDescriptor behind the C ABI.  Constructor arguments as TrackDescriptor.h:55-59."""

    def __init__(self, camera_calib: Dict[int, object], numfeats: int, numaruco: int, stereo: bool, histmethod: int, fast_threshold: int,
                 gridx: int, gridy: int, minpxdist: int, knnratio: float, device: int = 0):
        L = lib()
        p = UvioDescParams()
        L.uvio_desc_default_params(C.byref(p))
        p.num_features, p.num_aruco, p.use_stereo, p.histogram_method = int(numfeats), int(numaruco), int(bool(stereo)), int(histmethod)
        p.fast_threshold, p.grid_x, p.grid_y, p.min_px_dist, p.knn_ratio = int(fast_threshold), int(gridx), int(gridy), int(minpxdist), float(knnratio)
        cams = (UvioCamCalib * len(camera_calib))()
        for i, (cid, c) in enumerate(camera_calib.items()):
            cams[i].cam_id = int(cid)
            cams[i].model = int(c.model)
            for k in range(4):
                cams[i].intrinsics[k] = float(c.intrinsics[k])
                cams[i].distortion[k] = float(c.dist[k])
        self._h = C.c_void_p()
        rc = L.uvio_desc_create(C.byref(p), cams, len(camera_calib), int(device), C.byref(self._h))
        if rc != UVIO_OK:
            self._h = None
            raise UvioError(rc, "uvio_desc_create")
        self.camera_calib = dict(camera_calib)
        self.database = FeatureDatabase()
        self.num_features = int(numfeats)
        self._obs = np.zeros(8192, OBS_DTYPE)

    def close(self):
        if getattr(self, "_h", None):
            lib().uvio_desc_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def feed_new_camera(self, message: CameraData):
        keep: list = []
        f = _pack_frame(message, keep)
        rc = lib().uvio_desc_feed(self._h, C.byref(f))
        self._after_feed(rc, message.timestamp)

    def _after_feed(self, rc: int, timestamp: float):
        if rc == UVIO_E_BADMSG:
            raise SystemExit("[ERROR]: MESSAGE DATA SIZES DO NOT MATCH OR EMPTY!!!")   # std::exit(EXIT_FAILURE), TrackDescriptor.cpp:41
        if rc != UVIO_OK:
            raise UvioError(rc, "uvio_desc_feed", lib().uvio_desc_cuda_error(self._h))
        for o in self.new_observations():
            self.database.update_feature(int(o["id"]), float(timestamp), int(o["cam_id"]), float(o["u"]), float(o["v"]), float(o["un"]), float(o["vn"]))

    def new_observations(self) -> np.ndarray:
        n = C.c_int(0)
        L = lib()
        L.uvio_desc_get_new_observations(self._h, self._obs.ctypes.data, len(self._obs), C.byref(n))
        if n.value > len(self._obs):
            self._obs = np.zeros(n.value * 2, OBS_DTYPE)
            L.uvio_desc_get_new_observations(self._h, self._obs.ctypes.data, len(self._obs), C.byref(n))
        return self._obs[:n.value].copy()

    def get_feature_database(self) -> FeatureDatabase:
        return self.database

    def _last(self, cam_id: int):
        n = C.c_int(0)
        L = lib()
        rc = L.uvio_desc_get_last(self._h, int(cam_id), None, None, 0, C.byref(n))
        if rc not in (UVIO_OK, track.UVIO_E_SIZE):
            raise UvioError(rc, "uvio_desc_get_last")
        m = max(n.value, 1)
        ids, uv, desc = np.zeros(m, np.uint64), np.zeros((m, 2), np.float32), np.zeros((m, 32), np.uint8)
        rc = L.uvio_desc_get_last(self._h, int(cam_id), ids.ctypes.data, uv.ctypes.data, m, C.byref(n))
        if rc != UVIO_OK:
            raise UvioError(rc, "uvio_desc_get_last")
        rc = L.uvio_desc_get_last_descriptors(self._h, int(cam_id), desc.ctypes.data, m, C.byref(n))
        if rc != UVIO_OK:
            raise UvioError(rc, "uvio_desc_get_last_descriptors")
        return ids[:n.value], uv[:n.value], desc[:n.value]

    def get_last_obs(self) -> Dict[int, np.ndarray]:
        return {cid: self._last(cid)[1] for cid in self.camera_calib}

    def get_last_ids(self) -> Dict[int, np.ndarray]:
        return {cid: self._last(cid)[0] for cid in self.camera_calib}

    def get_last_descriptors(self) -> Dict[int, np.ndarray]:
        return {cid: self._last(cid)[2] for cid in self.camera_calib}

    def set_num_features(self, n: int):
        self.num_features = int(n)
        rc = lib().uvio_desc_set_num_features(self._h, int(n))
        if rc != UVIO_OK:
            raise UvioError(rc, "uvio_desc_set_num_features")

    def change_feat_id(self, id_old: int, id_new: int):
        """TrackBase.cpp:230-248: only feat->featid and ids_last change."""
        f = self.database.features_idlookup.get(id_old)
        if f is not None:
            f.featid = id_new
        rc = lib().uvio_desc_change_feat_id(self._h, int(id_old), int(id_new))
        if rc != UVIO_OK:
            raise UvioError(rc, "uvio_desc_change_feat_id")

    def last_feed_us(self) -> float:
        v = C.c_float(0)
        lib().uvio_desc_get_timing(self._h, C.byref(v))
        return float(v.value)


def feed_batch(trackers: Sequence[TrackDescriptor], messages: Sequence[CameraData]):
    """uvio_desc_feed_batch: one message for each of n different trackers of one device, in the same launches."""
    n = len(trackers)
    keep: list = []
    frames = (UvioFrame * n)()
    for i, m in enumerate(messages):
        frames[i] = _pack_frame(m, keep)
    hs = (C.c_void_p * n)(*[t._h for t in trackers])
    rc = lib().uvio_desc_feed_batch(hs, frames, n)
    for t, f, m in zip(trackers, frames, messages):
        t._after_feed(f.status, m.timestamp)
    return rc


def make_descriptor_tracker(cfg, knn_ratio: float = 0.70, device: int = 0, **over) -> TrackDescriptor:
    kw = dict(numfeats=cfg.num_features, numaruco=cfg.num_aruco, stereo=cfg.use_stereo, histmethod=cfg.histogram, fast_threshold=cfg.fast_threshold,
              gridx=cfg.grid_x, gridy=cfg.grid_y, minpxdist=cfg.min_px_dist, knnratio=knn_ratio)
    kw.update(over)
    return TrackDescriptor({i: c for i, c in enumerate(cfg.cams)}, device=device, **kw)
