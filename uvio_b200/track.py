"""ctypes binding of include/uvio_klt.h plus a Python mirror of the reference's tracker API, used by tests and bench.py.

Mirrors (names and argument meaning kept):
  ov_core::CameraData        utils/sensor_data.h:55-80
  ov_core::TrackBase         track/TrackBase.h:72-196   (feed_new_camera, get_feature_database, get_last_obs/ids, ...)
  ov_core::TrackKLT          track/TrackKLT.h:43-152
  ov_core::FeatureDatabase   feat/FeatureDatabase.cpp:59-85 (update_feature), feat/Feature.h:40-97

The product host side for C++ callers is uvio_b200/cpp/TrackKLTGpu.h; this module is the same boundary for Python.
There is no CPU fallback: loading fails loudly if libuvio_klt.so is missing, creating a tracker fails without a GPU.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass, field
from typing import Dict, List, Sequence, Tuple

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("UVIO_KLT_LIB") or os.path.join(_HERE, "libuvio_klt.so")   # the override is for A/B builds of the same library

UVIO_OK, UVIO_E_BADMSG, UVIO_E_CAMID, UVIO_E_CUDA, UVIO_E_ARG, UVIO_E_SIZE = range(6)
ERR_NAMES = {0: "UVIO_OK", 1: "UVIO_E_BADMSG", 2: "UVIO_E_CAMID", 3: "UVIO_E_CUDA", 4: "UVIO_E_ARG", 5: "UVIO_E_SIZE"}


class UvioKltParams(C.Structure):
    _fields_ = [("num_features", C.c_int), ("num_aruco", C.c_int), ("use_stereo", C.c_int), ("histogram_method", C.c_int),
                ("fast_threshold", C.c_int), ("grid_x", C.c_int), ("grid_y", C.c_int), ("min_px_dist", C.c_int),
                ("pyr_levels", C.c_int), ("win_size", C.c_int), ("lk_max_iters", C.c_int), ("lk_eps", C.c_double),
                ("downsample_cameras", C.c_int)]


class UvioCamCalib(C.Structure):
    _fields_ = [("cam_id", C.c_size_t), ("model", C.c_int), ("intrinsics", C.c_double * 4), ("distortion", C.c_double * 4)]


class UvioObs(C.Structure):
    _fields_ = [("id", C.c_size_t), ("cam_id", C.c_size_t), ("u", C.c_float), ("v", C.c_float), ("un", C.c_float),
                ("vn", C.c_float)]


class UvioFrame(C.Structure):
    _fields_ = [("timestamp", C.c_double), ("n_img", C.c_int), ("sensor_ids", C.POINTER(C.c_size_t)),
                ("img", C.POINTER(C.c_void_p)), ("img_stride", C.POINTER(C.c_size_t)), ("mask", C.POINTER(C.c_void_p)),
                ("mask_stride", C.POINTER(C.c_size_t)), ("width", C.POINTER(C.c_int)), ("height", C.POINTER(C.c_int)),
                ("status", C.c_int)]


OBS_DTYPE = np.dtype([("id", np.uint64), ("cam_id", np.uint64), ("u", np.float32), ("v", np.float32), ("un", np.float32),
                      ("vn", np.float32)])
assert OBS_DTYPE.itemsize == C.sizeof(UvioObs)

_lib = None


class UvioError(RuntimeError):
    def __init__(self, code: int, what: str, cuda: int = 0):
        super().__init__(f"{what}: {ERR_NAMES.get(code, code)}" + (f" (cudaError {cuda})" if cuda else ""))
        self.code = code


def lib():
    """The C-ABI library. Raises if it has not been built (python -m uvio_b200.build / __graft_entry__.build())."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} is missing: build it with `python -m uvio_b200.build` (no CPU fallback exists)")
        L = C.CDLL(LIB_PATH)
        vp, sz, ci = C.c_void_p, C.c_size_t, C.c_int
        L.uvio_klt_default_params.argtypes = [C.POINTER(UvioKltParams)]
        L.uvio_klt_default_params.restype = None
        L.uvio_klt_create.argtypes = [C.POINTER(UvioKltParams), C.POINTER(UvioCamCalib), ci, ci, C.POINTER(vp)]
        L.uvio_klt_destroy.argtypes = [vp]
        L.uvio_klt_destroy.restype = None
        L.uvio_klt_feed.argtypes = [vp, C.POINTER(UvioFrame)]
        L.uvio_klt_feed_batch.argtypes = [C.POINTER(vp), C.POINTER(UvioFrame), ci]
        L.uvio_klt_submit.argtypes = [vp, C.POINTER(UvioFrame)]
        L.uvio_klt_submit_batch.argtypes = [C.POINTER(vp), C.POINTER(UvioFrame), ci]
        L.uvio_klt_get_new_observations.argtypes = [vp, vp, ci, C.POINTER(ci)]
        L.uvio_klt_get_last.argtypes = [vp, sz, vp, vp, ci, C.POINTER(ci)]
        L.uvio_klt_set_num_features.argtypes = [vp, ci]
        L.uvio_klt_change_feat_id.argtypes = [vp, sz, sz]
        L.uvio_klt_get_timing.argtypes = [vp, C.POINTER(C.c_float)]
        L.uvio_klt_last_error.argtypes = [vp]
        L.uvio_klt_cuda_error.argtypes = [vp]
        L.uvio_klt_version.restype = C.c_char_p
        L.uvio_klt_stage_get_level.argtypes = [vp, sz, ci, ci, vp, sz, C.POINTER(ci), C.POINTER(ci), C.POINTER(ci)]
        L.uvio_klt_stage_fast.argtypes = [vp, sz, ci, ci, ci, ci, ci, ci, vp, vp, ci, C.POINTER(ci)]
        L.uvio_klt_stage_subpix.argtypes = [vp, sz, ci, vp, ci, vp]
        L.uvio_klt_stage_lk.argtypes = [vp, sz, ci, sz, ci, vp, vp, ci, vp, vp, vp, vp]
        L.uvio_klt_stage_ransac.argtypes = [vp, sz, sz, vp, vp, ci, vp, vp, vp]
        L.uvio_klt_device_stats.argtypes = [ci, C.POINTER(C.c_ulonglong)]
        L.uvio_klt_timer_start.argtypes = [ci]
        L.uvio_klt_timer_stop.argtypes = [ci, C.POINTER(C.c_float)]
        L.uvio_klt_profile.argtypes = [ci, ci, C.POINTER(C.c_double), C.POINTER(C.c_ulonglong)]
        L.uvio_klt_host_profile.argtypes = [ci, ci, C.POINTER(C.c_double)]
        L.uvio_klt_get_new_observations_batch.argtypes = [C.POINTER(vp), ci, vp, ci, C.POINTER(ci)]
        _lib = L
    return _lib


# ---------------------------------------------------------------------------------------------------------------------
# reference-shaped host types
# ---------------------------------------------------------------------------------------------------------------------
@dataclass
class CameraData:
    """utils/sensor_data.h:55-80"""
    timestamp: float = -1.0
    sensor_ids: List[int] = field(default_factory=list)
    images: List[np.ndarray] = field(default_factory=list)
    masks: List[np.ndarray] = field(default_factory=list)


@dataclass
class Feature:
    """feat/Feature.h:40-97"""
    featid: int
    to_delete: bool = False
    uvs: Dict[int, List[Tuple[float, float]]] = field(default_factory=dict)
    uvs_norm: Dict[int, List[Tuple[float, float]]] = field(default_factory=dict)
    timestamps: Dict[int, List[float]] = field(default_factory=dict)


class FeatureDatabase:
    """feat/FeatureDatabase.cpp:59-85 (the part the tracker writes) + the queries the estimator makes most."""

    def __init__(self):
        self.features_idlookup: Dict[int, Feature] = {}
        self.log: List[tuple] = []      # every update_feature call in order (tests compare it with the oracle's)

    def update_feature(self, fid: int, timestamp: float, cam_id: int, u: float, v: float, u_n: float, v_n: float):
        f = self.features_idlookup.get(fid)
        if f is None:
            f = Feature(fid)
            self.features_idlookup[fid] = f
        f.uvs.setdefault(cam_id, []).append((u, v))
        f.uvs_norm.setdefault(cam_id, []).append((u_n, v_n))
        f.timestamps.setdefault(cam_id, []).append(timestamp)
        self.log.append((fid, timestamp, cam_id, u, v, u_n, v_n))

    def get_feature(self, fid: int):
        return self.features_idlookup.get(fid)

    def size(self) -> int:
        return len(self.features_idlookup)

    def get_internal_data(self):
        return dict(self.features_idlookup)       # by value, as feat/FeatureDatabase.h:146-149

    def features_containing(self, timestamp: float):
        """feat/FeatureDatabase.cpp:169-209 (without removal)"""
        return [f for f in self.features_idlookup.values() if any(timestamp in t for t in f.timestamps.values())]

    def features_not_containing_newer(self, timestamp: float):
        """feat/FeatureDatabase.cpp:87-126 (without removal)"""
        return [f for f in self.features_idlookup.values()
                if not any(t and t[-1] > timestamp for t in f.timestamps.values())]


class DeviceImage:
    """A CV_8UC1 image that already lives in device memory (consumed in place by the tracker, no copy): raw pointer, size
    and row stride in bytes.  ``DeviceImage.of(x)`` wraps anything with ``__cuda_array_interface__`` (a torch CUDA tensor)."""

    def __init__(self, ptr: int, height: int, width: int, stride: int, owner=None):
        self.ptr, self.shape, self.strides, self.owner = int(ptr), (int(height), int(width)), (int(stride), 1), owner
        self.ctypes = type("_P", (), {"data": int(ptr)})()

    @staticmethod
    def of(x) -> "DeviceImage":
        ai = x.__cuda_array_interface__
        if ai["typestr"] not in ("|u1", "<u1") or len(ai["shape"]) != 2:
            raise TypeError("device images must be 2-D uint8")
        h, w = ai["shape"]
        st = ai.get("strides") or (w, 1)
        if st[1] != 1:
            raise TypeError("device images must have unit pixel stride")
        return DeviceImage(ai["data"][0], h, w, st[0], owner=x)


def _as_u8(a):
    if isinstance(a, DeviceImage):
        return a
    if hasattr(a, "__cuda_array_interface__"):
        return DeviceImage.of(a)
    if a.dtype != np.uint8 or a.ndim != 2:
        raise TypeError("images / masks must be 2-D uint8 arrays (CV_8UC1)")
    if a.strides[1] != 1:
        a = np.ascontiguousarray(a)
    return a


def _pack_frame(message: CameraData, keep: list) -> UvioFrame:
    n = len(message.images)
    ids = (C.c_size_t * max(n, 1))(*[int(i) for i in message.sensor_ids[:n]])
    imgs = [_as_u8(a) for a in message.images]
    masks = [None if m is None else _as_u8(m) for m in message.masks[:n]]
    masks += [None] * (n - len(masks))
    ip = (C.c_void_p * max(n, 1))(*[a.ctypes.data for a in imgs])
    istr = (C.c_size_t * max(n, 1))(*[a.strides[0] for a in imgs])
    mp = (C.c_void_p * max(n, 1))(*[None if m is None else m.ctypes.data for m in masks])
    mstr = (C.c_size_t * max(n, 1))(*[0 if m is None else m.strides[0] for m in masks])
    w = (C.c_int * max(n, 1))(*[a.shape[1] for a in imgs])
    h = (C.c_int * max(n, 1))(*[a.shape[0] for a in imgs])
    keep.extend([ids, imgs, masks, ip, istr, mp, mstr, w, h])
    f = UvioFrame()
    f.timestamp = float(message.timestamp)
    f.n_img = n if len(message.sensor_ids) == n and len(message.masks) == n else 0   # TrackKLT.cpp:37-43 size check
    f.sensor_ids = C.cast(ids, C.POINTER(C.c_size_t))
    f.img = C.cast(ip, C.POINTER(C.c_void_p))
    f.img_stride = C.cast(istr, C.POINTER(C.c_size_t))
    f.mask = C.cast(mp, C.POINTER(C.c_void_p))
    f.mask_stride = C.cast(mstr, C.POINTER(C.c_size_t))
    f.width = C.cast(w, C.POINTER(C.c_int))
    f.height = C.cast(h, C.POINTER(C.c_int))
    for m, a in zip(masks, imgs):
        if m is not None and m.shape != a.shape:
            f.n_img = 0
    return f


class TrackKLT:
    """GPU ov_core::TrackKLT behind the C ABI. Constructor arguments as TrackKLT.h:54-57; ``camera_calib`` maps
    camera id -> object with (model, intrinsics, dist) — e.g. uvio_b200.configs.CamCalib."""

    def __init__(self, camera_calib: Dict[int, object], numfeats: int, numaruco: int, stereo: bool, histmethod: int,
                 fast_threshold: int, gridx: int, gridy: int, minpxdist: int, pyr_levels: int = 5, device: int = 0,
                 downsample_cameras: bool = False):
        L = lib()
        p = UvioKltParams()
        L.uvio_klt_default_params(C.byref(p))
        p.num_features, p.num_aruco, p.use_stereo, p.histogram_method = int(numfeats), int(numaruco), int(bool(stereo)), int(histmethod)
        p.fast_threshold, p.grid_x, p.grid_y, p.min_px_dist, p.pyr_levels = int(fast_threshold), int(gridx), int(gridy), int(minpxdist), int(pyr_levels)
        p.downsample_cameras = int(bool(downsample_cameras))     # VioManager.cpp:268-278 in front of feed_new_camera
        cams = (UvioCamCalib * len(camera_calib))()
        for i, (cid, c) in enumerate(camera_calib.items()):
            cams[i].cam_id = int(cid)
            cams[i].model = int(c.model)
            for k in range(4):
                cams[i].intrinsics[k] = float(c.intrinsics[k])
                cams[i].distortion[k] = float(c.dist[k])
        self._h = C.c_void_p()
        rc = L.uvio_klt_create(C.byref(p), cams, len(camera_calib), int(device), C.byref(self._h))
        if rc != UVIO_OK:
            self._h = None
            raise UvioError(rc, "uvio_klt_create")
        self.camera_calib = dict(camera_calib)
        self.database = FeatureDatabase()                       # TrackBase.cpp:32
        self.num_features = int(numfeats)
        self._obs = np.zeros(4096, OBS_DTYPE)
        self._announced: list = []

    def close(self):
        if getattr(self, "_h", None):
            lib().uvio_klt_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- TrackBase API ----------------------------------------------------------------------------------------------
    def feed_new_camera(self, message: CameraData):
        keep: list = []
        f = _pack_frame(message, keep)
        rc = lib().uvio_klt_feed(self._h, C.byref(f))
        self._after_feed(rc, message.timestamp)

    def submit(self, message: CameraData):
        """uvio_klt_submit: announce the next message (its upload and pre-processing start now); feed the same message next."""
        keep: list = []
        f = _pack_frame(message, keep)
        rc = lib().uvio_klt_submit(self._h, C.byref(f))
        if rc != UVIO_OK:
            raise UvioError(rc, "uvio_klt_submit", lib().uvio_klt_cuda_error(self._h))
        self._announced.append(keep)          # the image memory must outlive the feed

    def _after_feed(self, rc: int, timestamp: float):
        if self._announced:
            self._announced.pop(0)
        if rc == UVIO_E_BADMSG:
            raise SystemExit("[ERROR]: MESSAGE DATA SIZES DO NOT MATCH OR EMPTY!!!")   # std::exit(EXIT_FAILURE), TrackKLT.cpp:42
        if rc != UVIO_OK:
            raise UvioError(rc, "uvio_klt_feed", lib().uvio_klt_cuda_error(self._h))
        for o in self.new_observations():
            self.database.update_feature(int(o["id"]), float(timestamp), int(o["cam_id"]), float(o["u"]), float(o["v"]),
                                         float(o["un"]), float(o["vn"]))

    def new_observations(self) -> np.ndarray:
        n = C.c_int(0)
        L = lib()
        L.uvio_klt_get_new_observations(self._h, self._obs.ctypes.data, len(self._obs), C.byref(n))
        if n.value > len(self._obs):
            self._obs = np.zeros(n.value * 2, OBS_DTYPE)
            L.uvio_klt_get_new_observations(self._h, self._obs.ctypes.data, len(self._obs), C.byref(n))
        return self._obs[:n.value].copy()

    def get_feature_database(self) -> FeatureDatabase:
        return self.database

    def _last(self, cam_id: int):
        n = C.c_int(0)
        L = lib()
        rc = L.uvio_klt_get_last(self._h, int(cam_id), None, None, 0, C.byref(n))
        if rc != UVIO_OK:
            raise UvioError(rc, "uvio_klt_get_last")
        ids = np.zeros(max(n.value, 1), np.uint64)
        uv = np.zeros((max(n.value, 1), 2), np.float32)
        L.uvio_klt_get_last(self._h, int(cam_id), ids.ctypes.data, uv.ctypes.data, n.value, C.byref(n))
        return ids[:n.value], uv[:n.value]

    def get_last_obs(self) -> Dict[int, np.ndarray]:
        return {cid: self._last(cid)[1] for cid in self.camera_calib}

    def get_last_ids(self) -> Dict[int, np.ndarray]:
        return {cid: self._last(cid)[0] for cid in self.camera_calib}

    def set_num_features(self, n: int):
        self.num_features = int(n)
        rc = lib().uvio_klt_set_num_features(self._h, int(n))
        if rc != UVIO_OK:
            raise UvioError(rc, "uvio_klt_set_num_features")

    def change_feat_id(self, id_old: int, id_new: int):
        """TrackBase.cpp:230-248: only feat->featid and ids_last change."""
        f = self.database.features_idlookup.get(id_old)
        if f is not None:
            f.featid = id_new
        lib().uvio_klt_change_feat_id(self._h, int(id_old), int(id_new))

    def get_timing(self) -> np.ndarray:
        t = (C.c_float * 5)()
        lib().uvio_klt_get_timing(self._h, t)
        return np.array(list(t), np.float32)

    # -- stage entry points -----------------------------------------------------------------------------------------
    def stage_level(self, cam_id: int, slot: int, level: int) -> np.ndarray:
        w, h, nl = C.c_int(), C.c_int(), C.c_int()
        L = lib()
        rc = L.uvio_klt_stage_get_level(self._h, cam_id, slot, level, None, 0, C.byref(w), C.byref(h), C.byref(nl))
        if rc != UVIO_OK:
            raise UvioError(rc, "stage_get_level")
        out = np.empty((h.value, w.value), np.uint8)
        rc = L.uvio_klt_stage_get_level(self._h, cam_id, slot, level, out.ctypes.data, out.strides[0], C.byref(w), C.byref(h), C.byref(nl))
        if rc != UVIO_OK:
            raise UvioError(rc, "stage_get_level")
        return out

    def stage_num_levels(self, cam_id: int, slot: int = 0) -> int:
        w, h, nl = C.c_int(), C.c_int(), C.c_int()
        rc = lib().uvio_klt_stage_get_level(self._h, cam_id, slot, 0, None, 0, C.byref(w), C.byref(h), C.byref(nl))
        if rc != UVIO_OK:
            raise UvioError(rc, "stage_get_level")
        return nl.value

    def stage_fast(self, cam_id: int, slot: int, x0: int, y0: int, cw: int, ch: int, threshold: int):
        cap = ((cw + 1) // 2) * ((ch + 1) // 2)
        xy = np.empty((cap, 2), np.int32)
        resp = np.empty(cap, np.float32)
        n = C.c_int()
        rc = lib().uvio_klt_stage_fast(self._h, cam_id, slot, x0, y0, cw, ch, threshold, xy.ctypes.data, resp.ctypes.data, cap, C.byref(n))
        if rc != UVIO_OK:
            raise UvioError(rc, "stage_fast", lib().uvio_klt_cuda_error(self._h))
        return xy[:n.value].copy(), resp[:n.value].copy()

    def stage_subpix(self, cam_id: int, slot: int, pts: np.ndarray) -> np.ndarray:
        p = np.ascontiguousarray(pts, np.float32).reshape(-1, 2)
        out = np.empty_like(p)
        rc = lib().uvio_klt_stage_subpix(self._h, cam_id, slot, p.ctypes.data, len(p), out.ctypes.data)
        if rc != UVIO_OK:
            raise UvioError(rc, "stage_subpix", lib().uvio_klt_cuda_error(self._h))
        return out

    def stage_lk(self, cam0: int, slot0: int, cam1: int, slot1: int, pts0, pts1_init, with_ransac: bool = False):
        p0 = np.ascontiguousarray(pts0, np.float32).reshape(-1, 2)
        p1 = np.ascontiguousarray(pts1_init, np.float32).reshape(-1, 2)
        out = np.empty_like(p0)
        st = np.zeros(len(p0), np.uint8)
        mask = np.zeros(len(p0), np.uint8) if with_ransac else None
        p1n = np.zeros_like(p0) if with_ransac else None
        rc = lib().uvio_klt_stage_lk(self._h, cam0, slot0, cam1, slot1, p0.ctypes.data, p1.ctypes.data, len(p0), out.ctypes.data,
                                     st.ctypes.data, None if mask is None else mask.ctypes.data, None if p1n is None else p1n.ctypes.data)
        if rc != UVIO_OK:
            raise UvioError(rc, "stage_lk", lib().uvio_klt_cuda_error(self._h))
        return (out, st, mask, p1n) if with_ransac else (out, st)

    def stage_ransac(self, cam0: int, cam1: int, pts0, pts1):
        p0 = np.ascontiguousarray(pts0, np.float32).reshape(-1, 2)
        p1 = np.ascontiguousarray(pts1, np.float32).reshape(-1, 2)
        mask = np.zeros(len(p0), np.uint8)
        p0n, p1n = np.zeros_like(p0), np.zeros_like(p0)
        rc = lib().uvio_klt_stage_ransac(self._h, cam0, cam1, p0.ctypes.data, p1.ctypes.data, len(p0), mask.ctypes.data,
                                         p0n.ctypes.data, p1n.ctypes.data)
        if rc != UVIO_OK:
            raise UvioError(rc, "stage_ransac", lib().uvio_klt_cuda_error(self._h))
        return mask, p0n, p1n


def feed_batch(trackers: Sequence[TrackKLT], messages: Sequence[CameraData]):
    """uvio_klt_feed_batch: one frame for each of n independent trackers (same device), GPU work batched."""
    n = len(trackers)
    keep: list = []
    frames = (UvioFrame * n)(*[_pack_frame(m, keep) for m in messages])
    hs = (C.c_void_p * n)(*[t._h for t in trackers])
    rc = lib().uvio_klt_feed_batch(hs, frames, n)
    for t, m, f in zip(trackers, messages, frames):
        t._after_feed(f.status if rc in (UVIO_OK, f.status) else rc, m.timestamp)


def submit_batch(trackers: Sequence[TrackKLT], messages: Sequence[CameraData]):
    """uvio_klt_submit_batch: announce the next frame of each of n trackers (one upload + one pre-processing launch set)."""
    n = len(trackers)
    keep: list = []
    frames = (UvioFrame * n)(*[_pack_frame(m, keep) for m in messages])
    hs = (C.c_void_p * n)(*[t._h for t in trackers])
    rc = lib().uvio_klt_submit_batch(hs, frames, n)
    if rc != UVIO_OK:
        raise UvioError(rc, "uvio_klt_submit_batch", lib().uvio_klt_cuda_error(trackers[0]._h))
    for t in trackers:
        t._announced.append(keep)


def make_tracker(cfg, device: int = 0, **over) -> TrackKLT:
    """Tracker for a uvio_b200.configs.TrackerConfig."""
    kw = dict(numfeats=cfg.num_features, numaruco=cfg.num_aruco, stereo=cfg.use_stereo, histmethod=cfg.histogram,
              fast_threshold=cfg.fast_threshold, gridx=cfg.grid_x, gridy=cfg.grid_y, minpxdist=cfg.min_px_dist,
              pyr_levels=cfg.pyr_levels, device=device)
    kw.update(over)
    return TrackKLT({i: c for i, c in enumerate(cfg.cams)}, **kw)


KERNEL_NAMES = ["k_hist", "k_lut_apply", "k_clahe_tiles", "k_clahe_apply", "k_pyrdown", "k_fast", "k_subpix", "k_lk", "k_ransac",
                "k_preprocess", "k_downsample", "k_dv_logic", "k_dv_topk"]


def device_stats(device: int = 0) -> dict:
    out = (C.c_ulonglong * 12)()
    rc = lib().uvio_klt_device_stats(device, out)
    if rc != UVIO_OK:
        raise UvioError(rc, "uvio_klt_device_stats")
    return {"frames": out[0], "launches": out[1], "h2d_bytes": out[2], "d2h_bytes": out[3], "feed_calls": out[4],
            "fast_pixels": out[5], "lk_point_levels": out[6], "lk_iterations": out[7], "lk_chain_iterations": out[8]}


def profile(device: int = 0, enable: int = -1) -> dict:
    """Per-kernel CUDA-event sums since profiling was last enabled: name -> (ms, launches)."""
    ms = (C.c_double * 16)()
    ln = (C.c_ulonglong * 16)()
    rc = lib().uvio_klt_profile(device, enable, ms, ln)
    if rc != UVIO_OK:
        raise UvioError(rc, "uvio_klt_profile")
    return {k: (ms[i], int(ln[i])) for i, k in enumerate(KERNEL_NAMES)}


HOST_SEGMENTS = ("issue_r1", "wait_r1", "host_r2", "wait_r2", "host_r3", "wait_r3", "host_merge", "lane_passes")


def host_profile(device: int = 0, reset: bool = False) -> dict:
    """Host wall-clock (us) the lanes of `device` spent in each segment of the feed, summed over lanes."""
    us = (C.c_double * 8)()
    rc = lib().uvio_klt_host_profile(device, 1 if reset else 0, us)
    if rc != UVIO_OK:
        raise UvioError(rc, "uvio_klt_host_profile")
    return {k: us[i] for i, k in enumerate(HOST_SEGMENTS)}
