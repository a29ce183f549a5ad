"""ORACLE — TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

Line-by-line restatement of the reference's KLT front-end control flow, driving the real OpenCV kernels through cv2.

  reference                                                  here
  ---------------------------------------------------------  -------------------------------------------
  TrackKLT::feed_new_camera          TrackKLT.cpp:34-94      TrackKLTOracle.feed_new_camera
  TrackKLT::feed_monocular           TrackKLT.cpp:96-200     TrackKLTOracle._feed_monocular
  TrackKLT::feed_stereo              TrackKLT.cpp:202-393    TrackKLTOracle._feed_stereo
  TrackKLT::perform_detection_monocular  :395-528            TrackKLTOracle._detect_mono
  TrackKLT::perform_detection_stereo     :530-827            TrackKLTOracle._detect_stereo
  TrackKLT::perform_matching             :829-886            TrackKLTOracle._perform_matching
  Grider_GRID::perform_griding       Grider_GRID.h:74-180    perform_griding
  CamBase::undistort_cv              cam/CamBase.h:108       Cam.undistort
  FeatureDatabase::update_feature    feat/FeatureDatabase.cpp:59   FeatureDatabaseOracle.update_feature

Pinning rules: cv2.ipp.setUseIPP(False); LK is given the level-0 images + maxLevel (the Python binding
rejects pyramid lists; OpenCV then builds the identical pyramids/derivatives internally); the per-cell top-K order comes
from the real libstdc++ std::sort (oracle/stdsort.cpp).
"""
from __future__ import annotations

import ctypes
import os
import subprocess
from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence, Tuple

import cv2
import numpy as np

cv2.ipp.setUseIPP(False)

HIST_NONE, HIST_HISTOGRAM, HIST_CLAHE = 0, 1, 2
CAM_RADTAN, CAM_EQUI = 0, 1

_HERE = os.path.dirname(os.path.abspath(__file__))
_BUILD = os.path.join(_HERE, "_build")
_f32 = np.float32


def build_helpers(force: bool = False) -> None:
    """Compile the oracle's native helpers (std::sort permutation, C stage restatements) with the committed Makefile."""
    targets = [os.path.join(_BUILD, "libstdsort.so"), os.path.join(_BUILD, "libcstages.so")]
    if force or not all(os.path.exists(t) for t in targets):
        subprocess.check_call(["make", "-s", "-C", _HERE, "all"])


_sort_lib = None


def std_sort_perm(responses: np.ndarray) -> np.ndarray:
    """Permutation libstdc++ std::sort(compare_response) applies to keypoints with these responses (Grider_GRID.h:128)."""
    global _sort_lib
    if _sort_lib is None:
        build_helpers()
        _sort_lib = ctypes.CDLL(os.path.join(_BUILD, "libstdsort.so"))
        _sort_lib.oracle_std_sort_perm.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p]
    r = np.ascontiguousarray(responses, dtype=np.float32)
    perm = np.empty(r.shape[0], np.int32)
    if r.shape[0]:
        _sort_lib.oracle_std_sort_perm(r.ctypes.data, int(r.shape[0]), perm.ctypes.data)
    return perm


# ---------------------------------------------------------------------------------------------------------------------
# camera models (cam/CamRadtan.h:89-111, cam/CamEqui.h:108-130)
# ---------------------------------------------------------------------------------------------------------------------
@dataclass
class Cam:
    model: int
    intrinsics: Tuple[float, float, float, float]
    dist: Tuple[float, float, float, float]

    def __post_init__(self):
        fx, fy, cx, cy = self.intrinsics
        self.K = np.array([[fx, 0, cx], [0, fy, cy], [0, 0, 1]], np.float64)
        self.D = np.array(self.dist, np.float64)

    def undistort(self, pts: np.ndarray) -> np.ndarray:
        """undistort_cv on each point. The reference calls OpenCV once per point on a 1x1 Mat; OpenCV treats every
        point independently, so one batched call returns the same floats (checked by tests/test_oracle_golden.py::test_batched_undistort_equals_per_point_calls)."""
        pts = np.asarray(pts, np.float32).reshape(-1, 1, 2)
        if pts.shape[0] == 0:
            return np.zeros((0, 2), np.float32)
        if self.model == CAM_RADTAN:
            out = cv2.undistortPoints(pts, self.K, self.D)
        else:
            out = cv2.fisheye.undistortPoints(pts, self.K, self.D)
        return out.reshape(-1, 2).astype(np.float32)

    def max_focal(self) -> float:
        return max(self.K[0, 0], self.K[1, 1])


# ---------------------------------------------------------------------------------------------------------------------
# stage wrappers: the exact OpenCV calls of the reference
# ---------------------------------------------------------------------------------------------------------------------
def equalize(img: np.ndarray, method: int) -> np.ndarray:
    """TrackKLT.cpp:56-67."""
    if method == HIST_HISTOGRAM:
        return cv2.equalizeHist(img)
    if method == HIST_CLAHE:
        return cv2.createCLAHE(10.0, (8, 8)).apply(img)
    return img


def build_pyramid(img: np.ndarray, pyr_levels: int, win: int = 15):
    """TrackKLT.cpp:71 — returns (maxLevel, [img0, deriv0, img1, deriv1, ...])."""
    max_level, pyr = cv2.buildOpticalFlowPyramid(img, (win, win), pyr_levels)
    return max_level, list(pyr)


def fast_detect(roi: np.ndarray, threshold: int, nonmax: bool = True):
    """cv::FAST(img(roi), pts, threshold, nonmaxSuppression) — Grider_GRID.h:125. Returns (xy int32 [n,2], response f32 [n])."""
    det = cv2.FastFeatureDetector_create(int(threshold), bool(nonmax))
    kps = det.detect(np.ascontiguousarray(roi))
    xy = np.array([[int(k.pt[0]), int(k.pt[1])] for k in kps], np.int32).reshape(-1, 2)
    resp = np.array([k.response for k in kps], np.float32)
    return xy, resp


def corner_subpix(img: np.ndarray, pts: np.ndarray) -> np.ndarray:
    """cv::cornerSubPix(img, pts, (5,5), (-1,-1), COUNT+EPS 20 / 0.001) — Grider_GRID.h:163-174."""
    p = np.ascontiguousarray(pts, np.float32).reshape(-1, 1, 2).copy()
    if p.shape[0] == 0:
        return p.reshape(-1, 2)
    crit = (cv2.TERM_CRITERIA_COUNT + cv2.TERM_CRITERIA_EPS, 20, 0.001)
    cv2.cornerSubPix(img, p, (5, 5), (-1, -1), crit)
    return p.reshape(-1, 2)


def lk_track(img0: np.ndarray, img1: np.ndarray, pts0: np.ndarray, pts1_init: np.ndarray, pyr_levels: int,
             win: int = 15):
    """cv::calcOpticalFlowPyrLK(pyr0, pyr1, pts0, pts1, status, err, win, pyr_levels, {COUNT|EPS,30,0.01},
    OPTFLOW_USE_INITIAL_FLOW) — TrackKLT.cpp:671,858. Returns (pts1 f32 [n,2], status u8 [n])."""
    p0 = np.ascontiguousarray(pts0, np.float32).reshape(-1, 1, 2)
    p1 = np.ascontiguousarray(pts1_init, np.float32).reshape(-1, 1, 2).copy()
    crit = (cv2.TERM_CRITERIA_COUNT | cv2.TERM_CRITERIA_EPS, 30, 0.01)
    p1, st, _err = cv2.calcOpticalFlowPyrLK(img0, img1, p0, p1, winSize=(win, win), maxLevel=int(pyr_levels),
                                            criteria=crit, flags=cv2.OPTFLOW_USE_INITIAL_FLOW)
    return p1.reshape(-1, 2), st.reshape(-1).astype(np.uint8)


def fundamental_ransac_mask(p0n: np.ndarray, p1n: np.ndarray, thr: float) -> np.ndarray:
    """cv::findFundamentalMat(pts0_n, pts1_n, FM_RANSAC, thr, 0.999, mask) — TrackKLT.cpp:873. Returns the mask (may be
    empty when OpenCV found no model; the reference then treats every point as an outlier, :877)."""
    a = np.ascontiguousarray(p0n, np.float32).reshape(-1, 1, 2)
    b = np.ascontiguousarray(p1n, np.float32).reshape(-1, 1, 2)
    _F, mask = cv2.findFundamentalMat(a, b, cv2.FM_RANSAC, float(thr), 0.999)
    if mask is None:
        return np.zeros((0,), np.uint8)
    return mask.reshape(-1).astype(np.uint8)


# ---------------------------------------------------------------------------------------------------------------------
# Grider_GRID::perform_griding (Grider_GRID.h:74-180)
# ---------------------------------------------------------------------------------------------------------------------
def perform_griding(img: np.ndarray, mask: np.ndarray, valid_locs: Sequence[Tuple[int, int]], num_features: int,
                    grid_x: int, grid_y: int, threshold: int, trace: Optional[dict] = None):
    """Returns (pts f32 [n,2] refined, resp f32 [n]). ``trace`` (optional dict) receives the intermediate lists."""
    if len(valid_locs) == 0:                                                     # :84-85
        return np.zeros((0, 2), np.float32), np.zeros((0,), np.float32)
    if num_features < grid_x * grid_y:                                           # :88-92
        ratio = float(grid_x) / float(grid_y)
        grid_y = int(np.ceil(np.sqrt(num_features / ratio)))
        grid_x = int(np.ceil(grid_y * ratio))
    num_features_grid = int(float(num_features) / float(grid_x * grid_y)) + 1     # :93
    rows, cols = img.shape
    size_x = cols // grid_x                                                      # :99-100 (integer division)
    size_y = rows // grid_y
    out_xy: List[Tuple[float, float]] = []
    out_resp: List[float] = []
    cells_trace = []
    for (gx, gy) in valid_locs:                                                  # :108-151 (order of valid_locs kept, :154)
        x = gx * size_x
        y = gy * size_y
        if x + size_x > cols or y + size_y > rows:                               # :117
            cells_trace.append(None)
            continue
        xy, resp = fast_detect(img[y:y + size_y, x:x + size_x], threshold, True)  # :125
        perm = std_sort_perm(resp)                                               # :128
        kept = []
        for i in range(min(num_features_grid, len(perm))):                       # :133
            px = _f32(xy[perm[i], 0]) + _f32(x)
            py = _f32(xy[perm[i], 1]) + _f32(y)
            if int(px) < 0 or int(px) > cols or int(py) < 0 or int(py) > rows:    # :142
                continue
            if mask[int(py), int(px)] > 127:                                     # :146
                continue
            kept.append((float(px), float(py), float(resp[perm[i]])))
        cells_trace.append({"cell": (gx, gy), "raw_xy": xy, "raw_resp": resp, "perm": perm, "kept": kept})
        for k in kept:
            out_xy.append((k[0], k[1]))
            out_resp.append(k[2])
    if trace is not None:
        trace["cells"] = cells_trace
        trace["selected"] = np.array(out_xy, np.float32).reshape(-1, 2)
    if not out_xy:                                                               # :159-160
        return np.zeros((0, 2), np.float32), np.zeros((0,), np.float32)
    pts = corner_subpix(img, np.array(out_xy, np.float32))                       # :163-174
    if trace is not None:
        trace["refined"] = pts.copy()
    return pts, np.array(out_resp, np.float32)


# ---------------------------------------------------------------------------------------------------------------------
# FeatureDatabase (feat/FeatureDatabase.cpp:59-85) — only what the tracker writes; insertion-ordered for comparison
# ---------------------------------------------------------------------------------------------------------------------
class FeatureDatabaseOracle:
    def __init__(self):
        self.features: Dict[int, dict] = {}
        self.log: List[Tuple[int, float, int, float, float, float, float]] = []   # every update_feature call, in order

    def update_feature(self, fid: int, timestamp: float, cam_id: int, u, v, un, vn):
        f = self.features.get(fid)
        if f is None:
            f = {"featid": fid, "uvs": {}, "uvs_norm": {}, "timestamps": {}}
            self.features[fid] = f
        f["uvs"].setdefault(cam_id, []).append((float(u), float(v)))
        f["uvs_norm"].setdefault(cam_id, []).append((float(un), float(vn)))
        f["timestamps"].setdefault(cam_id, []).append(float(timestamp))
        self.log.append((int(fid), float(timestamp), int(cam_id), float(u), float(v), float(un), float(vn)))


# ---------------------------------------------------------------------------------------------------------------------
# TrackKLT
# ---------------------------------------------------------------------------------------------------------------------
Pt = Tuple[np.float32, np.float32]


def _occupancy_pass(pts: List[Pt], ids: List[int], mask: np.ndarray, mask_updated: np.ndarray, cols: int, rows: int,
                    min_px_dist: int, grid_x: int, grid_y: int, stereo_ids=None):
    """The erase-while-iterating loop of TrackKLT.cpp:409-464 / :545-600 / :722-780.
    ``stereo_ids`` (right pass only): ids exempt from the closeness erase (:754-759).
    Returns (pts, ids, grid_2d_close, grid_2d_grid); paints ``mask_updated`` in place."""
    size_close = (int(_f32(cols) / _f32(min_px_dist)), int(_f32(rows) / _f32(min_px_dist)))       # (w, h) :401-402
    grid_close = np.zeros((size_close[1], size_close[0]), np.uint8)
    size_x = _f32(cols) / _f32(grid_x)                                                            # :404-405
    size_y = _f32(rows) / _f32(grid_y)
    grid_grid = np.zeros((grid_y, grid_x), np.uint8)
    out_pts: List[Pt] = []
    out_ids: List[int] = []
    for (px, py), fid in zip(pts, ids):
        px = _f32(px)
        py = _f32(py)
        x = int(px)
        y = int(py)
        edge = 10
        if x < edge or x >= cols - edge or y < edge or y >= rows - edge:                          # :417
            continue
        x_close = int(px / _f32(min_px_dist))                                                     # :423-424
        y_close = int(py / _f32(min_px_dist))
        if x_close < 0 or x_close >= size_close[0] or y_close < 0 or y_close >= size_close[1]:
            continue
        x_grid = int(np.floor(px / size_x))                                                       # :431-432
        y_grid = int(np.floor(py / size_y))
        if x_grid < 0 or x_grid >= grid_x or y_grid < 0 or y_grid >= grid_y:
            continue
        if stereo_ids is None:
            if grid_close[y_close, x_close] > 127:                                                # :439
                continue
        else:
            is_stereo = fid in stereo_ids                                                         # :754
            if grid_close[y_close, x_close] > 127 and not is_stereo:                              # :755
                continue
        if mask[y, x] > 127:                                                                      # :446 / :762
            continue
        grid_close[y_close, x_close] = 255                                                        # :452
        if grid_grid[y_grid, x_grid] < 255:                                                       # :453-455
            grid_grid[y_grid, x_grid] += 1
        if x - min_px_dist >= 0 and x + min_px_dist < cols and y - min_px_dist >= 0 and y + min_px_dist < rows:   # :457
            mask_updated[y - min_px_dist:y + min_px_dist + 1, x - min_px_dist:x + min_px_dist + 1] = 255        # :460
        out_pts.append((px, py))
        out_ids.append(fid)
    return out_pts, out_ids, grid_close, grid_grid


def _valid_cells(mask: np.ndarray, grid_grid: np.ndarray, num_features: int, grid_x: int, grid_y: int):
    """TrackKLT.cpp:479-492 (x-outer, y-inner)."""
    mask_grid = cv2.resize(mask, (grid_x, grid_y), interpolation=cv2.INTER_NEAREST)               # :480
    num_features_grid = int(float(num_features) / float(grid_x * grid_y)) + 1                     # :483
    num_features_grid_req = max(1, int(0.5 * num_features_grid))                                  # :484
    valid = []
    for x in range(grid_x):
        for y in range(grid_y):
            if int(grid_grid[y, x]) < num_features_grid_req and int(mask_grid[y, x]) != 255:
                valid.append((x, y))
    return valid


def _close_reject(pts_ext: np.ndarray, grid_close: np.ndarray, min_px_dist: int):
    """TrackKLT.cpp:497-512 / :635-650 / :811-825. Returns indices kept (grid_close updated in place)."""
    h, w = grid_close.shape
    keep = []
    for i in range(pts_ext.shape[0]):
        x_grid = int(_f32(pts_ext[i, 0]) / _f32(min_px_dist))
        y_grid = int(_f32(pts_ext[i, 1]) / _f32(min_px_dist))
        if x_grid < 0 or x_grid >= w or y_grid < 0 or y_grid >= h:
            continue
        if grid_close[y_grid, x_grid] > 127:
            continue
        grid_close[y_grid, x_grid] = 255
        keep.append(i)
    return keep


class TrackKLTOracle:
    """Restatement of ov_core::TrackKLT (+ the TrackBase state it uses). One instance = one tracker."""

    def __init__(self, cams: Dict[int, Cam], numfeats: int, numaruco: int, stereo: bool, histmethod: int,
                 fast_threshold: int, gridx: int, gridy: int, minpxdist: int, pyr_levels: int = 5, win: int = 15):
        self.camera_calib = cams
        self.database = FeatureDatabaseOracle()                     # TrackBase.cpp:32
        self.num_features = int(numfeats)
        self.use_stereo = bool(stereo)
        self.histogram_method = int(histmethod)
        self.currid = 4 * int(numaruco) + 1                         # TrackBase.cpp:34
        self.threshold = int(fast_threshold)
        self.grid_x, self.grid_y = int(gridx), int(gridy)
        self.min_px_dist = int(minpxdist)
        self.pyr_levels = int(pyr_levels)                           # TrackKLT.h:143 (5 in the reference)
        self.win = int(win)                                         # TrackKLT.h:144
        self.img_last: Dict[int, np.ndarray] = {}
        self.img_mask_last: Dict[int, np.ndarray] = {}
        self.pts_last: Dict[int, List[Pt]] = {}
        self.ids_last: Dict[int, List[int]] = {}
        self.img_curr: Dict[int, np.ndarray] = {}
        self.trace: Optional[dict] = None       # set to {} before a feed to capture that frame's intermediates
        self.stage_seconds = {"pyramid": 0.0, "detection": 0.0, "klt": 0.0, "db": 0.0}

    # -- TrackBase API ---------------------------------------------------------------------------------------------
    def get_feature_database(self):
        return self.database

    def get_last_obs(self):
        return {c: list(p) for c, p in self.pts_last.items()}

    def get_last_ids(self):
        return {c: list(i) for c, i in self.ids_last.items()}

    def set_num_features(self, n: int):
        self.num_features = int(n)

    def change_feat_id(self, id_old: int, id_new: int):
        """TrackBase.cpp:230-248. The DB part acts on by-value copies in the reference: only
        feat->featid changes; the ids_last rewrite sticks."""
        f = self.database.features.get(id_old)
        if f is not None:
            f["featid"] = id_new
        for cam in self.ids_last:
            self.ids_last[cam] = [id_new if i == id_old else i for i in self.ids_last[cam]]

    # -- feed_new_camera (TrackKLT.cpp:34-94) ----------------------------------------------------------------------
    def feed_new_camera(self, timestamp: float, sensor_ids: Sequence[int], images: Sequence[np.ndarray],
                        masks: Sequence[np.ndarray]):
        import time
        if len(sensor_ids) == 0 or len(sensor_ids) != len(images) or len(images) != len(masks):
            raise SystemExit("[ERROR]: MESSAGE DATA SIZES DO NOT MATCH OR EMPTY!!!")      # std::exit(EXIT_FAILURE) :42
        t0 = time.perf_counter()
        for msg_id, cam_id in enumerate(sensor_ids):                                      # :50-76
            self.img_curr[cam_id] = equalize(images[msg_id], self.histogram_method)
            # the pyramid itself is rebuilt inside cv2.calcOpticalFlowPyrLK (pinning rule 3); build it here too so the
            # CPU-baseline timing pays for it like the reference does (:71)
            if self.trace is not None or self._time_pyramid:
                ml, pyr = build_pyramid(self.img_curr[cam_id], self.pyr_levels, self.win)
                if self.trace is not None:
                    self.trace.setdefault("equalized", {})[cam_id] = self.img_curr[cam_id].copy()
                    self.trace.setdefault("pyramid", {})[cam_id] = (ml, pyr)
        self.stage_seconds["pyramid"] += time.perf_counter() - t0
        n = len(images)
        if n == 1:
            self._feed_monocular(timestamp, sensor_ids, masks, 0)
        elif n == 2 and self.use_stereo:
            self._feed_stereo(timestamp, sensor_ids, masks, 0, 1)
        elif not self.use_stereo:
            for i in range(n):                                                            # parallel_for_ :85-89
                self._feed_monocular(timestamp, sensor_ids, masks, i)
        else:
            raise SystemExit("[ERROR]: invalid number of images passed")                  # :91-92

    _time_pyramid = True

    # -- feed_monocular (TrackKLT.cpp:96-200) ----------------------------------------------------------------------
    def _feed_monocular(self, timestamp, sensor_ids, masks, msg_id):
        import time
        cam_id = sensor_ids[msg_id]
        img = self.img_curr[cam_id]
        mask = masks[msg_id]
        rows, cols = img.shape
        if len(self.pts_last.get(cam_id, [])) == 0:                                       # :110-123
            t0 = time.perf_counter()
            good_left, good_ids_left = self._detect_mono(img, mask, [], [])
            self.stage_seconds["detection"] += time.perf_counter() - t0
            self.img_last[cam_id] = img
            self.img_mask_last[cam_id] = mask
            self.pts_last[cam_id] = good_left
            self.ids_last[cam_id] = good_ids_left
            return
        t0 = time.perf_counter()
        pts_left_old, ids_left_old = self._detect_mono(self.img_last[cam_id], self.img_mask_last[cam_id],
                                                       list(self.pts_last[cam_id]), list(self.ids_last[cam_id]))   # :130
        t1 = time.perf_counter()
        self.stage_seconds["detection"] += t1 - t0
        pts_left_old, pts_left_new, mask_ll = self._perform_matching(self.img_last[cam_id], img, pts_left_old,
                                                                     list(pts_left_old), cam_id, cam_id)          # :138
        t2 = time.perf_counter()
        self.stage_seconds["klt"] += t2 - t1
        if len(mask_ll) == 0:                                                             # :143-152
            self.img_last[cam_id] = img
            self.img_mask_last[cam_id] = mask
            self.pts_last[cam_id] = []
            self.ids_last[cam_id] = []
            return
        good_left, good_ids_left = [], []
        for i, (px, py) in enumerate(pts_left_new):                                       # :159-173
            if px < 0 or py < 0 or int(px) >= cols or int(py) >= rows:
                continue
            if int(mask[int(py), int(px)]) > 127:                                         # :166 (current message mask)
                continue
            if mask_ll[i]:
                good_left.append((px, py))
                good_ids_left.append(ids_left_old[i])
        self._db_update(timestamp, cam_id, good_left, good_ids_left)                      # :176-179
        self.img_last[cam_id] = img                                                       # :182-189
        self.img_mask_last[cam_id] = mask
        self.pts_last[cam_id] = good_left
        self.ids_last[cam_id] = good_ids_left
        self.stage_seconds["db"] += time.perf_counter() - t2

    def _db_update(self, timestamp, cam_id, pts, ids):
        if not pts:
            return
        npts = self.camera_calib[cam_id].undistort(np.array(pts, np.float32))
        for i in range(len(pts)):
            self.database.update_feature(ids[i], timestamp, cam_id, pts[i][0], pts[i][1], npts[i, 0], npts[i, 1])

    # -- feed_stereo (TrackKLT.cpp:202-393) ------------------------------------------------------------------------
    def _feed_stereo(self, timestamp, sensor_ids, masks, msg_id_left, msg_id_right):
        import time
        cl, cr = sensor_ids[msg_id_left], sensor_ids[msg_id_right]
        img_left, img_right = self.img_curr[cl], self.img_curr[cr]
        mask_left, mask_right = masks[msg_id_left], masks[msg_id_right]
        if len(self.pts_last.get(cl, [])) == 0 and len(self.pts_last.get(cr, [])) == 0:   # :221-240
            t0 = time.perf_counter()
            gl, gr, gil, gir = self._detect_stereo(img_left, img_right, mask_left, mask_right, [], [], [], [])
            self.stage_seconds["detection"] += time.perf_counter() - t0
            self.img_last[cl], self.img_last[cr] = img_left, img_right
            self.img_mask_last[cl], self.img_mask_last[cr] = mask_left, mask_right
            self.pts_last[cl], self.pts_last[cr] = gl, gr
            self.ids_last[cl], self.ids_last[cr] = gil, gir
            return
        t0 = time.perf_counter()
        pts_left_old, pts_right_old, ids_left_old, ids_right_old = self._detect_stereo(                # :249
            self.img_last[cl], self.img_last[cr], self.img_mask_last[cl], self.img_mask_last[cr],
            list(self.pts_last[cl]), list(self.pts_last[cr]), list(self.ids_last[cl]), list(self.ids_last[cr]))
        t1 = time.perf_counter()
        self.stage_seconds["detection"] += t1 - t0
        pts_left_old, pts_left_new, mask_ll = self._perform_matching(self.img_last[cl], img_left, pts_left_old,
                                                                     list(pts_left_old), cl, cl)       # :260-268
        pts_right_old, pts_right_new, mask_rr = self._perform_matching(self.img_last[cr], img_right, pts_right_old,
                                                                       list(pts_right_old), cr, cr)
        t2 = time.perf_counter()
        self.stage_seconds["klt"] += t2 - t1
        if self.trace is not None:
            self.trace["temporal"] = {"pts_left_old": list(pts_left_old), "pts_left_new": list(pts_left_new),
                                      "mask_ll": list(mask_ll), "pts_right_old": list(pts_right_old),
                                      "pts_right_new": list(pts_right_new), "mask_rr": list(mask_rr),
                                      "ids_left_old": list(ids_left_old), "ids_right_old": list(ids_right_old)}
        if len(mask_ll) == 0 and len(mask_rr) == 0:                                       # :285-299
            self.img_last[cl], self.img_last[cr] = img_left, img_right
            self.img_mask_last[cl], self.img_mask_last[cr] = mask_left, mask_right
            self.pts_last[cl], self.pts_last[cr] = [], []
            self.ids_last[cl], self.ids_last[cr] = [], []
            return
        rows_l, cols_l = img_left.shape
        rows_r, cols_r = img_right.shape
        good_left, good_right, good_ids_left, good_ids_right = [], [], [], []
        for i, (px, py) in enumerate(pts_left_new):                                       # :306-338
            if px < 0 or py < 0 or int(px) > cols_l or int(py) > rows_l:                  # :308-309 ('>' on the left)
                continue
            found_right, index_right = False, 0
            for n_, rid in enumerate(ids_right_old):                                      # :314-320
                if ids_left_old[i] == rid:
                    found_right, index_right = True, n_
                    break
            # NB: mask_ll[i] / mask_rr[index_right] are read unchecked in the reference; an empty mask (fewer than 1
            # point on that side) cannot be indexed because the loop/lookup is then empty as well.
            if mask_ll[i] and found_right and mask_rr[index_right]:                        # :323
                qx, qy = pts_right_new[index_right]
                if qx < 0 or qy < 0 or int(qx) >= cols_r or int(qy) >= rows_r:            # :325-327
                    continue
                good_left.append((px, py))
                good_right.append((qx, qy))
                good_ids_left.append(ids_left_old[i])
                good_ids_right.append(ids_right_old[index_right])
            elif mask_ll[i]:                                                              # :333-337
                good_left.append((px, py))
                good_ids_left.append(ids_left_old[i])
        for i, (qx, qy) in enumerate(pts_right_new):                                      # :341-354
            if qx < 0 or qy < 0 or int(qx) >= cols_r or int(qy) >= rows_r:
                continue
            added_already = ids_right_old[i] in good_ids_right                            # :347
            if mask_rr[i] and not added_already:
                good_right.append((qx, qy))
                good_ids_right.append(ids_right_old[i])
        self._db_update(timestamp, cl, good_left, good_ids_left)                          # :357-366
        self._db_update(timestamp, cr, good_right, good_ids_right)
        self.img_last[cl], self.img_last[cr] = img_left, img_right                        # :369-381
        self.img_mask_last[cl], self.img_mask_last[cr] = mask_left, mask_right
        self.pts_last[cl], self.pts_last[cr] = good_left, good_right
        self.ids_last[cl], self.ids_last[cr] = good_ids_left, good_ids_right
        self.stage_seconds["db"] += time.perf_counter() - t2

    # -- perform_detection_monocular (TrackKLT.cpp:395-528) --------------------------------------------------------
    def _detect_mono(self, img0, mask0, pts0: List[Pt], ids0: List[int]):
        rows, cols = img0.shape
        mask0_updated = mask0.copy()                                                      # :408
        pts0, ids0, grid_close, grid_grid = _occupancy_pass(pts0, ids0, mask0, mask0_updated, cols, rows,
                                                            self.min_px_dist, self.grid_x, self.grid_y)
        num_featsneeded = self.num_features - len(pts0)                                   # :469
        if num_featsneeded < min(20, int(0.5 * self.num_features)):                       # :470
            return pts0, ids0
        valid_locs = _valid_cells(mask0, grid_grid, self.num_features, self.grid_x, self.grid_y)
        tr = {} if self.trace is not None else None
        pts0_ext, _ = perform_griding(img0, mask0_updated, valid_locs, self.num_features, self.grid_x, self.grid_y,
                                      self.threshold, tr)                                 # :494
        if tr is not None:
            tr["valid_locs"] = valid_locs
            self.trace.setdefault("detect", []).append(tr)
        keep = _close_reject(pts0_ext, grid_close, self.min_px_dist)                      # :497-512
        for i in keep:                                                                    # :519-527
            pts0.append((_f32(pts0_ext[i, 0]), _f32(pts0_ext[i, 1])))
            self.currid += 1
            ids0.append(self.currid)
        return pts0, ids0

    # -- perform_detection_stereo (TrackKLT.cpp:530-827) -----------------------------------------------------------
    def _detect_stereo(self, img0, img1, mask0, mask1, pts0, pts1, ids0, ids1):
        rows0, cols0 = img0.shape
        rows1, cols1 = img1.shape
        mask0_updated = mask0.copy()                                                      # :544
        pts0, ids0, grid_close0, grid_grid0 = _occupancy_pass(pts0, ids0, mask0, mask0_updated, cols0, rows0,
                                                              self.min_px_dist, self.grid_x, self.grid_y)
        num_featsneeded_0 = self.num_features - len(pts0)                                 # :604
        if num_featsneeded_0 > min(20, int(0.5 * self.num_features)):                     # :609
            valid_locs = _valid_cells(mask0, grid_grid0, self.num_features, self.grid_x, self.grid_y)
            tr = {} if self.trace is not None else None
            pts0_ext, _ = perform_griding(img0, mask0_updated, valid_locs, self.num_features, self.grid_x,
                                          self.grid_y, self.threshold, tr)                # :632
            keep = _close_reject(pts0_ext, grid_close0, self.min_px_dist)                 # :635-650
            pts0_new = pts0_ext[keep].astype(np.float32).reshape(-1, 2)
            if tr is not None:
                tr["valid_locs"] = valid_locs
                tr["side"] = "left"
                tr["pts0_new"] = pts0_new.copy()
                self.trace.setdefault("detect", []).append(tr)
            if pts0_new.shape[0] > 0:                                                     # :662
                pts1_new, status = lk_track(img0, img1, pts0_new, pts0_new.copy(), self.pyr_levels, self.win)   # :671
                if tr is not None:
                    tr["pts1_new"] = pts1_new.copy()
                    tr["stereo_status"] = status.copy()
                for i in range(pts0_new.shape[0]):                                        # :675-708
                    x0, y0 = _f32(pts0_new[i, 0]), _f32(pts0_new[i, 1])
                    x1, y1 = _f32(pts1_new[i, 0]), _f32(pts1_new[i, 1])
                    oob_left = int(x0) < 0 or int(x0) >= cols0 or int(y0) < 0 or int(y0) >= rows0
                    oob_right = int(x1) < 0 or int(x1) >= cols1 or int(y1) < 0 or int(y1) >= rows1
                    if (not oob_left) and (not oob_right) and status[i] == 1:
                        pts0.append((x0, y0))
                        pts1.append((x1, y1))
                        self.currid += 1
                        ids0.append(self.currid)
                        ids1.append(self.currid)
                    elif not oob_left:
                        pts0.append((x0, y0))
                        self.currid += 1
                        ids0.append(self.currid)
        # RIGHT (:712-826)
        mask1_updated = mask0.copy()                                                      # :721 (sic: mask0)
        pts1, ids1, grid_close1, grid_grid1 = _occupancy_pass(pts1, ids1, mask1, mask1_updated, cols1, rows1,
                                                              self.min_px_dist, self.grid_x, self.grid_y,
                                                              stereo_ids=set(ids0))
        num_featsneeded_1 = self.num_features - len(pts1)                                 # :784
        if num_featsneeded_1 > min(20, int(0.5 * self.num_features)):                     # :785
            valid_locs = _valid_cells(mask1, grid_grid1, self.num_features, self.grid_x, self.grid_y)   # :794-806
            tr = {} if self.trace is not None else None
            pts1_ext, _ = perform_griding(img1, mask1_updated, valid_locs, self.num_features, self.grid_x,
                                          self.grid_y, self.threshold, tr)                # :808
            if tr is not None:
                tr["valid_locs"] = valid_locs
                tr["side"] = "right"
                self.trace.setdefault("detect", []).append(tr)
            keep = _close_reject(pts1_ext, grid_close1, self.min_px_dist)                 # :811-825
            for i in keep:
                pts1.append((_f32(pts1_ext[i, 0]), _f32(pts1_ext[i, 1])))
                self.currid += 1
                ids1.append(self.currid)
        return pts0, pts1, ids0, ids1

    # -- perform_matching (TrackKLT.cpp:829-886) -------------------------------------------------------------------
    def _perform_matching(self, img0, img1, kpts0: List[Pt], kpts1: List[Pt], id0: int, id1: int):
        """Returns (kpts0, kpts1, mask_out). mask_out == [] when there were no points (:836-837)."""
        assert len(kpts0) == len(kpts1)
        if len(kpts0) == 0:
            return kpts0, kpts1, []
        if len(kpts0) < 10:                                                               # :848-852
            return kpts0, kpts1, [0] * len(kpts0)
        pts0 = np.array(kpts0, np.float32).reshape(-1, 2)
        pts1_init = np.array(kpts1, np.float32).reshape(-1, 2)
        pts1, mask_klt = lk_track(img0, img1, pts0, pts1_init, self.pyr_levels, self.win)  # :858
        pts0_n = self.camera_calib[id0].undistort(pts0)                                   # :863-866
        pts1_n = self.camera_calib[id1].undistort(pts1)
        max_focallength = max(self.camera_calib[id0].max_focal(), self.camera_calib[id1].max_focal())   # :870-872
        mask_rsc = fundamental_ransac_mask(pts0_n, pts1_n, 2.0 / max_focallength)         # :873
        mask_out = []
        for i in range(len(mask_klt)):                                                    # :876-879
            mask_out.append(1 if (mask_klt[i] and i < len(mask_rsc) and mask_rsc[i]) else 0)
        if self.trace is not None:
            self.trace.setdefault("matching", []).append({"pts0": pts0.copy(), "pts1": pts1.copy(),
                                                          "klt": mask_klt.copy(), "rsc": mask_rsc.copy(),
                                                          "pts0_n": pts0_n.copy(), "pts1_n": pts1_n.copy()})
        kpts1 = [(_f32(pts1[i, 0]), _f32(pts1[i, 1])) for i in range(pts1.shape[0])]      # :882-885
        return kpts0, kpts1, mask_out



def downsample_message(images: Sequence[np.ndarray], masks: Sequence[np.ndarray]):
    """The step VioManager::track_image_and_update runs on every message just before trackFEATS->feed_new_camera when
    params.downsample_cameras is set (ov_msckf/src/core/VioManager.cpp:268-278): cv::pyrDown of each image and of each
    mask to cv::Size(cols / 2.0, rows / 2.0) (doubles truncated by the cv::Size(int, int) constructor, so an odd
    dimension loses its last row / column instead of rounding up as pyrDown's default size would).  The camera
    intrinsics handed to the tracker are halved where the options are parsed (VioManagerOptions.h:251-260)."""
    out_i, out_m = [], []
    for img, mask in zip(images, masks):
        rows, cols = img.shape[:2]
        out_i.append(cv2.pyrDown(img, dstsize=(int(cols / 2.0), int(rows / 2.0))))
        rows, cols = mask.shape[:2]
        out_m.append(cv2.pyrDown(mask, dstsize=(int(cols / 2.0), int(rows / 2.0))))
    return out_i, out_m

def make_oracle(cfg, **over) -> "TrackKLTOracle":
    """Build an oracle tracker from a uvio_b200.configs.TrackerConfig-shaped object (duck-typed: no product import)."""
    cams = {i: Cam(c.model, tuple(c.intrinsics), tuple(c.dist)) for i, c in enumerate(cfg.cams)}
    kw = dict(numfeats=cfg.num_features, numaruco=cfg.num_aruco, stereo=cfg.use_stereo, histmethod=cfg.histogram,
              fast_threshold=cfg.fast_threshold, gridx=cfg.grid_x, gridy=cfg.grid_y, minpxdist=cfg.min_px_dist,
              pyr_levels=cfg.pyr_levels)
    kw.update(over)
    return TrackKLTOracle(cams, **kw)
