"""ORACLE — TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  Parity unpinned by the reference (it holds no vectors for
this class and cannot be built here); pinned by the real OpenCV kernels at the reference's call sites.

Line-by-line restatement of the reference's descriptor tracker, driving the real OpenCV kernels through cv2 4.13:

  reference                                                        here
  ---------------------------------------------------------------  ----------------------------------------------
  TrackDescriptor::feed_new_camera      TrackDescriptor.cpp:33-60   TrackDescriptorOracle.feed_new_camera
  TrackDescriptor::feed_monocular       :62-174                     ._feed_monocular
  TrackDescriptor::feed_stereo          :176-356                    ._feed_stereo
  TrackDescriptor::perform_detection_monocular  :358-405            ._detect_mono
  TrackDescriptor::perform_detection_stereo     :407-485            ._detect_stereo
  TrackDescriptor::robust_match         :487-540                    ._robust_match
  robust_ratio_test / robust_symmetry_test  :542-575                _ratio_test / _symmetry_test
  Grider_FAST::perform_griding          Grider_FAST.h:74-173        perform_griding_fast
  cv::ORB::create()->compute            TrackDescriptor.h:150-151   orb_compute  (default ORB: 31x31 patch, WTA_K 2, edge 31)
  DescriptorMatcher "BruteForce-Hamming" knnMatch(k=2)  :154        knn_match

The parallel_for_ bodies of the reference (two cameras, grid cells) write disjoint outputs, so their order is immaterial.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence, Tuple

import cv2
import numpy as np

from .trackklt_cv2 import (Cam, FeatureDatabaseOracle, corner_subpix, equalize, fast_detect, std_sort_perm)

cv2.ipp.setUseIPP(False)
_f32 = np.float32

Match = Tuple[int, int, float]          # (queryIdx, trainIdx, distance)


def perform_griding_fast(img: np.ndarray, mask: np.ndarray, num_features: int, grid_x: int, grid_y: int, threshold: int,
                         trace: Optional[dict] = None):
    """Grider_FAST::perform_griding (Grider_FAST.h:74-173). Returns (pts f32 [n,2] refined, resp f32 [n])."""
    if num_features < grid_x * grid_y:                                           # :83-87
        ratio = float(grid_x) / float(grid_y)
        grid_y = int(np.ceil(np.sqrt(num_features / ratio)))
        grid_x = int(np.ceil(grid_y * ratio))
    num_features_grid = int(float(num_features) / float(grid_x * grid_y)) + 1     # :88
    rows, cols = img.shape
    size_x = cols // grid_x                                                      # :94-95
    size_y = rows // grid_y
    ct_cols = cols // size_x                                                     # :102-103 (floor of an integer division)
    ct_rows = rows // size_y
    out_xy: List[Tuple[float, float]] = []
    out_resp: List[float] = []
    for r in range(ct_cols * ct_rows):                                           # :105-147; collections joined in r order :150-152
        x = r % ct_cols * size_x
        y = r // ct_cols * size_y
        if x + size_x > cols or y + size_y > rows:                               # :112
            continue
        xy, resp = fast_detect(img[y:y + size_y, x:x + size_x], threshold, True)  # :120
        perm = std_sort_perm(resp)                                               # :123
        for i in range(min(num_features_grid, len(perm))):                       # :128
            px = _f32(xy[perm[i], 0]) + _f32(x)
            py = _f32(xy[perm[i], 1]) + _f32(y)
            if int(px) < 0 or int(px) > cols or int(py) < 0 or int(py) > rows:    # :136
                continue
            if mask[int(py), int(px)] > 127:                                     # :141
                continue
            out_xy.append((float(px), float(py)))
            out_resp.append(float(resp[perm[i]]))
    if trace is not None:
        trace.setdefault("selected", []).append(np.array(out_xy, np.float32).reshape(-1, 2))
    if not out_xy:                                                               # :155-156
        return np.zeros((0, 2), np.float32), np.zeros((0,), np.float32)
    pts = corner_subpix(img, np.array(out_xy, np.float32))                       # :158-172
    return pts, np.array(out_resp, np.float32)


_orb = None


def orb_compute(img: np.ndarray, pts: np.ndarray, resp: np.ndarray):
    """cv::ORB::create()->compute(img, keypoints, desc) (TrackDescriptor.cpp:368, :424) on FAST keypoints (size 7, angle -1,
    octave 0).  compute() drops keypoints within 31 px of the border and returns the survivors with their descriptors.
    Returns (pts f32 [m,2], resp f32 [m], desc u8 [m,32])."""
    global _orb
    if _orb is None:
        _orb = cv2.ORB_create()
    kps = [cv2.KeyPoint(float(p[0]), float(p[1]), 7.0, -1.0, float(r), 0, -1) for p, r in zip(pts, resp)]
    if not kps:
        return np.zeros((0, 2), np.float32), np.zeros((0,), np.float32), np.zeros((0, 32), np.uint8)
    kps, desc = _orb.compute(img, kps)
    if desc is None or len(kps) == 0:
        return np.zeros((0, 2), np.float32), np.zeros((0,), np.float32), np.zeros((0, 32), np.uint8)
    return (np.array([k.pt for k in kps], np.float32).reshape(-1, 2), np.array([k.response for k in kps], np.float32),
            np.ascontiguousarray(desc))


_matcher = None


def knn_match(dq: np.ndarray, dt: np.ndarray) -> List[List[Match]]:
    """matcher->knnMatch(desc_q, desc_t, matches, 2) (TrackDescriptor.cpp:494-495)."""
    global _matcher
    if _matcher is None:
        _matcher = cv2.DescriptorMatcher_create("BruteForce-Hamming")
    if len(dq) == 0:
        return []
    if len(dt) == 0:
        return [[] for _ in range(len(dq))]
    res = _matcher.knnMatch(np.ascontiguousarray(dq), np.ascontiguousarray(dt), 2)
    return [[(m.queryIdx, m.trainIdx, float(m.distance)) for m in mm] for mm in res]


def _ratio_test(matches: List[List[Match]], knn_ratio: float) -> None:
    """robust_ratio_test (TrackDescriptor.cpp:542-556): float / float compared with the double ratio."""
    for k, match in enumerate(matches):
        if len(match) > 1:
            with np.errstate(divide="ignore", invalid="ignore"):
                if float(_f32(match[0][2]) / _f32(match[1][2])) > knn_ratio:
                    matches[k] = []
        else:
            matches[k] = []


def _symmetry_test(m1: List[List[Match]], m2: List[List[Match]]) -> List[Match]:
    """robust_symmetry_test (TrackDescriptor.cpp:558-575)."""
    good: List[Match] = []
    for a in m1:
        if len(a) < 2:
            continue
        for b in m2:
            if len(b) < 2:
                continue
            if a[0][0] == b[0][1] and b[0][0] == a[0][1]:
                good.append((a[0][0], a[0][1], a[0][2]))
                break
    return good


class TrackDescriptorOracle:
    """Restatement of ov_core::TrackDescriptor (+ the TrackBase state it uses)."""

    def __init__(self, cams: Dict[int, Cam], numfeats: int, numaruco: int, stereo: bool, histmethod: int, fast_threshold: int,
                 gridx: int, gridy: int, minpxdist: int, knnratio: float):
        self.camera_calib = cams
        self.database = FeatureDatabaseOracle()
        self.num_features = int(numfeats)
        self.use_stereo = bool(stereo)
        self.histogram_method = int(histmethod)
        self.currid = 4 * int(numaruco) + 1                         # TrackBase.cpp:34
        self.threshold = int(fast_threshold)
        self.grid_x, self.grid_y = int(gridx), int(gridy)
        self.min_px_dist = int(minpxdist)
        self.knn_ratio = float(knnratio)
        self.img_last: Dict[int, np.ndarray] = {}
        self.img_mask_last: Dict[int, np.ndarray] = {}
        self.pts_last: Dict[int, np.ndarray] = {}
        self.ids_last: Dict[int, List[int]] = {}
        self.desc_last: Dict[int, np.ndarray] = {}
        self.trace: Optional[dict] = None
        self.ransac_sizes: List[int] = []          # symmetric matches handed to every robust_match so far, in call order

    def get_feature_database(self):
        return self.database

    def get_last_obs(self):
        return {c: [(_f32(p[0]), _f32(p[1])) for p in v] for c, v in self.pts_last.items()}

    def get_last_ids(self):
        return {c: list(i) for c, i in self.ids_last.items()}

    def set_num_features(self, n: int):
        self.num_features = int(n)

    # -- feed_new_camera (TrackDescriptor.cpp:33-60) ------------------------------------------------------------------
    def feed_new_camera(self, timestamp: float, sensor_ids: Sequence[int], images: Sequence[np.ndarray], masks: Sequence[np.ndarray]):
        if len(sensor_ids) == 0 or len(sensor_ids) != len(images) or len(images) != len(masks):
            raise SystemExit("[ERROR]: MESSAGE DATA SIZES DO NOT MATCH OR EMPTY!!!")      # :36-42
        n = len(images)
        if n == 1:
            self._feed_monocular(timestamp, sensor_ids, images, masks, 0)
        elif n == 2 and self.use_stereo:
            self._feed_stereo(timestamp, sensor_ids, images, masks, 0, 1)
        elif not self.use_stereo:
            for i in range(n):
                self._feed_monocular(timestamp, sensor_ids, images, masks, i)
        else:
            raise SystemExit("[ERROR]: invalid number of images passed")                  # :57-58

    # -- feed_monocular (:62-174) ---------------------------------------------------------------------------------------
    def _feed_monocular(self, timestamp, sensor_ids, images, masks, msg_id):
        cam_id = sensor_ids[msg_id]
        img = equalize(images[msg_id], self.histogram_method)                             # :72-82
        mask = masks[msg_id]
        if cam_id not in self.pts_last or len(self.pts_last[cam_id]) == 0:                # :86-98
            pts, desc, ids = self._detect_mono(img, mask)
            self._store(cam_id, img, mask, pts, ids, desc)
            return
        pts_new, desc_new, ids_new = self._detect_mono(img, mask)                         # :106
        matches_ll = self._robust_match(self.pts_last[cam_id], pts_new, self.desc_last[cam_id], desc_new, cam_id, cam_id)   # :113
        good_ids = []
        for i in range(len(pts_new)):                                                     # :127-148
            idll = -1
            for (q, t, _d) in matches_ll:
                if t == i:
                    idll = q
            good_ids.append(self.ids_last[cam_id][idll] if idll != -1 else ids_new[i])
        npt = self.camera_calib[cam_id].undistort(pts_new)                                # :152-155
        for i in range(len(pts_new)):
            self.database.update_feature(good_ids[i], timestamp, cam_id, pts_new[i][0], pts_new[i][1], npt[i][0], npt[i][1])
        self._store(cam_id, img, mask, pts_new, good_ids, desc_new)                       # :161-168

    def _store(self, cam_id, img, mask, pts, ids, desc):
        self.img_last[cam_id] = img
        self.img_mask_last[cam_id] = mask
        self.pts_last[cam_id] = np.asarray(pts, np.float32).reshape(-1, 2)
        self.ids_last[cam_id] = list(ids)
        self.desc_last[cam_id] = np.asarray(desc, np.uint8).reshape(-1, 32)

    # -- feed_stereo (:176-356) -------------------------------------------------------------------------------------------
    def _feed_stereo(self, timestamp, sensor_ids, images, masks, ml, mr):
        cl, cr = sensor_ids[ml], sensor_ids[mr]
        img_l, img_r = equalize(images[ml], self.histogram_method), equalize(images[mr], self.histogram_method)   # :187-203
        mask_l, mask_r = masks[ml], masks[mr]
        if len(self.pts_last.get(cl, ())) == 0 or len(self.pts_last.get(cr, ())) == 0:    # :208-227
            p0, p1, d0, d1, i0, i1 = self._detect_stereo(img_l, img_r, mask_l, mask_r, cl, cr)
            self._store(cl, img_l, mask_l, p0, i0, d0)
            self._store(cr, img_r, mask_r, p1, i1, d1)
            return
        p0, p1, d0, d1, i0, i1 = self._detect_stereo(img_l, img_r, mask_l, mask_r, cl, cr)   # :235
        m_ll = self._robust_match(self.pts_last[cl], p0, self.desc_last[cl], d0, cl, cl)  # :241-250
        m_rr = self._robust_match(self.pts_last[cr], p1, self.desc_last[cr], d1, cr, cr)
        assert len(self.pts_last[cl]) == len(self.pts_last[cr]) and len(p0) == len(p1)    # :259-260
        if self.trace is not None:
            self.trace["m_ll"], self.trace["m_rr"] = list(m_ll), list(m_rr)
        good_l, good_r = [], []
        for i in range(len(p0)):                                                          # :268-307
            idll = idrr = -1
            for (q, t, _d) in m_ll:
                if t == i:
                    idll = q
            for (q, t, _d) in m_rr:
                if t == i:
                    idrr = q
            if idll != -1 and idrr != -1 and self.ids_last[cl][idll] == self.ids_last[cr][idrr]:
                good_l.append(self.ids_last[cl][idll])
                good_r.append(self.ids_last[cr][idrr])
            else:
                good_l.append(i0[i])
                good_r.append(i0[i])                                                      # :305 (ids_left_new for both)
        nl = self.camera_calib[cl].undistort(p0)                                          # :314-326
        nr = self.camera_calib[cr].undistort(p1)
        for i in range(len(p0)):
            self.database.update_feature(good_l[i], timestamp, cl, p0[i][0], p0[i][1], nl[i][0], nl[i][1])
            self.database.update_feature(good_l[i], timestamp, cr, p1[i][0], p1[i][1], nr[i][0], nr[i][1])
        self._store(cl, img_l, mask_l, p0, good_l, d0)                                    # :333-345
        self._store(cr, img_r, mask_r, p1, good_r, d1)

    # -- perform_detection_monocular (:358-405) -----------------------------------------------------------------------------
    def _detect_mono(self, img0, mask0):
        pts_ext, resp = perform_griding_fast(img0, mask0, self.num_features, self.grid_x, self.grid_y, self.threshold, self.trace)   # :365
        pts_ext, resp, desc_ext = orb_compute(img0, pts_ext, resp)                        # :368-369
        if self.trace is not None:
            self.trace.setdefault("ext", []).append((pts_ext.copy(), desc_ext.copy()))
        rows, cols = img0.shape
        mpd = _f32(self.min_px_dist)
        gw, gh = int(_f32(cols) / mpd), int(_f32(rows) / mpd)                             # :374
        grid = np.zeros((gh, gw), np.uint8)
        pts, desc, ids = [], [], []
        for i in range(len(pts_ext)):                                                     # :382-404
            px, py = _f32(pts_ext[i][0]), _f32(pts_ext[i][1])
            x, y = int(px), int(py)
            xg, yg = int(px / mpd), int(py / mpd)
            if xg < 0 or xg >= gw or yg < 0 or yg >= gh or x < 0 or x >= cols or y < 0 or y >= rows:
                continue
            if grid[yg, xg] > 127:
                continue
            pts.append(pts_ext[i])
            desc.append(desc_ext[i])
            self.currid += 1
            ids.append(self.currid)
            grid[yg, xg] = 255
        return np.array(pts, np.float32).reshape(-1, 2), np.array(desc, np.uint8).reshape(-1, 32), ids

    # -- perform_detection_stereo (:407-485) ----------------------------------------------------------------------------------
    def _detect_stereo(self, img0, img1, mask0, mask1, cam_id0, cam_id1):
        p0e, r0 = perform_griding_fast(img0, mask0, self.num_features, self.grid_x, self.grid_y, self.threshold, self.trace)   # :419-427
        p0e, r0, d0e = orb_compute(img0, p0e, r0)
        p1e, r1 = perform_griding_fast(img1, mask1, self.num_features, self.grid_x, self.grid_y, self.threshold, self.trace)
        p1e, r1, d1e = orb_compute(img1, p1e, r1)
        matches = self._robust_match(p0e, p1e, d0e, d1e, cam_id0, cam_id1)                # :431
        if self.trace is not None:
            self.trace["ext"] = [(p0e.copy(), d0e.copy()), (p1e.copy(), d1e.copy())]
            self.trace["m_lr"] = list(matches)
        mpd = _f32(self.min_px_dist)
        rows0, cols0 = img0.shape
        rows1, cols1 = img1.shape
        gw0, gh0 = int(_f32(cols0) / mpd), int(_f32(rows0) / mpd)                         # :436-439
        gw1, gh1 = int(_f32(cols1) / mpd), int(_f32(rows1) / mpd)
        # NOTE the reference allocates both grids and tests them (:469) but never marks a cell: the stereo path has no
        # minimum-distance rejection.  Restated as written.
        pts0, pts1, desc0, desc1, ids0, ids1 = [], [], [], [], [], []
        for (q, t, _d) in matches:                                                        # :442-484
            a, b = p0e[q], p1e[t]
            x0, y0 = int(_f32(a[0])), int(_f32(a[1]))
            x0g, y0g = int(_f32(a[0]) / mpd), int(_f32(a[1]) / mpd)
            if x0g < 0 or x0g >= gw0 or y0g < 0 or y0g >= gh0 or x0 < 0 or x0 >= cols0 or y0 < 0 or y0 >= rows0:
                continue
            x1, y1 = int(_f32(b[0])), int(_f32(b[1]))
            x1g, y1g = int(_f32(b[0]) / mpd), int(_f32(b[1]) / mpd)
            if x1g < 0 or x1g >= gw1 or y1g < 0 or y1g >= gh1 or x1 < 0 or x1 >= cols0 or y1 < 0 or y1 >= rows0:   # :462-463 (img0's size, as written)
                continue
            pts0.append(a)
            pts1.append(b)
            desc0.append(d0e[q])
            desc1.append(d1e[t])
            self.currid += 1                                                              # :481-483
            ids0.append(self.currid)
            ids1.append(self.currid)
        f = lambda v, w: np.array(v, np.float32).reshape(-1, w) if w == 2 else np.array(v, np.uint8).reshape(-1, 32)
        return f(pts0, 2), f(pts1, 2), f(desc0, 32), f(desc1, 32), ids0, ids1

    # -- robust_match (:487-540) ------------------------------------------------------------------------------------------------
    def _robust_match(self, pts0, pts1, desc0, desc1, id0, id1) -> List[Match]:
        m01 = knn_match(desc0, desc1)                                                     # :494-495
        m10 = knn_match(desc1, desc0)
        _ratio_test(m01, self.knn_ratio)                                                  # :498-499
        _ratio_test(m10, self.knn_ratio)
        good = _symmetry_test(m01, m10)                                                   # :502-503
        self.ransac_sizes.append(len(good))        # (test aid: 10..14 correspondences send cv::findFundamentalMat to LMedS)
        if len(good) < 10:                                                                # :517-518
            return []
        a = np.array([pts0[q] for (q, _t, _d) in good], np.float32)
        b = np.array([pts1[t] for (_q, t, _d) in good], np.float32)
        n0 = self.camera_calib[id0].undistort(a)                                          # :522-526
        n1 = self.camera_calib[id1].undistort(b)
        fmax = max(self.camera_calib[id0].max_focal(), self.camera_calib[id1].max_focal())   # :530-533
        _F, mask = cv2.findFundamentalMat(n0.reshape(-1, 1, 2), n1.reshape(-1, 1, 2), cv2.FM_RANSAC, 1.0 / fmax, 0.999)
        if mask is None:
            # the reference indexes an empty mask_rsc here (:536-539, undefined behaviour); no match survives in this restatement
            return []
        mask = mask.reshape(-1)
        return [g for g, m in zip(good, mask) if m == 1]


def make_descriptor_oracle(cfg, knn_ratio: float = 0.70, **over) -> TrackDescriptorOracle:
    cams = {i: Cam(c.model, tuple(c.intrinsics), tuple(c.dist)) for i, c in enumerate(cfg.cams)}
    kw = dict(numfeats=cfg.num_features, numaruco=cfg.num_aruco, stereo=cfg.use_stereo, histmethod=cfg.histogram,
              fast_threshold=cfg.fast_threshold, gridx=cfg.grid_x, gridy=cfg.grid_y, minpxdist=cfg.min_px_dist, knnratio=knn_ratio)
    kw.update(over)
    return TrackDescriptorOracle(cams, **kw)
