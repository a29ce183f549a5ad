"""GPU parity, stage by stage, through the C ABI: CUDA kernels vs OpenCV (cv2 4.13, IPP off) on the same inputs.
Integer stages must be bit-exact; LK / cornerSubPix within 0.01 px with identical status (they come out bit-exact)."""
import cv2
import numpy as np
import pytest

from uvio_b200 import configs
from uvio_b200.track import CameraData, make_tracker

from util import frames_for, masks_for

cv2.ipp.setUseIPP(False)
pytestmark = pytest.mark.gpu

CFGS = [configs.EUROC_MAV, configs.TUM_VI, configs.RS_D455, configs.UZHFPV_INDOOR,
        configs.EUROC_MAV.with_(name="euroc_none", histogram=configs.HIST_NONE),
        configs.EUROC_MAV.with_(name="euroc_pyr3", pyr_levels=3), configs.HD_1000, configs.HD_1000_PYR5]


def _equalize(img, method):
    if method == configs.HIST_HISTOGRAM:
        return cv2.equalizeHist(img)
    if method == configs.HIST_CLAHE:
        return cv2.createCLAHE(10.0, (8, 8)).apply(img)
    return img


def _feed(trk, cfg, frame, k=0):
    trk.feed_new_camera(CameraData(0.05 * k, [0, 1], [frame[0], frame[1]], masks_for(cfg)))


@pytest.mark.parametrize("cfg", CFGS, ids=lambda c: c.name)
def test_equalize_and_pyramid_bit_exact(built_lib, cfg):
    frames = frames_for(cfg, 2)
    trk = make_tracker(cfg)
    _feed(trk, cfg, frames[0], 0)
    _feed(trk, cfg, frames[1], 1)
    for cam in (0, 1):
        for slot, fr in ((0, frames[1]), (1, frames[0])):
            ref = _equalize(fr[cam], cfg.histogram)
            max_level, pyr = cv2.buildOpticalFlowPyramid(ref, (15, 15), cfg.pyr_levels)
            assert trk.stage_num_levels(cam, slot) == max_level + 1
            for lv in range(max_level + 1):
                got = trk.stage_level(cam, slot, lv)
                assert got.shape == ref.shape
                assert np.array_equal(got, ref), f"{cfg.name} cam {cam} slot {slot} level {lv}"
                ref = cv2.pyrDown(ref)


def test_equalize_degenerate_images(built_lib):
    cfg = configs.EUROC_MAV
    trk = make_tracker(cfg)
    const = np.full((cfg.height, cfg.width), 77, np.uint8)
    two = const.copy()
    two[:10] = 3
    _feed(trk, cfg, (const, two))
    assert np.array_equal(trk.stage_level(0, 0, 0), cv2.equalizeHist(const))
    assert np.array_equal(trk.stage_level(1, 0, 0), cv2.equalizeHist(two))


@pytest.mark.parametrize("cfg", [configs.EUROC_MAV, configs.RS_D455, configs.HD_1000], ids=lambda c: c.name)
def test_fast_cells_bit_exact(built_lib, cfg):
    frames = frames_for(cfg, 1)
    trk = make_tracker(cfg)
    _feed(trk, cfg, frames[0])
    img = trk.stage_level(0, 0, 0)
    sx, sy = cfg.width // cfg.grid_x, cfg.height // cfg.grid_y
    det = cv2.FastFeatureDetector_create(cfg.fast_threshold, True)
    total = 0
    for gx in range(cfg.grid_x):
        for gy in range(cfg.grid_y):
            kps = det.detect(np.ascontiguousarray(img[gy * sy:(gy + 1) * sy, gx * sx:(gx + 1) * sx]))
            ref_xy = np.array([[int(k.pt[0]), int(k.pt[1])] for k in kps], np.int32).reshape(-1, 2)
            ref_r = np.array([k.response for k in kps], np.float32)
            xy, r = trk.stage_fast(0, 0, gx * sx, gy * sy, sx, sy, cfg.fast_threshold)
            assert np.array_equal(xy, ref_xy) and np.array_equal(r, ref_r), f"cell {gx},{gy}"
            total += len(kps)
    assert total > 100
    # odd rectangle, whole image
    kps = det.detect(img)
    xy, r = trk.stage_fast(0, 0, 0, 0, cfg.width, cfg.height, cfg.fast_threshold)
    assert np.array_equal(xy, np.array([[int(k.pt[0]), int(k.pt[1])] for k in kps], np.int32).reshape(-1, 2))


def _corner_points(img, thr, step=3):
    kps = cv2.FastFeatureDetector_create(thr, True).detect(img)
    pts = np.array([k.pt for k in kps], np.float32)[::step]
    h, w = img.shape
    border = np.array([[2, 3], [w - 3, h - 3], [5.5, h / 2], [w / 2, 1], [w - 1, 7], [0, 0], [6, 6], [w - 8, h - 8]], np.float32)
    return np.vstack([pts, border]).astype(np.float32)


@pytest.mark.parametrize("cfg", [configs.EUROC_MAV, configs.TUM_VI], ids=lambda c: c.name)
def test_corner_subpix(built_lib, cfg):
    frames = frames_for(cfg, 1)
    trk = make_tracker(cfg)
    _feed(trk, cfg, frames[0])
    img = trk.stage_level(0, 0, 0)
    pts = _corner_points(img, cfg.fast_threshold)
    ref = pts.reshape(-1, 1, 2).copy()
    cv2.cornerSubPix(img, ref, (5, 5), (-1, -1), (cv2.TERM_CRITERIA_COUNT + cv2.TERM_CRITERIA_EPS, 20, 0.001))
    got = trk.stage_subpix(0, 0, pts)
    err = np.abs(got - ref.reshape(-1, 2)).max()
    assert err <= 0.01, err                     # north_star tolerance
    assert np.array_equal(got, ref.reshape(-1, 2)), f"not bit-exact: max err {err}"


@pytest.mark.parametrize("cfg", [configs.EUROC_MAV, configs.TUM_VI, configs.EUROC_MAV.with_(name="euroc_pyr3", pyr_levels=3)],
                         ids=lambda c: c.name)
def test_lk_temporal_and_stereo(built_lib, cfg):
    frames = frames_for(cfg, 2, events=False, start=7)
    trk = make_tracker(cfg)
    _feed(trk, cfg, frames[0], 0)
    _feed(trk, cfg, frames[1], 1)
    prev_l, cur_l, cur_r = trk.stage_level(0, 1, 0), trk.stage_level(0, 0, 0), trk.stage_level(1, 0, 0)
    pts = _corner_points(prev_l, cfg.fast_threshold, step=2)
    w, h = cfg.width, cfg.height
    pts = np.vstack([pts, [[-3, 50], [w + 8, 100], [w / 2, -5], [w / 2, h + 10]]]).astype(np.float32)
    crit = (cv2.TERM_CRITERIA_COUNT | cv2.TERM_CRITERIA_EPS, 30, 0.01)
    for (a, b, sa, sb, ca, cb) in ((prev_l, cur_l, 1, 0, 0, 0), (cur_l, cur_r, 0, 0, 0, 1)):
        ref, st, _ = cv2.calcOpticalFlowPyrLK(a, b, pts.reshape(-1, 1, 2), pts.reshape(-1, 1, 2).copy(), winSize=(15, 15),
                                              maxLevel=cfg.pyr_levels, criteria=crit, flags=cv2.OPTFLOW_USE_INITIAL_FLOW)
        got, gst = trk.stage_lk(ca, sa, cb, sb, pts, pts)
        st = st.reshape(-1)
        assert np.array_equal(gst, st), f"status differs at {np.nonzero(gst != st)[0]}"
        assert st.sum() > 50
        err = np.abs(got - ref.reshape(-1, 2))[st == 1].max()
        assert err <= 0.01, err                 # north_star tolerance
        assert np.array_equal(got[st == 1], ref.reshape(-1, 2)[st == 1]), f"not bit-exact: max err {err}"


def _two_view(N, out_frac, seed, f, cx, cy, noise=0.3):
    r = np.random.default_rng(seed)
    X = np.c_[r.uniform(-3, 3, N), r.uniform(-2, 2, N), r.uniform(3, 10, N)]
    R = cv2.Rodrigues(r.normal(0, 0.02, 3))[0]
    t = r.normal(0, 0.1, 3)
    X2 = X @ R.T + t
    a = X[:, :2] / X[:, 2:] * f + [cx, cy]
    b = X2[:, :2] / X2[:, 2:] * f + [cx, cy]
    a += r.normal(0, noise, a.shape)
    b += r.normal(0, noise, b.shape)
    k = int(out_frac * N)
    b[:k] += r.normal(0, 30, (k, 2))
    return a.astype(np.float32), b.astype(np.float32)


@pytest.mark.parametrize("cfg", [configs.EUROC_MAV, configs.TUM_VI], ids=lambda c: c.name)
def test_undistort_and_ransac_masks(built_lib, cfg):
    from oracle.trackklt_cv2 import Cam
    trk = make_tracker(cfg)
    cams = [Cam(c.model, tuple(c.intrinsics), tuple(c.dist)) for c in cfg.cams]
    fx, fy, cx, cy = cfg.cams[0].intrinsics
    bad = 0
    for seed in range(60):
        r = np.random.default_rng(seed + 5000)
        N = int(r.choice([14, 15, 16, 20, 50, 100, 200, 400, 1000]))
        a, b = _two_view(N, r.choice([0, 0.05, 0.1, 0.3, 0.5]), seed, 0.5 * fx, cx, cy, r.choice([0.1, 0.3, 1.0]))
        an, bn = cams[0].undistort(a), cams[1].undistort(b)
        thr = 2.0 / max(cams[0].max_focal(), cams[1].max_focal())
        _F, m = cv2.findFundamentalMat(an.reshape(-1, 1, 2), bn.reshape(-1, 1, 2), cv2.FM_RANSAC, thr, 0.999)
        m = np.zeros(N, np.uint8) if m is None else m.reshape(-1)
        mask, p0n, p1n = trk.stage_ransac(0, 1, a, b)
        assert np.abs(p0n - an).max() <= 1e-6 and np.abs(p1n - bn).max() <= 1e-6
        if cfg.cams[0].model == configs.CAM_RADTAN:
            assert np.array_equal(p0n, an) and np.array_equal(p1n, bn)
        bad += int(not np.array_equal(mask, m))
    assert bad == 0, f"{bad}/60 inlier masks differ from cv2"
