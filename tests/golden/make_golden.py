"""Generates the golden fixtures of tests/golden/ from the oracle (cv2-driven restatement of the reference, IPP off).

    python tests/golden/make_golden.py

The reference ships no golden vectors for this path, so these pin the parity target to OpenCV's
behaviour at the reference's call sites: cv2 version and flags are stored in every file.  Inputs are not stored: they are
regenerated from uvio_b200.synth by seed, and their checksum is stored to detect generator drift.
"""
import os
import sys

import cv2
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from uvio_b200 import configs  # noqa: E402
from uvio_b200.synth import frames_checksum  # noqa: E402
from util import frames_for, masks_for, run_oracle, log_arrays  # noqa: E402

cv2.ipp.setUseIPP(False)
META = dict(cv2_version=cv2.__version__, ipp=False)

SEQUENCES = [("euroc_mav", configs.EUROC_MAV, 24, 0, False), ("tum_vi", configs.TUM_VI, 16, 0, False),
             ("rs_d455", configs.RS_D455, 12, 0, False), ("uzhfpv_indoor", configs.UZHFPV_INDOOR, 12, 0, False),
             ("euroc_mono", configs.EUROC_MAV, 16, 3, True), ("hd_1000", configs.HD_1000, 12, 0, False)]


def sequence_fixture(name, cfg, n, stream_id, mono):
    frames = frames_for(cfg, n, stream_id=stream_id)
    masks = masks_for(cfg)
    _o, ref = run_oracle(cfg, frames, masks, mono=mono)
    out = dict(meta=np.array(str(META)), n_frames=n, stream_id=stream_id, mono=mono,
               frames_checksum=np.uint64(frames_checksum([f for lr in frames for f in lr])))
    for k, (log, pts_last, ids_last) in enumerate(ref):
        ids, cams, uv, uvn = log_arrays(log)
        out[f"ids_{k}"], out[f"cams_{k}"], out[f"uv_{k}"], out[f"uvn_{k}"] = ids, cams.astype(np.int8), uv, uvn
        for c, want in ids_last.items():
            out[f"last_ids_{k}_{c}"] = np.array(want, np.int64)
            out[f"last_uv_{k}_{c}"] = np.array(pts_last[c], np.float32).reshape(-1, 2)
    np.savez_compressed(os.path.join(HERE, f"seq_{name}.npz"), **out)
    print(name, n, "frames,", sum(len(r[0]) for r in ref), "observations")


def stage_fixture():
    """Stage-level vectors on frames 7/8 of euroc stream 0 (no events): FAST per cell, cornerSubPix, LK, undistort, RANSAC."""
    cfg = configs.EUROC_MAV
    frames = frames_for(cfg, 2, events=False, start=7)
    a, b = cv2.equalizeHist(frames[0][0]), cv2.equalizeHist(frames[1][0])
    out = dict(meta=np.array(str(META)), frames_checksum=np.uint64(frames_checksum([f for lr in frames for f in lr])))
    lv = a
    for k in range(5):
        out[f"pyr_sum_{k}"] = np.uint64(frames_checksum([lv]))
        lv = cv2.pyrDown(lv)
    out["clahe_sum"] = np.uint64(frames_checksum([cv2.createCLAHE(10.0, (8, 8)).apply(frames[0][0])]))
    det = cv2.FastFeatureDetector_create(cfg.fast_threshold, True)
    sx, sy = cfg.width // cfg.grid_x, cfg.height // cfg.grid_y
    for (gx, gy) in ((0, 0), (2, 3), (4, 4)):
        kps = det.detect(np.ascontiguousarray(a[gy * sy:(gy + 1) * sy, gx * sx:(gx + 1) * sx]))
        out[f"fast_xy_{gx}_{gy}"] = np.array([[int(k.pt[0]), int(k.pt[1])] for k in kps], np.int16).reshape(-1, 2)
        out[f"fast_resp_{gx}_{gy}"] = np.array([k.response for k in kps], np.uint8)
    kps = det.detect(a)
    pts = np.array([k.pt for k in kps], np.float32)[::9]
    h, w = a.shape
    pts = np.vstack([pts, [[2, 3], [w - 3, h - 3], [5.5, h / 2], [w / 2, 1], [-3, 50], [w + 8, 100]]]).astype(np.float32)
    inside = (pts[:, 0] >= 0) & (pts[:, 0] < w)
    sp = pts[inside].reshape(-1, 1, 2).copy()
    cv2.cornerSubPix(a, sp, (5, 5), (-1, -1), (cv2.TERM_CRITERIA_COUNT + cv2.TERM_CRITERIA_EPS, 20, 0.001))
    out["subpix_in"], out["subpix_out"] = pts[inside], sp.reshape(-1, 2)
    crit = (cv2.TERM_CRITERIA_COUNT | cv2.TERM_CRITERIA_EPS, 30, 0.01)
    nxt, st, _ = cv2.calcOpticalFlowPyrLK(a, b, pts.reshape(-1, 1, 2), pts.reshape(-1, 1, 2).copy(), winSize=(15, 15), maxLevel=5,
                                          criteria=crit, flags=cv2.OPTFLOW_USE_INITIAL_FLOW)
    out["lk_in"], out["lk_out"], out["lk_status"] = pts, nxt.reshape(-1, 2), st.reshape(-1).astype(np.uint8)
    from oracle.trackklt_cv2 import Cam
    for cname, c in (("radtan", configs.EUROC_MAV.cams[0]), ("equi", configs.TUM_VI.cams[0])):
        cam = Cam(c.model, tuple(c.intrinsics), tuple(c.dist))
        uv = (np.random.default_rng(5).random((300, 2)) * [500, 470]).astype(np.float32)
        out[f"undist_in_{cname}"], out[f"undist_out_{cname}"] = uv, cam.undistort(uv)
    # RANSAC on the LK result (tracked points only), as perform_matching does
    cam = Cam(cfg.cams[0].model, tuple(cfg.cams[0].intrinsics), tuple(cfg.cams[0].dist))
    good = st.reshape(-1) == 1
    p0, p1 = pts[good], nxt.reshape(-1, 2)[good]
    p1[:7] += 25.0                                   # a few gross outliers
    _F, m = cv2.findFundamentalMat(cam.undistort(p0).reshape(-1, 1, 2), cam.undistort(p1).reshape(-1, 1, 2), cv2.FM_RANSAC,
                                   2.0 / cam.max_focal(), 0.999)
    out["ransac_p0"], out["ransac_p1"], out["ransac_mask"] = p0, p1, m.reshape(-1).astype(np.uint8)
    np.savez_compressed(os.path.join(HERE, "stages_euroc.npz"), **out)
    print("stages:", len(pts), "points,", int(good.sum()), "tracked,", int(m.sum()), "inliers")


if __name__ == "__main__":
    only = sys.argv[1:]                      # e.g. `make_golden.py hd_1000` regenerates one sequence
    for args in SEQUENCES:
        if not only or args[0] in only:
            sequence_fixture(*args)
    if not only or "stages" in only:
        stage_fixture()
