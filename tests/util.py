"""Shared helpers of the parity tests (test infrastructure; the only place besides bench/smoke that touches oracle/)."""
import numpy as np

from uvio_b200 import configs
from uvio_b200.synth import SynthStereoStream, circular_mask


def frames_for(cfg, n, stream_id=0, events=True, start=0):
    s = SynthStereoStream(cfg.width, cfg.height, stream_id, events=events)
    return [s.frame(start + k) for k in range(n)]


def masks_for(cfg):
    if cfg.use_mask:
        m = circular_mask(cfg.width, cfg.height)
        return [m, m.copy()]
    z = np.zeros((cfg.height, cfg.width), np.uint8)
    return [z, z.copy()]


def run_oracle(cfg, frames, masks, mono=False, num_features=None, set_nf_at=None):
    """Feeds the cv2-driven oracle; returns (oracle, per-frame list of (log slice, pts_last, ids_last))."""
    from oracle.trackklt_cv2 import make_oracle
    o = make_oracle(cfg) if num_features is None else make_oracle(cfg, numfeats=num_features)
    out = []
    for k, (l, r) in enumerate(frames):
        if set_nf_at and k in set_nf_at:
            o.set_num_features(set_nf_at[k])
        n0 = len(o.database.log)
        if mono:
            o.feed_new_camera(float(k) * 0.05, [0], [l], [masks[0]])
        else:
            o.feed_new_camera(float(k) * 0.05, [0, 1], [l, r], masks)
        out.append((list(o.database.log[n0:]), {c: list(p) for c, p in o.pts_last.items()},
                    {c: list(i) for c, i in o.ids_last.items()}))
    return o, out


def log_arrays(log):
    """update_feature log -> (ids, cams, uv, uvn)"""
    if not log:
        return (np.zeros(0, np.int64), np.zeros(0, np.int64), np.zeros((0, 2), np.float32), np.zeros((0, 2), np.float32))
    ids = np.array([e[0] for e in log], np.int64)
    cams = np.array([e[2] for e in log], np.int64)
    uv = np.array([[e[3], e[4]] for e in log], np.float32)
    uvn = np.array([[e[5], e[6]] for e in log], np.float32)
    return ids, cams, uv, uvn


# ---- cv::findFundamentalMat in its least-median regime (10 <= N <= 13) ------------------------------------
def fm_errors(F, p0, p1):
    """FMEstimatorCallback::computeError (OpenCV fundam.cpp): max of the two squared point-to-epipolar-line distances."""
    F = np.asarray(F, np.float64).reshape(3, 3)
    x0, y0, x1, y1 = (np.asarray(v, np.float64) for v in (p0[:, 0], p0[:, 1], p1[:, 0], p1[:, 1]))
    a = F[0, 0] * x0 + F[0, 1] * y0 + F[0, 2]
    b = F[1, 0] * x0 + F[1, 1] * y0 + F[1, 2]
    c = F[2, 0] * x0 + F[2, 1] * y0 + F[2, 2]
    s2 = 1.0 / (a * a + b * b)
    d2 = x1 * a + y1 * b + c
    a = F[0, 0] * x1 + F[1, 0] * y1 + F[2, 0]
    b = F[0, 1] * x1 + F[1, 1] * y1 + F[2, 1]
    c = F[0, 2] * x1 + F[1, 2] * y1 + F[2, 2]
    s1 = 1.0 / (a * a + b * b)
    d1 = x0 * a + y0 * b + c
    return np.maximum(d1 * d1 * s1, d2 * d2 * s2)


def lmeds_mask_is_valid_outcome(p0n, p1n, mask):
    """With 10..13 correspondences OpenCV's LMedS takes the median error at an index that lies inside the 7 points every
    hypothesis fits exactly: all medians are rounding noise (~1e-30), the winner is whichever hypothesis has the smallest
    noise, sigma clamps to 0.001 and the mask is that hypothesis' inlier set at err <= 1e-6.  Which hypothesis wins cannot be
    reproduced (cv2's own mask changes in 60-80 % of cases when one input coordinate moves by one ulp:
    profiles/r2_lmeds_probe.json), but WHAT a result must look like can be checked: the mask is the exact inlier set, at that
    threshold, of a fundamental matrix fitted to 7 of its own inliers.  Returns (ok, reason)."""
    import itertools
    import cv2
    mask = np.asarray(mask).astype(bool)
    n = len(mask)
    idx = np.nonzero(mask)[0]
    if len(idx) < 7:
        return False, f"only {len(idx)} inliers: the 7 fitted points are always inliers of their own model"
    p0 = np.ascontiguousarray(p0n, np.float32)
    p1 = np.ascontiguousarray(p1n, np.float32)
    thr = float(np.float32(0.001 * 0.001))
    best = n
    for sub in itertools.combinations(idx.tolist(), 7):
        s = list(sub)
        F, _m = cv2.findFundamentalMat(p0[s].reshape(-1, 1, 2), p1[s].reshape(-1, 1, 2), cv2.FM_7POINT)
        if F is None:
            continue
        for k in range(F.shape[0] // 3):
            e = fm_errors(F[3 * k:3 * k + 3], p0, p1)
            if np.sort(e)[n // 2] > 1e-18:                 # not an exact-fit model at the median index: LMedS would not pick it
                continue
            inl = e <= thr
            sure = np.abs(e - thr) > 1e-6 * thr            # points whose side of the threshold does not hang on the last bits
            bad = int(np.count_nonzero((inl != mask) & sure))
            best = min(best, bad)
            if bad == 0:
                return True, "ok"
    return False, f"no 7-point model through the mask's own inliers reproduces it (closest differs in {best} points)"
