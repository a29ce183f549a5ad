"""How often does TrackKLT reach cv::findFundamentalMat's least-median regime (10 <= N <= 14 correspondences, decided by
rounding noise for N <= 13: DESIGN.md section 5)?  Runs the oracle over 200-frame synthetic sequences (events on: jumps and
blank frames) of the five BASELINE configurations and over a num_features sweep, and counts the perform_matching calls
(TrackKLT.cpp:829-879) by the number of points they hand to RANSAC.

    python tools/lmeds_census.py [frames]      -> one JSON line (committed as profiles/r2_lmeds_census.json)
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import cv2  # noqa: E402
import numpy as np  # noqa: E402

from uvio_b200 import configs  # noqa: E402
from util import frames_for, masks_for  # noqa: E402
from oracle import trackklt_cv2 as T  # noqa: E402

cv2.ipp.setUseIPP(False)


def census(cfg, n_frames, stream_id=0):
    counts = {"calls": 0, "lt10": 0, "n10_13": 0, "n14": 0, "ge15": 0, "raised": 0}
    real = T.fundamental_ransac_mask

    def counting(p0n, p1n, thr):
        n = len(p0n)
        counts["calls"] += 1
        counts["n10_13" if 10 <= n <= 13 else "n14" if n == 14 else "ge15" if n >= 15 else "lt10"] += 1
        try:
            return real(p0n, p1n, thr)
        except cv2.error:                      # every correspondence identical: the reference would terminate here
            counts["raised"] += 1
            return np.ones(n, np.uint8)
    T.fundamental_ransac_mask = counting
    try:
        o = T.make_oracle(cfg)
        masks = masks_for(cfg)
        frames = frames_for(cfg, n_frames, stream_id=stream_id, events=True)
        for k, (l, r) in enumerate(frames):
            o.feed_new_camera(0.05 * k, [0, 1], [l, r], masks)
    finally:
        T.fundamental_ransac_mask = real
    return counts


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 200
    out = {"frames_per_sequence": n, "events": True, "sequences": {}, "num_features_sweep": {}}
    for cfg in (configs.EUROC_MAV, configs.TUM_VI, configs.RS_D455, configs.HD_1000, configs.UZHFPV_INDOOR):
        out["sequences"][cfg.name] = census(cfg, n)
    for nf in (12, 16, 24):
        out["num_features_sweep"][str(nf)] = census(configs.EUROC_MAV.with_(num_features=nf, grid_x=2, grid_y=2), n, stream_id=51)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
