"""Latency of one undistort + fundamental-matrix RANSAC job (uvio_klt_stage_ransac) on synthetic correspondences: median
wall-clock per call (launch + kernel + sync) for a few inlier ratios.  Development probe for A/B builds (UVIO_KLT_LIB)."""
import json
import sys
import time
import os

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from uvio_b200 import configs
from uvio_b200.track import make_tracker, CameraData

cfg = configs.EUROC_MAV
t = make_tracker(cfg, device=0)
z = np.zeros((cfg.height, cfg.width), np.uint8)
img = (np.random.default_rng(0).integers(0, 255, (cfg.height, cfg.width))).astype(np.uint8)
t.feed_new_camera(CameraData(0.0, [0, 1], [img, img], [z, z]))      # gives the cameras their device state
out = {}
rng = np.random.default_rng(1)
for ratio in (0.95, 0.8, 0.6):
    n = 200
    p0 = np.c_[rng.uniform(40, 700, n), rng.uniform(40, 440, n)].astype(np.float32)
    depth = rng.uniform(2, 10, n)
    p1 = p0 + np.c_[30.0 / depth, np.zeros(n)].astype(np.float32) + rng.normal(0, 0.2, (n, 2)).astype(np.float32)
    bad = rng.random(n) > ratio
    p1[bad] += rng.uniform(-40, 40, (int(bad.sum()), 2)).astype(np.float32)
    p1 = p1.astype(np.float32)
    for _ in range(20):
        mask, _a, _b = t.stage_ransac(0, 0, p0, p1)
    ts = []
    for _ in range(200):
        t0 = time.perf_counter()
        mask, _a, _b = t.stage_ransac(0, 0, p0, p1)
        ts.append(1e6 * (time.perf_counter() - t0))
    out[f"inliers_{ratio}"] = {"us_median": round(float(np.median(ts)), 1), "us_p10": round(float(np.percentile(ts, 10)), 1),
                               "mask_sum": int(mask.sum()), "mask_crc": int(np.dot(mask.astype(np.int64), np.arange(n) + 1))}
print(json.dumps(out))
