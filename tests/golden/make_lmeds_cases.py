"""Generates tests/golden/lmeds_cases.npz: normalised correspondences with N = 10..13 points (cv::findFundamentalMat's LMedS
regime, reached from TrackKLT.cpp:848-873 when a camera holds that few tracks) and the masks cv2 4.13 (IPP off) returns
for them on the generating machine.  The masks are kept for the record and to validate the checker
(util.lmeds_mask_is_valid_outcome): they are NOT expected to be reproduced (profiles/r2_lmeds_probe.json).

    python tests/golden/make_lmeds_cases.py
"""
import os
import sys

import cv2
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(HERE)), "tools"))
import lmeds_probe  # noqa: E402

cv2.ipp.setUseIPP(False)
PER_N = 6
out = {"meta": np.array(str(dict(cv2_version=cv2.__version__, ipp=False, per_n=PER_N)))}
cs = lmeds_probe.cases(PER_N, seed=11)
k = 0
for j, n in enumerate(lmeds_probe.NS):
    if n > 13:
        break
    for i in range(j * PER_N, (j + 1) * PER_N):
        p0, p1 = cs[i]
        _F, m = cv2.findFundamentalMat(p0.reshape(-1, 1, 2), p1.reshape(-1, 1, 2), cv2.FM_RANSAC, 2.0 / 460.0, 0.999)
        out[f"p0_{k}"], out[f"p1_{k}"], out[f"cv2_mask_{k}"] = p0, p1, m.reshape(-1).astype(np.uint8)
        k += 1
out["count"] = k
np.savez_compressed(os.path.join(HERE, "lmeds_cases.npz"), **out)
print(k, "cases")
