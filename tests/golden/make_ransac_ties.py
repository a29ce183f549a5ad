"""Generator of tests/golden/ransac_ties.npz: two-view point sets on which two models of ONE 7-point sample reach the same
best inlier count, so that cv::findFundamentalMat's answer depends on the order run7Point returns its roots in (Hartley
normalisation + the null-space basis of OpenCV's Jacobi SVD).  The point sets were captured from a 161x127 stereo sequence
(tests/test_gpu_edge.py, odd_161x127, frames 3 and 9, camera 0); this script undistorts them and asks cv2 for the masks.

    python tests/golden/make_ransac_ties.py            # rewrites the .npz next to it (cv2 4.13.0, IPP off)
"""
import os
import sys

import cv2
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from oracle.trackklt_cv2 import Cam            # noqa: E402
from uvio_b200 import configs                  # noqa: E402

W, H = 161, 127


def camera():
    c = configs.EUROC_MAV.cams[0]
    return (c.model, (c.intrinsics[0] * W / 752, c.intrinsics[1] * H / 480, W / 2.0 + 1.5, H / 2.0 - 2.0), tuple(c.dist))


def main():
    cv2.ipp.setUseIPP(False)
    src = np.load(os.path.join(HERE, "ransac_ties.npz"))
    model, intr, dist = camera()
    cam = Cam(model, intr, dist)
    out = {"intrinsics": np.array(intr), "dist": np.array(dist), "model": np.array(model), "cv2_version": np.array(cv2.__version__)}
    for k in ("a", "b"):
        p0, p1 = src[f"p0_{k}"], src[f"p1_{k}"]
        _F, m = cv2.findFundamentalMat(cam.undistort(p0).reshape(-1, 1, 2), cam.undistort(p1).reshape(-1, 1, 2), cv2.FM_RANSAC,
                                       2.0 / cam.max_focal(), 0.999)
        out[f"p0_{k}"], out[f"p1_{k}"], out[f"mask_{k}"] = p0, p1, m.reshape(-1).astype(np.uint8)
    np.savez_compressed(os.path.join(HERE, "ransac_ties.npz"), **out)


if __name__ == "__main__":
    main()
