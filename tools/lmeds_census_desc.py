"""How often the descriptor tracker's robust_match (TrackDescriptor.cpp:487-540) hands cv::findFundamentalMat 10..13
correspondences -- the LMedS regime whose outcome OpenCV itself does not reproduce -- on the synthetic
sequences of the BASELINE configurations.  CPU only (the cv2-driven oracle).

    python tools/lmeds_census_desc.py > profiles/r2_lmeds_census_desc.json
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import cv2

from oracle.trackdescriptor_cv2 import make_descriptor_oracle
from uvio_b200 import configs
from util import frames_for, masks_for

out = {}
for name in ("EUROC_MAV", "TUM_VI", "RS_D455", "HD_1000", "UZHFPV_INDOOR"):
    cfg = getattr(configs, name)
    n = 60 if name != "HD_1000" else 30
    frames = frames_for(cfg, n, stream_id=1)
    masks = masks_for(cfg)
    o = make_descriptor_oracle(cfg)
    raised = 0
    for k, (l, r) in enumerate(frames):
        try:
            o.feed_new_camera(0.05 * k, [0, 1], [l, r], masks)
        except cv2.error:
            raised += 1                      # identical correspondences: cv2 asserts (the reference would terminate)
    sizes = o.ransac_sizes
    stereo = sizes[0::3] if False else None
    out[name] = {"frames": n, "robust_match_calls": len(sizes), "calls_with_fewer_than_10": sum(s < 10 for s in sizes),
                 "calls_in_lmeds_regime_10_to_13": sum(10 <= s <= 13 for s in sizes), "calls_with_14": sum(s == 14 for s in sizes),
                 "median_symmetric_matches": sorted(sizes)[len(sizes) // 2] if sizes else 0, "cv2_raised": raised,
                 "features_per_frame": len(o.database.log) / 2 / max(1, n - 1)}
print(json.dumps(out, indent=1))
