"""Golden fixtures of the descriptor tracker (tests/golden/desc_*.npz) from oracle/trackdescriptor_cv2.py (cv2 kernels, IPP off):

    python tests/golden/make_desc_golden.py

Per frame: the update_feature stream (ids, cameras, positions, normalised positions) and, per camera, ids_last / pts_last /
desc_last; plus `ids_ok_until` = the first frame whose robust_match was left with 10..13 symmetric matches (cv::findFundamentalMat
then runs LMedS, which OpenCV itself does not reproduce: DESIGN.md section 5) -- ids are compared only before it.  Inputs are
regenerated from uvio_b200.synth by seed; their checksum is stored."""
import os
import sys

import cv2
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from oracle.trackdescriptor_cv2 import make_descriptor_oracle  # noqa: E402
from uvio_b200 import configs  # noqa: E402
from uvio_b200.synth import frames_checksum  # noqa: E402
from util import frames_for, log_arrays, masks_for  # noqa: E402

cv2.ipp.setUseIPP(False)
SEQUENCES = [("euroc_mav", configs.EUROC_MAV, 10, 3, False), ("euroc_mono", configs.EUROC_MAV, 8, 5, True), ("uzhfpv_indoor", configs.UZHFPV_INDOOR, 8, 4, False)]


def main():
    for name, cfg, n, stream_id, mono in SEQUENCES:
        frames = frames_for(cfg, n, stream_id=stream_id)
        masks = masks_for(cfg)
        o = make_descriptor_oracle(cfg)
        out = dict(meta=np.array(str(dict(cv2_version=cv2.__version__, ipp=False, knn_ratio=0.70))), n_frames=n, stream_id=stream_id, mono=mono,
                   frames_checksum=np.uint64(frames_checksum([f for lr in frames for f in lr])))
        ids_ok_until = n
        for k, (l, r) in enumerate(frames):
            n0, r0 = len(o.database.log), len(o.ransac_sizes)
            cams = [0] if mono else [0, 1]
            o.feed_new_camera(0.05 * k, cams, [l] if mono else [l, r], masks[:len(cams)])
            sizes = o.ransac_sizes[r0:]
            if any(10 <= s <= 13 for s in sizes):
                ids_ok_until = min(ids_ok_until, k)
            out[f"ransac_sizes_{k}"] = np.array(sizes, np.int32)
            ids, cc, uv, uvn = log_arrays(o.database.log[n0:])
            out[f"ids_{k}"], out[f"cams_{k}"], out[f"uv_{k}"], out[f"uvn_{k}"] = ids, cc.astype(np.int8), uv, uvn
            for c in cams:
                out[f"last_ids_{k}_{c}"] = np.array(o.ids_last[c], np.int64)
                out[f"last_uv_{k}_{c}"] = np.asarray(o.pts_last[c], np.float32).reshape(-1, 2)
                out[f"last_desc_{k}_{c}"] = np.asarray(o.desc_last[c], np.uint8).reshape(-1, 32)
        out["ids_ok_until"] = ids_ok_until
        path = os.path.join(HERE, f"desc_{name}.npz")
        np.savez_compressed(path, **out)
        print(name, "frames", n, "ids comparable before frame", ids_ok_until, "observations", len(o.database.log), os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
