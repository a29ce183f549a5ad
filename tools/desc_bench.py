"""Throughput of the descriptor tracker: S synthetic euroc_mav stereo streams fed through
uvio_desc_feed_batch from pinned host frames (upload + results inside the timed region), next to the cv2-driven oracle
(oracle/trackdescriptor_cv2.py = the reference's TrackDescriptor control flow over the real OpenCV kernels) on one host core.

    python tools/desc_bench.py [--streams 64] [--feeds 40] > profiles/r2_desc_bench.json
"""
import argparse
import ctypes as C
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch

import bench
from uvio_b200 import desc, track


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--streams", type=int, default=64)
    ap.add_argument("--feeds", type=int, default=40)
    ap.add_argument("--warm", type=int, default=6)
    ap.add_argument("--pool", type=int, default=12)
    ap.add_argument("--config", default="euroc_mav")
    ap.add_argument("--cpu-frames", type=int, default=60)
    a = ap.parse_args()
    cfg = bench.workload_config(a.config)
    S, NP = a.streams, a.pool
    h, w = cfg.height, cfg.width
    frames = bench.make_frames(cfg, 0, S, 8, NP)
    hv = bench.pinned_hugepage_buffer(NP * S * 2 * h * w).reshape(NP, S, 2, h, w)
    for t in range(NP):
        for s in range(S):
            hv[t, s, 0], hv[t, s, 1] = frames[t][s]
    keep = []
    pool = bench._frame_structs(track, hv.ctypes.data, NP, S, h, w, keep)
    L = desc.lib()
    trackers = [desc.make_descriptor_tracker(cfg) for _ in range(S)]
    hs = (C.c_void_p * S)(*[t._h for t in trackers])
    obs = np.zeros(8192, track.OBS_DTYPE)
    n = C.c_int()

    def feed(t):
        rc = L.uvio_desc_feed_batch(hs, pool[bench.pool_index(t, NP)], S)
        assert rc == 0, rc
        tot = 0
        for x in trackers:
            L.uvio_desc_get_new_observations(x._h, obs.ctypes.data, len(obs), C.byref(n))
            tot += n.value
        return tot

    for t in range(a.warm):
        feed(t)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    tot = 0
    dev_us = 0.0
    for t in range(a.warm, a.warm + a.feeds):
        tot += feed(t)
        dev_us += trackers[0].last_feed_us()
    torch.cuda.synchronize()
    sec = time.perf_counter() - t0
    out = {"what": "ov_core TrackDescriptor on one B200: uvio_desc_feed_batch from pinned host frames, results read back, wall clock",
           "config": bench.config_dict(cfg, S, 1)["workload"].replace("TrackKLT", "TrackDescriptor"), "knn_ratio": 0.70,
           "streams": S, "feeds": a.feeds, "stereo_frames_per_sec": S * a.feeds / sec, "ms_per_feed": 1e3 * sec / a.feeds,
           "device_ms_per_feed": dev_us / a.feeds / 1e3, "observations_per_frame": tot / (S * a.feeds),
           "h2d_bytes_per_feed": S * 2 * h * w}
    # the CPU arm: the oracle on one stream, one thread (cv2.setNumThreads(1))
    import cv2
    from oracle.trackdescriptor_cv2 import make_descriptor_oracle
    cv2.setNumThreads(1)
    o = make_descriptor_oracle(cfg)
    z = [np.zeros((h, w), np.uint8)] * 2
    from uvio_b200.synth import SynthStereoStream
    st = SynthStereoStream(w, h, 0)
    fr = [st.frame(k) for k in range(a.cpu_frames + 5)]
    for k in range(5):
        o.feed_new_camera(0.05 * k, [0, 1], list(fr[k]), z)
    t0 = time.perf_counter()
    for k in range(5, 5 + a.cpu_frames):
        o.feed_new_camera(0.05 * k, [0, 1], list(fr[k]), z)
    cs = time.perf_counter() - t0
    out["cpu_baseline"] = {"value": a.cpu_frames / cs, "unit": "frames/s", "cores": 1, "kind": "port",
                           "what": "oracle/trackdescriptor_cv2.py (cv2 %s kernels, Python control flow), one stream, one thread" % cv2.__version__,
                           "host_cores": os.cpu_count()}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
