"""Runs the same sequences repeatedly (single tracker, and a 19-tracker batch) and counts runs whose observation streams
differ from the first run.  Development probe for a rare nondeterminism seen in tests/test_ahead.py on the GPU."""
import os, sys, hashlib
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
from uvio_b200 import configs
from uvio_b200.track import CameraData, feed_batch, make_tracker
from util import frames_for, masks_for

cfg = configs.EUROC_MAV.with_(num_features=400, min_px_dist=6)
masks = masks_for(cfg)
n_trk, n = 19, 7
seqs = [frames_for(cfg, n, stream_id=80 + s) for s in range(n_trk)]

def single(s):
    t = make_tracker(cfg, device=0); h = []
    for k in range(n):
        t.feed_new_camera(CameraData(0.05 * k, [0, 1], list(seqs[s][k]), masks)); h.append(hashlib.md5(t.new_observations().tobytes()).hexdigest())
    t.close(); return h

def batch():
    trks = [make_tracker(cfg, device=0) for _ in range(n_trk)]; h = []
    for k in range(n):
        feed_batch(trks, [CameraData(0.05 * k, [0, 1], list(seqs[s][k]), masks) for s in range(n_trk)])
        h.append([hashlib.md5(t.new_observations().tobytes()).hexdigest() for t in trks])
    for t in trks: t.close()
    return h

reps = int(os.environ.get("REPS", "30"))
skip_single = os.environ.get("SKIP_SINGLE") == "1"
ref = [] if skip_single else [single(s) for s in range(4)]
bad_single = 0
for r in range(0 if skip_single else reps):
    for s in range(4):
        h = single(s)
        if h != ref[s]:
            bad_single += 1
            print("single mismatch: run", r, "tracker", s, "first frame", [i for i in range(n) if h[i] != ref[s][i]][:1])
b0 = batch(); bad_batch = 0
for r in range(reps // 2):
    b = batch()
    if b != b0:
        bad_batch += 1
        d = [(k, s) for k in range(n) for s in range(n_trk) if b[k][s] != b0[k][s]]
        print("batch mismatch: run", r, "first (frame, tracker)", d[:2], "count", len(d))
print("RESULT", os.environ.get("UVIO_KLT_AHEAD", "1"), "single bad", bad_single, "of", 4 * reps, "batch bad", bad_batch, "of", reps // 2)
