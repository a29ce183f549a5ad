"""A short, fixed workload for ncu: S euroc_mav stereo streams fed n frames through uvio_klt_feed_batch with the frames resident
in HBM (what one bench step does), so that a launch list / a --set full capture of the steady-state kernels can be taken:

    ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/launches.csv python tools/profile_step.py
    ncu --set full --clock-control none --import-source on -k regex:k_lk -s 6 -c 1 -o gpurun_out/lk python tools/profile_step.py
"""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch

import bench
from uvio_b200 import track

S = int(os.environ.get("UVIO_PROFILE_STREAMS", "64"))
N = int(os.environ.get("UVIO_PROFILE_FRAMES", "8"))
cfg = bench.workload_config(os.environ.get("UVIO_PROFILE_CONFIG", "euroc_mav"))
L = track.lib()
frames = bench.make_frames(cfg, 0, S, 8, N)
h, w = cfg.height, cfg.width
hv = np.zeros((N, S, 2, h, w), np.uint8)
for t in range(N):
    for s in range(S):
        hv[t, s, 0], hv[t, s, 1] = frames[t][s]
dev = torch.from_numpy(hv).cuda()
keep = []
pool = bench._frame_structs(track, dev.data_ptr(), N, S, h, w, keep)
trackers = [track.make_tracker(cfg) for _ in range(S)]
hs = (C.c_void_p * S)(*[t._h for t in trackers])
tot = 0
for t in range(N):
    rc = L.uvio_klt_feed_batch(hs, pool[t], S)
    assert rc == 0, rc
    tot += sum(len(x.new_observations()) for x in trackers)
torch.cuda.synchronize()
st = track.device_stats(0)
print("observations", tot, "lk point-levels", st["lk_point_levels"], "lk iterations", st["lk_iterations"])
