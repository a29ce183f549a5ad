"""Why is the end-to-end leg slower than the resident one?  Separates (a) the host-side mask scan, (b) PCIe contention
between the frame upload and the small host<->device traffic of the rounds.  Development probe, not part of the bench."""
import ctypes as C
import json
import os
import sys
import threading

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from uvio_b200 import build, track  # noqa: E402

build.build()
L = track.lib()
cfg = bench.workload_config()
S, W, K = int(os.environ.get("S", "32")), 5, 30
T = W + K + 1
h, w = cfg.height, cfg.width
frames = bench.make_frames(cfg, 0, S, 8, T)
host = torch.empty((T, S, 2, h, w), dtype=torch.uint8).pin_memory()
hv = host.numpy()
for t in range(T):
    for s in range(S):
        hv[t, s, 0], hv[t, s, 1] = frames[t][s]
dev = host.to("cuda:0")
zero = np.zeros((h, w), np.uint8)
keep = []
dev_nomask = bench._frame_structs(track, dev.data_ptr(), T, S, h, w, None, keep)
dev_mask = bench._frame_structs(track, dev.data_ptr(), T, S, h, w, zero.ctypes.data, keep)
host_nomask = bench._frame_structs(track, host.data_ptr(), T, S, h, w, None, keep)
host_mask = bench._frame_structs(track, host.data_ptr(), T, S, h, w, zero.ctypes.data, keep)
obs = np.zeros((S, 1024), track.OBS_DTYPE)
counts = (C.c_int * S)()


class Background:
    def __init__(self, chunk):
        self.chunk, self.stop = chunk, False
        self.src = torch.empty(S * 2 * h * w, dtype=torch.uint8).pin_memory()
        self.dst = torch.empty_like(self.src, device="cuda:0")
        self.st = torch.cuda.Stream()
        self.copied = 0
        self.th = threading.Thread(target=self.run, daemon=True)

    def run(self):
        n = self.src.numel()
        ev = torch.cuda.Event()
        with torch.cuda.stream(self.st):
            while not self.stop:
                for o in range(0, n, self.chunk):
                    self.dst[o:o + self.chunk].copy_(self.src[o:o + self.chunk], non_blocking=True)
                ev.record(self.st)
                ev.synchronize()
                self.copied += n

    def __enter__(self):
        self.th.start()
        return self

    def __exit__(self, *a):
        self.stop = True
        self.th.join()


def leg(name, steps, announce, bg_chunk=0):
    if bg_chunk:
        with Background(bg_chunk) as b:
            sec, n_obs, _c, st, _p, _w = bench._run_leg(track, L, cfg, 0, S, steps, W, K, None, obs, counts, False, 0, announce)
            gb = b.copied / 1e9
    else:
        sec, n_obs, _c, st, _p, _w = bench._run_leg(track, L, cfg, 0, S, steps, W, K, None, obs, counts, False, 0, announce)
        gb = 0
    print(json.dumps({"leg": name, "frames_per_s": round(S * K / sec), "ms_per_step": round(1e3 * sec / K, 3),
                      "bg_GB": round(gb, 2), "host": st["host_us_per_lane_pass"]}), flush=True)


leg("resident nomask", dev_nomask, True)
leg("resident hostmask", dev_mask, True)
leg("resident nomask + bg h2d whole", dev_nomask, True, S * 2 * h * w)
leg("resident nomask + bg h2d 1MB chunks", dev_nomask, True, 1 << 20)
leg("resident nomask + bg h2d 256KB chunks", dev_nomask, True, 256 << 10)
leg("e2e nomask announce", host_nomask, True)
leg("e2e hostmask announce", host_mask, True)
leg("e2e nomask plain", host_nomask, False)
