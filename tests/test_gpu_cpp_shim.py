"""GPU: the C++ TrackBase shim (uvio_b200/cpp/TrackKLTGpu.h over the C ABI), driven like ov_core/src/test_tracking.cpp drives
TrackKLT, fills the FeatureDatabase with exactly the golden observation stream."""
import os
import subprocess

import numpy as np
import pytest

from golden_util import golden_inputs, load

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_cpp_trackbase_shim_matches_golden(built_lib, tmp_path):
    exe = str(tmp_path / "shim_test")
    subprocess.check_call(["g++", "-std=c++17", "-O2", "-o", exe, os.path.join(ROOT, "tests", "cpp", "test_trackbase_shim.cpp"),
                           built_lib, "-Wl,-rpath," + os.path.dirname(built_lib)])
    g = load("euroc_mav")
    n = 10
    cfg, frames, _masks = golden_inputs("euroc_mav", g, n)
    raw = tmp_path / "frames.bin"
    with open(raw, "wb") as f:
        for l, r in frames:
            f.write(l.tobytes())
            f.write(r.tobytes())
    out = subprocess.run([exe, str(raw), str(cfg.width), str(cfg.height), str(n)], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr
    got = {}
    per_frame = {}
    for line in out.stdout.splitlines():
        t = line.split()
        if t[0] == "obs":
            got[(int(t[1]), int(t[2]), round(float(t[3]) / 0.05))] = (np.float32(t[4]), np.float32(t[5]))
        elif t[0] == "frame":
            per_frame[int(t[1])] = int(t[7])
    want = {}
    for k in range(n):
        for i, c, uv in zip(g[f"ids_{k}"], g[f"cams_{k}"], g[f"uv_{k}"]):
            want[(int(i), int(c), k)] = (uv[0], uv[1])
        assert per_frame[k] == len(g[f"ids_{k}"]), f"frame {k}: observation count"
    assert set(got) == set(want)
    assert all(got[key] == want[key] for key in want)
