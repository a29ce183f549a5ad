"""GPU: the C++ TrackBase shim (uvio_b200/cpp/TrackKLTGpu.h over the C ABI), driven like ov_core/src/test_tracking.cpp drives
TrackKLT, fills the FeatureDatabase with exactly the golden observation stream."""
import os
import subprocess

import numpy as np
import pytest

from golden_util import golden_inputs, load

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _build(built_lib, tmp_path):
    exe = str(tmp_path / "shim_test")
    subprocess.check_call(["g++", "-std=c++17", "-O2", "-I", os.path.join(ROOT, "tests", "cpp"), "-o", exe,
                           os.path.join(ROOT, "tests", "cpp", "test_trackbase_shim.cpp"), built_lib, "-Wl,-rpath," + os.path.dirname(built_lib)])
    return exe


def _sim_lib():
    """engine.cpp + abi.cpp over the CPU stand-ins of the launch wrappers (tests/hostsim): the same C ABI without a GPU."""
    from conftest import SIM, SIM_SOURCES, sim_build
    csrc = os.path.join(ROOT, "uvio_b200", "csrc")
    return sim_build(os.path.join(SIM, "_sim", "libuvio_klt_sim.so"),
                     SIM_SOURCES(), ["-I", SIM])


def test_cpp_shim_nonvirtual_members_host_simulation(tmp_path):
    """The same check as the GPU test below, on the host-simulation build of the library (runs without a GPU)."""
    _nonvirtual_members(_sim_lib(), tmp_path, n=11)


@pytest.mark.gpu
def test_cpp_shim_nonvirtual_members_through_the_base_pointer(built_lib, tmp_path):
    _nonvirtual_members(built_lib, tmp_path, n=14)


def _nonvirtual_members(built_lib, tmp_path, n):
    """set_num_features(25 -> 150), get_last_obs(), get_last_ids(), change_feat_id() called on std::shared_ptr<TrackBase> --
    plain members of the reference's TrackBase (TrackBase.h:134-152), used that way at VioManagerHelper.cpp:124, :209-210 --
    reach the GPU tracker: per-frame ids_last / pts_last equal the oracle's run with the same calls."""
    from uvio_b200 import configs
    from oracle.trackklt_cv2 import make_oracle
    from util import frames_for, masks_for
    exe = _build(built_lib, tmp_path)
    cfg = configs.EUROC_MAV
    frames = frames_for(cfg, n, stream_id=5, events=False)
    masks = masks_for(cfg)
    raw = tmp_path / "frames.bin"
    with open(raw, "wb") as f:
        for l, r in frames:
            f.write(l.tobytes())
            f.write(r.tobytes())
    out = subprocess.run([exe, str(raw), str(cfg.width), str(cfg.height), str(n), "base"], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr
    got = {}
    nf = {}
    for line in out.stdout.splitlines():
        t = line.split()
        if t[0] == "last":
            got.setdefault((int(t[1]), int(t[2])), []).append((int(t[3]), np.float32(t[4]), np.float32(t[5])))
        elif t[0] == "nf":
            nf[int(t[1])] = int(t[2])
    ora = make_oracle(cfg, numfeats=25)
    grew = False
    for k, (l, r) in enumerate(frames):
        if k == 6:
            ora.set_num_features(150)
        if k == 9 and ora.ids_last[0]:
            ora.change_feat_id(ora.ids_last[0][0], ora.ids_last[0][0] + 100000)
        ora.feed_new_camera(0.05 * k, [0, 1], [l, r], masks)
        assert nf[k] == (25 if k < 6 else 150)
        for c in (0, 1):
            want = [(int(i), np.float32(p[0]), np.float32(p[1])) for i, p in zip(ora.ids_last[c], ora.pts_last[c])]
            assert got.get((k, c), []) == want, f"frame {k} cam {c}: the base's ids_last / pts_last differ from the oracle's"
        grew |= k > 6 and len(ora.ids_last[0]) > 60
        if k >= 9:
            assert any(i >= 100000 for i, _x, _y in got.get((k, 0), [])) == any(i >= 100000 for i in ora.ids_last[0])
    assert grew, "set_num_features never took effect in the oracle run: the test would prove nothing"


@pytest.mark.gpu
def test_cpp_trackbase_shim_matches_golden(built_lib, tmp_path):
    exe = _build(built_lib, tmp_path)
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
