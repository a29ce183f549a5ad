"""CPU: the oracle reproduces the committed golden fixtures (pins cv2's behaviour on this machine to the fixtures), and
the plain-C restatement of every OpenCV routine (oracle/stages.c) agrees with cv2 and with the fixtures."""
import os

import cv2
import numpy as np
import pytest

from uvio_b200 import configs
from uvio_b200.synth import frames_checksum

from golden_util import GOLDEN, golden_inputs, load
from util import frames_for, log_arrays, run_oracle

cv2.ipp.setUseIPP(False)


@pytest.mark.parametrize("name,n", [("euroc_mav", 10), ("tum_vi", 8), ("euroc_mono", 8)])
def test_oracle_reproduces_golden_sequences(name, n):
    g = load(name)
    cfg, frames, masks = golden_inputs(name, g, n)
    _o, ref = run_oracle(cfg, frames, masks, mono=bool(g["mono"]))
    for k, (log, _p, ids_last) in enumerate(ref):
        ids, cams, uv, uvn = log_arrays(log)
        assert np.array_equal(ids, g[f"ids_{k}"]) and np.array_equal(cams.astype(np.int8), g[f"cams_{k}"])
        assert np.array_equal(uv, g[f"uv_{k}"]) and np.array_equal(uvn, g[f"uvn_{k}"])


def test_c_restatement_against_cv2_and_goldens():
    from oracle import cstages
    g = np.load(os.path.join(GOLDEN, "stages_euroc.npz"))
    cfg = configs.EUROC_MAV
    frames = frames_for(cfg, 2, events=False, start=7)
    raw = frames[0][0]
    a = cstages.equalize_hist(raw)
    assert np.array_equal(a, cv2.equalizeHist(raw))
    assert np.array_equal(cstages.clahe(raw), cv2.createCLAHE(10.0, (8, 8)).apply(raw))
    assert frames_checksum([cstages.clahe(raw)]) == int(g["clahe_sum"])
    odd = np.ascontiguousarray(raw[:451, :645])
    assert np.array_equal(cstages.clahe(odd), cv2.createCLAHE(10.0, (8, 8)).apply(odd))
    lv = cstages.build_pyramid(a, 5)
    assert len(lv) == 5
    for k, im in enumerate(lv):
        assert frames_checksum([im]) == int(g[f"pyr_sum_{k}"])
    assert np.array_equal(cstages.scharr(a)[..., 0], cv2.Scharr(a, cv2.CV_16S, 1, 0))
    assert np.array_equal(cstages.scharr(a)[..., 1], cv2.Scharr(a, cv2.CV_16S, 0, 1))
    sx, sy = cfg.width // cfg.grid_x, cfg.height // cfg.grid_y
    for (gx, gy) in ((0, 0), (2, 3), (4, 4)):
        xy, r = cstages.fast(a[gy * sy:(gy + 1) * sy, gx * sx:(gx + 1) * sx], cfg.fast_threshold)
        assert np.array_equal(xy, g[f"fast_xy_{gx}_{gy}"].astype(np.int32)) and np.array_equal(r, g[f"fast_resp_{gx}_{gy}"].astype(np.float32))
    assert np.array_equal(cstages.corner_subpix(a, g["subpix_in"]), g["subpix_out"])
    b = cstages.equalize_hist(frames[1][0])
    out, st = cstages.lk(lv, cstages.build_pyramid(b, 5), g["lk_in"], g["lk_in"])
    assert np.array_equal(st, g["lk_status"]) and np.array_equal(out[st == 1], g["lk_out"][st == 1])
    c = cfg.cams[0]
    assert np.array_equal(cstages.undistort(g["undist_in_radtan"], 0, c.intrinsics, c.dist), g["undist_out_radtan"])
    c = configs.TUM_VI.cams[0]
    assert np.abs(cstages.undistort(g["undist_in_equi"], 1, c.intrinsics, c.dist) - g["undist_out_equi"]).max() <= 1e-6


def test_std_sort_permutation_is_not_stable_but_is_a_permutation():
    from oracle.trackklt_cv2 import std_sort_perm
    r = np.random.default_rng(0).integers(20, 40, 500).astype(np.float32)
    p = std_sort_perm(r)
    assert sorted(p.tolist()) == list(range(500)) and np.all(np.diff(r[p]) <= 0)
    assert not np.array_equal(p, np.argsort(-r, kind="stable"))          # the tie order is libstdc++'s, not a stable sort's


def test_batched_undistort_equals_per_point_calls():
    """The reference undistorts one point per OpenCV call on a 1x1 Mat (cam/CamBase.h:108-116); the oracle makes one batched call.
    OpenCV treats every point independently: the floats are the same."""
    from oracle.trackklt_cv2 import Cam
    rng = np.random.default_rng(8)
    for cfg in (configs.EUROC_MAV, configs.TUM_VI):
        c = cfg.cams[0]
        cam = Cam(c.model, tuple(c.intrinsics), tuple(c.dist))
        pts = np.c_[rng.uniform(0, cfg.width, 300), rng.uniform(0, cfg.height, 300)].astype(np.float32)
        batched = cam.undistort(pts)
        single = np.concatenate([cam.undistort(p.reshape(1, 2)) for p in pts])
        assert np.array_equal(batched, single)
