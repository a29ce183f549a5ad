"""GPU parity against the committed golden fixtures (tests/golden/, produced by the oracle; no cv2 needed at run time)."""
import os

import numpy as np
import pytest

from uvio_b200 import configs
from uvio_b200.synth import frames_checksum
from uvio_b200.track import CameraData, make_tracker

from golden_util import GOLDEN, check_tracker_against_golden
from util import frames_for, masks_for

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", ["euroc_mav", "tum_vi", "rs_d455", "uzhfpv_indoor", "euroc_mono", "hd_1000"])
def test_sequence_matches_golden(built_lib, name):
    assert check_tracker_against_golden(name, lambda cfg: make_tracker(cfg)) > 1000


def test_stage_goldens(built_lib):
    g = np.load(os.path.join(GOLDEN, "stages_euroc.npz"))
    cfg = configs.EUROC_MAV
    frames = frames_for(cfg, 2, events=False, start=7)
    assert frames_checksum([f for lr in frames for f in lr]) == int(g["frames_checksum"])
    trk = make_tracker(cfg)
    for k, fr in enumerate(frames):
        trk.feed_new_camera(CameraData(0.05 * k, [0, 1], [fr[0], fr[1]], masks_for(cfg)))
    for lv in range(5):                                                    # equalizeHist + pyramid, bit-exact
        assert frames_checksum([trk.stage_level(0, 1, lv)]) == int(g[f"pyr_sum_{lv}"]), f"level {lv}"
    clahe = make_tracker(cfg.with_(histogram=configs.HIST_CLAHE))
    clahe.feed_new_camera(CameraData(0.0, [0, 1], [frames[0][0], frames[0][1]], masks_for(cfg)))
    assert frames_checksum([clahe.stage_level(0, 0, 0)]) == int(g["clahe_sum"])
    sx, sy = cfg.width // cfg.grid_x, cfg.height // cfg.grid_y
    for (gx, gy) in ((0, 0), (2, 3), (4, 4)):                              # FAST: positions, order, responses
        xy, r = trk.stage_fast(0, 1, gx * sx, gy * sy, sx, sy, cfg.fast_threshold)
        assert np.array_equal(xy, g[f"fast_xy_{gx}_{gy}"].astype(np.int32)) and np.array_equal(r, g[f"fast_resp_{gx}_{gy}"].astype(np.float32))
    sp = trk.stage_subpix(0, 1, g["subpix_in"])
    assert np.abs(sp - g["subpix_out"]).max() <= 0.01 and np.array_equal(sp, g["subpix_out"])
    out, st = trk.stage_lk(0, 1, 0, 0, g["lk_in"], g["lk_in"])
    assert np.array_equal(st, g["lk_status"])
    ok = st == 1
    assert np.abs(out - g["lk_out"])[ok].max() <= 0.01 and np.array_equal(out[ok], g["lk_out"][ok])
    mask, p0n, _p1n = trk.stage_ransac(0, 0, g["ransac_p0"], g["ransac_p1"])
    assert np.array_equal(mask, g["ransac_mask"])
    _m, un, _ = trk.stage_ransac(0, 0, g["undist_in_radtan"], g["undist_in_radtan"])
    assert np.array_equal(un, g["undist_out_radtan"])
    equi = make_tracker(configs.TUM_VI)
    _m, un, _ = equi.stage_ransac(0, 0, g["undist_in_equi"], g["undist_in_equi"])
    assert np.abs(un - g["undist_out_equi"]).max() <= 1e-6


def test_size_independent_properties_at_full_size(built_lib):
    """1280x720 / 1000 features (BASELINE configs[3]): properties that need no oracle."""
    cfg = configs.HD_1000
    frames = frames_for(cfg, 6, events=False)
    trk = make_tracker(cfg)
    seen = set()
    for k, (l, r) in enumerate(frames):
        trk.feed_new_camera(CameraData(0.05 * k, [0, 1], [l, r], masks_for(cfg)))
        obs = trk.new_observations()
        ids, last = trk.get_last_ids(), trk.get_last_obs()
        for c in (0, 1):
            assert len(set(ids[c].tolist())) == len(ids[c])                # ids unique per camera
            assert len(ids[c]) <= cfg.num_features + cfg.grid_x * cfg.grid_y * (cfg.num_features // 25 + 1)
            if len(ids[c]):
                assert last[c][:, 0].min() >= 0 and last[c][:, 0].max() <= cfg.width and last[c][:, 1].max() <= cfg.height
                cells = (last[c] / cfg.min_px_dist).astype(np.int32)      # min-distance grid: no two NEW points share a cell
        if k > 0:
            assert len(obs) > 800
            assert set(obs["id"].tolist()) <= seen | set(ids[0].tolist()) | set(ids[1].tolist())
        seen |= set(ids[0].tolist()) | set(ids[1].tolist())
        # pyramid idempotence: level l+1 is a pure function of level l -> identical for identical inputs
    a = make_tracker(cfg)
    a.feed_new_camera(CameraData(0.0, [0, 1], [frames[0][0], frames[0][0]], masks_for(cfg)))
    for lv in range(a.stage_num_levels(0)):
        assert np.array_equal(a.stage_level(0, 0, lv), a.stage_level(1, 0, lv))
    eq = a.stage_level(0, 0, 0)                                            # equalisation is monotone in the input grey level
    src = frames[0][0]
    order = np.argsort(src.ravel(), kind="stable")
    assert np.all(np.diff(eq.ravel()[order].astype(np.int32)) >= 0)
