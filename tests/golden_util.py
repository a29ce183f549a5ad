"""Comparison of a tracker (GPU through the C ABI, or the host-simulation build) with the committed golden sequences."""
import os

import numpy as np

from uvio_b200 import configs
from uvio_b200.synth import frames_checksum
from uvio_b200.track import CameraData

from util import frames_for, masks_for

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
SEQ_CFG = {"euroc_mav": configs.EUROC_MAV, "tum_vi": configs.TUM_VI, "rs_d455": configs.RS_D455,
           "uzhfpv_indoor": configs.UZHFPV_INDOOR, "euroc_mono": configs.EUROC_MAV, "hd_1000": configs.HD_1000}


def load(name):
    return np.load(os.path.join(GOLDEN, f"seq_{name}.npz"))


def golden_inputs(name, g, n=None):
    cfg = SEQ_CFG[name]
    n_all = int(g["n_frames"])
    frames = frames_for(cfg, n_all, stream_id=int(g["stream_id"]))
    assert frames_checksum([f for lr in frames for f in lr]) == int(g["frames_checksum"]), "synthetic generator drifted"
    return cfg, frames[:n or n_all], masks_for(cfg)


def check_tracker_against_golden(name, make, n=None, exact=True):
    """make(cfg) -> tracker with the TrackBase-shaped API of uvio_b200.track.TrackKLT."""
    g = load(name)
    cfg, frames, masks = golden_inputs(name, g, n)
    mono = bool(g["mono"])
    trk = make(cfg)
    total = 0
    for k, (l, r) in enumerate(frames):
        if mono:
            trk.feed_new_camera(CameraData(0.05 * k, [0], [l], [masks[0]]))
        else:
            trk.feed_new_camera(CameraData(0.05 * k, [0, 1], [l, r], masks))
        obs = trk.new_observations()
        assert np.array_equal(obs["id"].astype(np.int64), g[f"ids_{k}"]), f"{name} frame {k}: id stream differs"
        assert np.array_equal(obs["cam_id"].astype(np.int8), g[f"cams_{k}"]), f"{name} frame {k}: camera stream differs"
        uv, uvn = np.c_[obs["u"], obs["v"]], np.c_[obs["un"], obs["vn"]]
        if len(uv):
            assert np.abs(uv - g[f"uv_{k}"]).max() <= 0.01                      # north_star tolerance
            assert np.abs(uvn - g[f"uvn_{k}"]).max() <= 1e-5
            if exact:
                assert np.array_equal(uv, g[f"uv_{k}"]), f"{name} frame {k}: positions not bit-exact"
        ids_last, obs_last = trk.get_last_ids(), trk.get_last_obs()
        for c in ([0] if mono else [0, 1]):
            assert np.array_equal(ids_last[c].astype(np.int64), g[f"last_ids_{k}_{c}"]), f"{name} frame {k} cam {c}: ids_last"
            if exact and len(ids_last[c]):
                assert np.array_equal(obs_last[c], g[f"last_uv_{k}_{c}"])
        total += len(obs)
    return total
