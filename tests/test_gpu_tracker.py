"""GPU parity of the whole tracker through the C ABI: free-running TrackKLT vs the cv2-driven oracle on the same frames.
The update_feature stream (ids, cameras, order) must be identical and positions within 0.01 px (they are bit-exact)."""
import numpy as np
import pytest

from uvio_b200 import configs
from uvio_b200.track import CameraData, UvioError, feed_batch, make_tracker

from util import frames_for, log_arrays, masks_for, run_oracle

pytestmark = pytest.mark.gpu


def _compare_frame(k, trk, obs, ref, cfg_name):
    log, pts_last, ids_last = ref
    ids, cams, uv, uvn = log_arrays(log)
    g_ids, g_cams = obs["id"].astype(np.int64), obs["cam_id"].astype(np.int64)
    assert len(g_ids) == len(ids), f"{cfg_name} frame {k}: {len(g_ids)} observations, oracle has {len(ids)}"
    assert np.array_equal(g_ids, ids) and np.array_equal(g_cams, cams), f"{cfg_name} frame {k}: id/cam stream differs"
    g_uv = np.c_[obs["u"], obs["v"]]
    g_uvn = np.c_[obs["un"], obs["vn"]]
    if len(ids):
        assert np.abs(g_uv - uv).max() <= 0.01, f"{cfg_name} frame {k}: {np.abs(g_uv - uv).max()} px"
        assert np.abs(g_uvn - uvn).max() <= 1e-5
        assert np.array_equal(g_uv, uv), f"{cfg_name} frame {k}: positions not bit-exact"
    last_ids, last_obs = trk.get_last_ids(), trk.get_last_obs()
    for cam, want in ids_last.items():
        assert np.array_equal(last_ids[cam].astype(np.int64), np.array(want, np.int64)), f"{cfg_name} frame {k} cam {cam}: ids_last"
        if want:
            assert np.array_equal(last_obs[cam], np.array(pts_last[cam], np.float32).reshape(-1, 2)), f"frame {k} cam {cam}: pts_last"


SEQ = [(configs.EUROC_MAV, 40), (configs.TUM_VI, 30), (configs.RS_D455, 25), (configs.UZHFPV_INDOOR, 25),
       (configs.EUROC_MAV.with_(name="euroc_pyr3", pyr_levels=3), 20), (configs.HD_1000, 12), (configs.HD_1000_PYR5, 6)]


@pytest.mark.parametrize("cfg,n", SEQ, ids=lambda v: getattr(v, "name", str(v)))
def test_stereo_sequence_matches_oracle(built_lib, cfg, n):
    frames = frames_for(cfg, n)
    masks = masks_for(cfg)
    _o, ref = run_oracle(cfg, frames, masks)
    trk = make_tracker(cfg)
    nobs = 0
    for k, (l, r) in enumerate(frames):
        trk.feed_new_camera(CameraData(0.05 * k, [0, 1], [l, r], masks))
        obs = trk.new_observations()
        _compare_frame(k, trk, obs, ref[k], cfg.name)
        nobs += len(obs)
    assert nobs > 50 * (n // 4)
    assert trk.get_feature_database().size() > 0


def test_monocular_sequence_matches_oracle(built_lib):
    cfg = configs.EUROC_MAV
    frames = frames_for(cfg, 30, stream_id=3)
    masks = masks_for(cfg)
    _o, ref = run_oracle(cfg, frames, masks, mono=True)
    trk = make_tracker(cfg)
    for k, (l, _r) in enumerate(frames):
        trk.feed_new_camera(CameraData(0.05 * k, [0], [l], [masks[0]]))
        _compare_frame(k, trk, trk.new_observations(), ref[k], "euroc_mono")


def test_set_num_features_mid_sequence(built_lib):
    cfg = configs.EUROC_MAV
    frames = frames_for(cfg, 16, stream_id=5, events=False)
    masks = masks_for(cfg)
    _o, ref = run_oracle(cfg, frames, masks, num_features=25, set_nf_at={6: 150})     # VioManagerHelper.cpp:124
    trk = make_tracker(cfg, numfeats=25)
    for k, (l, r) in enumerate(frames):
        if k == 6:
            trk.set_num_features(150)
        trk.feed_new_camera(CameraData(0.05 * k, [0, 1], [l, r], masks))
        _compare_frame(k, trk, trk.new_observations(), ref[k], "euroc_nf")


def test_batched_feed_equals_single_feeds(built_lib):
    cfg = configs.UZHFPV_INDOOR
    n_streams, n = 4, 10
    masks = masks_for(cfg)
    seqs = [frames_for(cfg, n, stream_id=10 + s) for s in range(n_streams)]
    single = [make_tracker(cfg) for _ in range(n_streams)]
    batched = [make_tracker(cfg) for _ in range(n_streams)]
    for k in range(n):
        msgs = [CameraData(0.05 * k, [0, 1], [seqs[s][k][0], seqs[s][k][1]], masks) for s in range(n_streams)]
        for s in range(n_streams):
            single[s].feed_new_camera(msgs[s])
        feed_batch(batched, msgs)
        for s in range(n_streams):
            a, b = single[s].new_observations(), batched[s].new_observations()
            assert a.tobytes() == b.tobytes(), f"stream {s} frame {k}"
    assert all(t.get_feature_database().size() > 50 for t in batched)


def test_error_codes(built_lib):
    cfg = configs.EUROC_MAV
    trk = make_tracker(cfg)
    img = np.zeros((cfg.height, cfg.width), np.uint8)
    with pytest.raises(SystemExit):                                   # TrackKLT.cpp:37-43
        trk.feed_new_camera(CameraData(0.0, [0, 1], [img], [img]))
    with pytest.raises(UvioError) as e:                               # unknown camera id (.at throws in the reference)
        trk.feed_new_camera(CameraData(0.0, [7], [img], [img]))
    assert e.value.code == 2
    with pytest.raises(SystemExit):                                   # 3 images with use_stereo (TrackKLT.cpp:90-93)
        make_tracker(cfg.with_(cams=cfg.cams + (cfg.cams[0],))).feed_new_camera(CameraData(0.0, [0, 1, 2], [img] * 3, [img] * 3))
    trk.feed_new_camera(CameraData(0.0, [0, 1], [img, img], [img, img]))
    with pytest.raises(UvioError) as e:
        trk.feed_new_camera(CameraData(0.1, [0, 1], [img[:100], img[:100]], [img[:100], img[:100]]))
    assert e.value.code == 5
