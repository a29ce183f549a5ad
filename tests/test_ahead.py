"""Detection issued ahead of the next feed (engine.cpp, Lane::Ahead): the occupancy pass and the FAST launch of feed t+1 are
issued when feed t ends.  Whatever they read may still change before feed t+1 -- set_num_features, change_feat_id, another
batch composition, a tracker that lost every point -- and must then be redone; results never differ from the oracle's."""
import numpy as np
import pytest

from uvio_b200 import configs
from uvio_b200.track import CameraData, feed_batch, make_tracker

from util import frames_for, log_arrays, masks_for


def _run(make, n=9):
    from oracle.trackklt_cv2 import make_oracle
    cfg = configs.EUROC_MAV.with_(num_features=40)
    frames = frames_for(cfg, n, stream_id=17, events=True)
    masks = masks_for(cfg)
    ora, trk = make_oracle(cfg), make(cfg)
    total = 0
    for k, (l, r) in enumerate(frames):
        if k == 3:                                  # VioManagerHelper.cpp:124: the feature budget is raised after initialisation
            ora.set_num_features(120)
            trk.set_num_features(120)
        if k == 5 and ora.ids_last[0]:              # TrackBase.cpp:230-248
            old = ora.ids_last[0][0]
            ora.change_feat_id(old, old + 100000)
            trk.change_feat_id(old, old + 100000)
        if k == 7:                                  # a blank frame: tracks are lost, the next feed may start over
            l, r = np.zeros_like(l), np.zeros_like(r)
        n0 = len(ora.database.log)
        ora.feed_new_camera(0.05 * k, [0, 1], [l, r], masks)
        trk.feed_new_camera(CameraData(0.05 * k, [0, 1], [l, r], masks))
        oi, oc, ouv, _ = log_arrays(ora.database.log[n0:])
        obs = trk.new_observations()
        assert np.array_equal(obs["id"].astype(np.int64), oi), f"frame {k}: id stream differs"
        assert np.array_equal(obs["cam_id"].astype(np.int64), oc), f"frame {k}: camera stream differs"
        if len(oi):
            assert np.array_equal(np.c_[obs["u"], obs["v"]], ouv), f"frame {k}: positions not bit-exact"
        last = trk.get_last_ids()
        for c in (0, 1):
            assert np.array_equal(last[c].astype(np.int64), np.array(ora.ids_last[c], np.int64)), f"frame {k} cam {c}: ids_last"
        total += len(oi)
    return total


def test_ahead_invalidation_sim(sim_tracker):
    assert _run(sim_tracker) > 200


@pytest.mark.gpu
def test_ahead_invalidation_gpu():
    assert _run(lambda c: make_tracker(c, device=0), n=14) > 400


def _changing_batches(make, cfg, n_trk, n):
    """The trackers of a lane change from call to call (a batch, then singles, then another order): each stream must
    still equal its own single-tracker run.  Every feed detects (the feature budget is never reached), so a detection
    that was issued ahead for a tracker and went stale -- the tracker was fed through another lane in between -- shows."""
    masks = masks_for(cfg)
    seqs = [frames_for(cfg, n, stream_id=60 + s) for s in range(n_trk)]
    ref = []
    for s in range(n_trk):
        t = make(cfg)
        out = []
        for k in range(n):
            t.feed_new_camera(CameraData(0.05 * k, [0, 1], list(seqs[s][k]), masks))
            out.append(t.new_observations().tobytes())
        ref.append(out)
    trks = [make(cfg) for _ in range(n_trk)]
    order = [3, 1, 4, 0, 2][:n_trk] if n_trk == 5 else list(range(n_trk))[::-1]
    for k in range(n):
        msgs = [CameraData(0.05 * k, [0, 1], list(seqs[s][k]), masks) for s in range(n_trk)]
        if k % 3 == 0:
            feed_batch(trks, msgs)
        elif k % 3 == 1:
            for s in range(n_trk):
                trks[s].feed_new_camera(msgs[s])
        else:
            feed_batch([trks[i] for i in order], [msgs[i] for i in order])
        for s in range(n_trk):
            assert trks[s].new_observations().tobytes() == ref[s][k], f"frame {k} tracker {s}"


def test_ahead_changing_batches_sim(sim_tracker, monkeypatch):
    monkeypatch.setenv("UVIO_KLT_LANES", "3")
    cams = tuple(configs.CamCalib(c.model, (c.intrinsics[0] * 320 / 752, c.intrinsics[1] * 240 / 480, 160.0, 120.0), c.dist)
                 for c in configs.EUROC_MAV.cams)
    cfg = configs.EUROC_MAV.with_(name="small", width=320, height=240, num_features=300, min_px_dist=5, cams=cams)
    _changing_batches(sim_tracker, cfg, 3, 7)


@pytest.mark.gpu
def test_ahead_changing_batches_gpu():
    _changing_batches(lambda c: make_tracker(c, device=0), configs.EUROC_MAV.with_(num_features=400, min_px_dist=6), 5, 9)


def _destroy_mid_sequence(make, cfg, n_trk, n, drop_at, drop):
    """More trackers than lanes (several per lane); one is destroyed between two batched feeds, which shifts every later
    tracker to another place of its lane or to another lane.  The survivors must not notice."""
    masks = masks_for(cfg)
    seqs = [frames_for(cfg, n, stream_id=80 + s) for s in range(n_trk)]
    ref = []
    for s in range(n_trk):
        t = make(cfg)
        out = []
        for k in range(n):
            t.feed_new_camera(CameraData(0.05 * k, [0, 1], list(seqs[s][k]), masks))
            out.append(t.new_observations().tobytes())
        ref.append(out)
        t.close()
    trks = {s: make(cfg) for s in range(n_trk)}
    for k in range(n):
        if k == drop_at:
            trks.pop(drop).close()
        alive = sorted(trks)
        feed_batch([trks[s] for s in alive], [CameraData(0.05 * k, [0, 1], list(seqs[s][k]), masks) for s in alive])
        for s in alive:
            got = trks[s].new_observations()
            if got.tobytes() != ref[s][k]:
                want = np.frombuffer(ref[s][k], dtype=got.dtype)
                bad = [i for i in range(min(len(got), len(want))) if got[i].tobytes() != want[i].tobytes()]
                raise AssertionError(f"frame {k} tracker {s}: {len(got)} vs {len(want)} observations, {len(bad)} differ, first "
                                     f"{bad[:3]}: got {[got[i] for i in bad[:3]]} want {[want[i] for i in bad[:3]]}")


def _small_cfg():
    cams = tuple(configs.CamCalib(c.model, (c.intrinsics[0] * 320 / 752, c.intrinsics[1] * 240 / 480, 160.0, 120.0), c.dist)
                 for c in configs.EUROC_MAV.cams)
    return configs.EUROC_MAV.with_(name="small", width=320, height=240, num_features=300, min_px_dist=5, cams=cams)


def test_ahead_destroy_mid_sequence_sim(sim_tracker):
    _destroy_mid_sequence(sim_tracker, _small_cfg(), 11, 5, 2, 3)


@pytest.mark.gpu
def test_ahead_destroy_mid_sequence_gpu():
    _destroy_mid_sequence(lambda c: make_tracker(c, device=0), configs.EUROC_MAV.with_(num_features=400, min_px_dist=6), 19, 7, 3, 5)


def _camera_subsets(make, n=8):
    """One tracker fed [0]-only and [1]-only messages in turn, then the pair in swapped order (TrackBase.h:60-65 allows all
    of it: the reference keys its state by camera id, TrackKLT.cpp:80-89).  A detection issued ahead for the unit that held
    the OTHER camera must not be taken over (round 1 accepted it: units are indexed by message position)."""
    from oracle.trackklt_cv2 import make_oracle
    for stereo, plan in ((False, [[0], [1]] * (n // 2)), (False, [[0, 1], [1, 0]] * (n // 2)), (True, [[0, 1], [1, 0]] * (n // 2))):
        cfg = configs.EUROC_MAV.with_(num_features=300, min_px_dist=6, use_stereo=stereo)   # budget never reached: every feed detects
        frames = frames_for(cfg, n, stream_id=23, events=False)
        masks = masks_for(cfg)
        ora, trk = make_oracle(cfg), make(cfg)
        for k, cams in enumerate(plan):
            imgs = [frames[k][c] for c in cams]
            n0 = len(ora.database.log)
            ora.feed_new_camera(0.05 * k, cams, imgs, [masks[c] for c in cams])
            trk.feed_new_camera(CameraData(0.05 * k, cams, imgs, [masks[c] for c in cams]))
            oi, oc, ouv, _ = log_arrays(ora.database.log[n0:])
            obs = trk.new_observations()
            what = f"stereo={stereo} plan={plan[:2]} frame {k}"
            assert np.array_equal(obs["id"].astype(np.int64), oi), f"{what}: id stream differs"
            assert np.array_equal(obs["cam_id"].astype(np.int64), oc), f"{what}: camera stream differs"
            if len(oi):
                assert np.array_equal(np.c_[obs["u"], obs["v"]], ouv), f"{what}: positions not bit-exact"
            last = trk.get_last_ids()
            for c in cams:
                assert np.array_equal(last[c].astype(np.int64), np.array(ora.ids_last.get(c, []), np.int64)), f"{what} cam {c}: ids_last"


def test_ahead_camera_subsets_sim(sim_tracker):
    _camera_subsets(sim_tracker, n=6)


@pytest.mark.gpu
def test_ahead_camera_subsets_gpu():
    _camera_subsets(lambda c: make_tracker(c, device=0), n=10)
