"""the descriptor tracker (ov_core::TrackDescriptor) against oracle/trackdescriptor_cv2.py -- the
reference's control flow restated line by line over the real cv2 kernels (ORB.compute, BFMatcher.knnMatch, findFundamentalMat).

Bar: key-point positions, descriptors (256 bits each), camera ids and feature ids identical; normalised coordinates within
1e-5.  One regime cannot be compared id for id: a robust_match that is left with 10..13 symmetric matches sends
cv::findFundamentalMat to LMedS, whose outcome OpenCV itself does not reproduce under a one-ulp change of an input
(profiles/r2_lmeds_probe.json).  Positions and descriptors of a frame do not depend on the temporal
matches, so from the first such feed on the test keeps comparing everything but the ids.

The same cases run (a) over the host simulation of the kernels (tests/hostsim, `not gpu`) and (b) on the GPU through the C ABI."""
import os

import numpy as np
import pytest

from uvio_b200 import configs, track
from util import frames_for, log_arrays, masks_for


def _lmeds(sizes):
    return any(10 <= n <= 13 for n in sizes)


def _run_case(make, cfg, n_frames, stream_id, mono=False, set_nf_at=None, knn_ratio=0.70):
    from oracle.trackdescriptor_cv2 import make_descriptor_oracle
    from uvio_b200 import desc
    frames = frames_for(cfg, n_frames, stream_id=stream_id)
    masks = masks_for(cfg)
    o = make_descriptor_oracle(cfg, knn_ratio=knn_ratio)
    t = make(cfg, knn_ratio=knn_ratio)
    ids_comparable = True
    compared_ids = 0
    total_obs = 0
    for k, (l, r) in enumerate(frames):
        if set_nf_at and k in set_nf_at:
            o.set_num_features(set_nf_at[k])
            t.set_num_features(set_nf_at[k])
        n0, g0, r0 = len(o.database.log), len(t.database.log), len(o.ransac_sizes)
        cams = [0] if mono else [0, 1]
        imgs = [l] if mono else [l, r]
        o.feed_new_camera(0.05 * k, cams, imgs, masks[:len(cams)])
        t.feed_new_camera(track.CameraData(0.05 * k, cams, imgs, masks[:len(cams)]))
        sizes = o.ransac_sizes[r0:]
        lr_lmeds = (not mono) and len(sizes) > 0 and 10 <= sizes[0] <= 13     # the stereo match of this frame: its points differ
        if _lmeds(sizes):
            ids_comparable = False
        if lr_lmeds:
            continue
        oi, oc, ouv, oun = log_arrays(o.database.log[n0:])
        gi, gc, guv, gun = log_arrays(t.database.log[g0:])
        assert len(oi) == len(gi), f"frame {k}: {len(oi)} vs {len(gi)} observations"
        assert (oc == gc).all() and (ouv == guv).all(), f"frame {k}: observation positions differ"
        if len(oi):
            assert np.abs(oun - gun).max() <= 1e-5
        total_obs += len(oi)
        if ids_comparable:
            assert (oi == gi).all(), f"frame {k}: feature ids differ"
            compared_ids += len(oi)
        for cam in cams:
            ids, uv, d = t._last(cam)
            assert len(ids) == len(o.ids_last[cam])
            assert (uv == o.pts_last[cam]).all(), f"frame {k} cam {cam}: pts_last differs"
            assert (d == o.desc_last[cam]).all(), f"frame {k} cam {cam}: descriptors differ"
            if ids_comparable:
                assert (ids == np.array(o.ids_last[cam], np.uint64)).all(), f"frame {k} cam {cam}: ids_last differs"
    t.close()
    return total_obs, compared_ids


CASES = [
    ("EUROC_MAV", 8, 3, False),
    ("EUROC_MAV", 6, 5, True),
    ("TUM_VI", 5, 3, False),          # masks, equidistant model, reaches the LMedS regime at its 4th frame
    ("RS_D455", 4, 2, False),         # CLAHE
]


@pytest.fixture()
def sim_desc(sim_tracker):
    """make(cfg) -> descriptor tracker over the host simulation (sim_tracker has pointed uvio_b200.track at the sim library)."""
    from uvio_b200 import desc
    yield lambda cfg, **kw: desc.make_descriptor_tracker(cfg, **kw)


@pytest.mark.parametrize("name,n,stream,mono", CASES)
def test_descriptor_tracker_matches_oracle_on_the_host_simulation(sim_desc, name, n, stream, mono):
    obs, ids = _run_case(sim_desc, getattr(configs, name), n, stream, mono=mono)
    assert obs > 100 and ids > 50


def test_set_num_features_and_first_frames(sim_desc):
    """25 features while initialising, then the full budget (VioManager.cpp:131, VioManagerHelper.cpp:124): the extraction grid,
    K and the capacity of every list change between two feeds; pts_last / desc_last / ids_last survive the re-allocation."""
    obs, _ = _run_case(sim_desc, configs.EUROC_MAV.with_(num_features=60), 6, 7, set_nf_at={3: 300})
    assert obs > 100


def test_error_codes_match_the_reference(sim_desc):
    cfg = configs.EUROC_MAV
    t = sim_desc(cfg)
    (l, r), = frames_for(cfg, 1, stream_id=1)
    z = np.zeros_like(l)
    with pytest.raises(SystemExit):                         # sizes do not match: std::exit (TrackDescriptor.cpp:36-42)
        t.feed_new_camera(track.CameraData(0.0, [0, 1], [l], [z, z]))
    with pytest.raises(SystemExit):                         # three images on a stereo tracker (:57-58)
        t.feed_new_camera(track.CameraData(0.0, [0, 1, 2], [l, r, l], [z, z, z]))
    with pytest.raises(track.UvioError) as e:               # camera_calib.at() throws
        t.feed_new_camera(track.CameraData(0.0, [0, 9], [l, r], [z, z]))
    assert e.value.code == track.UVIO_E_CAMID
    t.feed_new_camera(track.CameraData(0.0, [0, 1], [l, r], [z, z]))
    with pytest.raises(track.UvioError) as e:
        t.feed_new_camera(track.CameraData(0.05, [0, 1], [l[:-2], r[:-2]], [z[:-2], z[:-2]]))
    assert e.value.code == track.UVIO_E_SIZE
    t.close()


# ---- the same on the GPU ------------------------------------------------------------------------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("name,n,stream,mono", CASES + [("UZHFPV_INDOOR", 6, 4, False), ("HD_1000", 3, 1, False)])
def test_descriptor_tracker_matches_oracle_on_the_gpu(name, n, stream, mono):
    from uvio_b200 import desc
    obs, ids = _run_case(lambda cfg, **kw: desc.make_descriptor_tracker(cfg, **kw), getattr(configs, name), n, stream, mono=mono)
    assert obs > 100


@pytest.mark.gpu
def test_gpu_set_num_features():
    from uvio_b200 import desc
    obs, _ = _run_case(lambda cfg, **kw: desc.make_descriptor_tracker(cfg, **kw), configs.EUROC_MAV.with_(num_features=60), 6, 7, set_nf_at={3: 300})
    assert obs > 100


@pytest.mark.gpu
def test_gpu_batch_equals_single_feeds():
    """uvio_desc_feed_batch of 7 trackers (stereo and mono mixed) = the same trackers fed one at a time."""
    from uvio_b200 import desc
    cfg = configs.EUROC_MAV
    masks = masks_for(cfg)
    n_trk, n = 7, 4
    seqs = [frames_for(cfg, n, stream_id=20 + s) for s in range(n_trk)]
    mono = [s % 3 == 2 for s in range(n_trk)]

    def msg(s, k):
        l, r = seqs[s][k]
        return track.CameraData(0.05 * k, [0], [l], masks[:1]) if mono[s] else track.CameraData(0.05 * k, [0, 1], [l, r], masks)

    a = [desc.make_descriptor_tracker(cfg) for _ in range(n_trk)]
    b = [desc.make_descriptor_tracker(cfg) for _ in range(n_trk)]
    for k in range(n):
        desc.feed_batch(a, [msg(s, k) for s in range(n_trk)])
        for s in range(n_trk):
            b[s].feed_new_camera(msg(s, k))
    for s in range(n_trk):
        assert a[s].database.log == b[s].database.log and len(a[s].database.log) > 50
        for cam in ([0] if mono[s] else [0, 1]):
            for x, y in zip(a[s]._last(cam), b[s]._last(cam)):
                assert (x == y).all()
    for t in a + b:
        t.close()


# ---- committed golden sequences (tests/golden/desc_*.npz, made by tests/golden/make_desc_golden.py): no cv2 at run time ----------
def _check_golden(make, name):
    from uvio_b200.synth import frames_checksum
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", f"desc_{name}.npz"))
    cfg = {"euroc_mav": configs.EUROC_MAV, "euroc_mono": configs.EUROC_MAV, "uzhfpv_indoor": configs.UZHFPV_INDOOR}[name]
    n, mono = int(g["n_frames"]), bool(g["mono"])
    frames = frames_for(cfg, n, stream_id=int(g["stream_id"]))
    assert frames_checksum([f for lr in frames for f in lr]) == int(g["frames_checksum"]), "synthetic generator drifted"
    masks = masks_for(cfg)
    t = make(cfg)
    ok_until = int(g["ids_ok_until"])
    total = 0
    for k, (l, r) in enumerate(frames):
        cams = [0] if mono else [0, 1]
        t.feed_new_camera(track.CameraData(0.05 * k, cams, [l] if mono else [l, r], masks[:len(cams)]))
        sizes = g[f"ransac_sizes_{k}"]
        if (not mono) and len(sizes) and 10 <= sizes[0] <= 13:
            continue
        obs = t.new_observations()
        assert np.array_equal(obs["cam_id"].astype(np.int8), g[f"cams_{k}"])
        assert np.array_equal(np.c_[obs["u"], obs["v"]].reshape(-1, 2), g[f"uv_{k}"].reshape(-1, 2)), f"{name} frame {k}: positions differ"
        if len(obs):
            assert np.abs(np.c_[obs["un"], obs["vn"]] - g[f"uvn_{k}"]).max() <= 1e-5
        if k < ok_until:
            assert np.array_equal(obs["id"].astype(np.int64), g[f"ids_{k}"]), f"{name} frame {k}: ids differ"
        for c in cams:
            ids, uv, d = t._last(c)
            assert np.array_equal(uv, g[f"last_uv_{k}_{c}"]) and np.array_equal(d, g[f"last_desc_{k}_{c}"]), f"{name} frame {k} cam {c}"
            if k < ok_until:
                assert np.array_equal(ids.astype(np.int64), g[f"last_ids_{k}_{c}"])
        total += len(obs)
    t.close()
    assert total > 300 and ok_until >= 3


@pytest.mark.parametrize("name", ["euroc_mav", "euroc_mono", "uzhfpv_indoor"])
def test_golden_sequences_on_the_host_simulation(sim_desc, name):
    _check_golden(sim_desc, name)


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["euroc_mav", "euroc_mono", "uzhfpv_indoor"])
def test_golden_sequences_on_the_gpu(name):
    from uvio_b200 import desc
    _check_golden(lambda cfg: desc.make_descriptor_tracker(cfg), name)


# ---- edge cases: the same code paths the reference takes on degenerate input -----------------------------------------------------
def _edge_sequence(make):
    """blank frames (no corners at all), a frame masked out completely, then normal frames again: every list is empty for a
    while (knnMatch on empty sets, robust_match below 10 matches, 'first frame' taken again after the track is lost)."""
    from oracle.trackdescriptor_cv2 import make_descriptor_oracle
    cfg = configs.EUROC_MAV
    (a0, b0), (a1, b1), (a2, b2), (a3, b3), (a4, b4) = frames_for(cfg, 5, stream_id=9)
    z = np.zeros_like(a0)
    full = np.full_like(a0, 255)
    blank = np.full_like(a0, 90)
    seq = [(a0, b0, z, z), (blank, blank, z, z), (a1, b1, z, z), (a2, b2, full, full), (a2, b2, z, full), (a3, b3, z, z), (a4, b4, z, z)]
    o = make_descriptor_oracle(cfg)
    t = make(cfg)
    for k, (l, r, ml, mr) in enumerate(seq):
        n0, g0, r0 = len(o.database.log), len(t.database.log), len(o.ransac_sizes)
        o.feed_new_camera(0.05 * k, [0, 1], [l, r], [ml, mr])
        t.feed_new_camera(track.CameraData(0.05 * k, [0, 1], [l, r], [ml, mr]))
        assert not any(10 <= s <= 13 for s in o.ransac_sizes[r0:]), "the edge sequence must stay out of the LMedS regime"
        assert o.database.log[n0:] == [tuple(x) for x in t.database.log[g0:]] or _same_log(o.database.log[n0:], t.database.log[g0:]), f"frame {k}"
        for cam in (0, 1):
            ids, uv, d = t._last(cam)
            assert len(ids) == len(o.ids_last[cam]), f"frame {k} cam {cam}: {len(ids)} vs {len(o.ids_last[cam])}"
            assert (ids == np.array(o.ids_last[cam], np.uint64)).all() and (uv == o.pts_last[cam]).all() and (d == o.desc_last[cam]).all()
    assert len(o.database.log) > 50
    t.close()


def _same_log(a, b):
    oi, oc, ouv, oun = log_arrays(a)
    gi, gc, guv, gun = log_arrays(b)
    return len(oi) == len(gi) and (oi == gi).all() and (oc == gc).all() and (ouv == guv).all() and (len(oi) == 0 or np.abs(oun - gun).max() <= 1e-5)


def test_blank_and_masked_frames_on_the_host_simulation(sim_desc):
    _edge_sequence(sim_desc)


@pytest.mark.gpu
def test_blank_and_masked_frames_on_the_gpu():
    from uvio_b200 import desc
    _edge_sequence(lambda cfg: desc.make_descriptor_tracker(cfg))


def test_tiny_image_keeps_nothing(sim_desc):
    """An image no larger than twice ORB's 31-px edge: compute() drops every key point, nothing is ever tracked."""
    cfg = configs.EUROC_MAV.with_(width=60, height=48, num_features=20)
    t = sim_desc(cfg)
    rng = np.random.default_rng(5)
    for k in range(3):
        l, r = (rng.integers(0, 256, (48, 60), dtype=np.uint8) for _ in range(2))
        t.feed_new_camera(track.CameraData(0.05 * k, [0, 1], [l, r], [np.zeros_like(l), np.zeros_like(l)]))
        assert len(t.new_observations()) == 0 and len(t._last(0)[0]) == 0
    t.close()


def _identical_frames(make):
    """The same frame delivered twice: every temporal match joins identical positions, cv::findFundamentalMat asserts
    (matrix.cpp:764, as in the KLT tracker's case: DESIGN.md section 5) and the reference, which does not catch it
    (TrackDescriptor.cpp:533), terminates.  This tracker returns UVIO_OK and hands every id over (the zero-motion model fits all)."""
    import cv2
    from oracle.trackdescriptor_cv2 import make_descriptor_oracle
    cfg = configs.EUROC_MAV
    (l, r), = frames_for(cfg, 1, stream_id=11)
    z = np.zeros_like(l)
    o = make_descriptor_oracle(cfg)
    o.feed_new_camera(0.0, [0, 1], [l, r], [z, z])
    with pytest.raises(cv2.error):
        o.feed_new_camera(0.05, [0, 1], [l, r], [z, z])
    t = make(cfg)
    t.feed_new_camera(track.CameraData(0.0, [0, 1], [l, r], [z, z]))
    ids0 = t._last(0)[0].copy()
    t.feed_new_camera(track.CameraData(0.05, [0, 1], [l, r], [z, z]))
    obs = t.new_observations()
    assert len(obs) == 2 * len(ids0) and len(ids0) > 20
    kept = np.isin(obs["id"][::2], ids0).mean()
    assert kept > 0.5, "identical frames: most features keep their ids"
    t.close()


def test_identical_frames_where_the_reference_would_terminate(sim_desc):
    _identical_frames(sim_desc)


@pytest.mark.gpu
def test_identical_frames_on_the_gpu():
    from uvio_b200 import desc
    _identical_frames(lambda cfg: desc.make_descriptor_tracker(cfg))


# ---- a non-stereo tracker fed two cameras in one message: feed_monocular per camera (TrackDescriptor.cpp:50-55) -------------------
def _two_mono_cameras(make):
    """The reference runs the cameras through cv::parallel_for_, each drawing ids from the shared counter; camera by camera in
    message order is the order that does not change from run to run (the oracle's, and this tracker's)."""
    from oracle.trackdescriptor_cv2 import make_descriptor_oracle
    cfg = configs.EUROC_MAV.with_(use_stereo=False)
    frames = frames_for(cfg, 5, stream_id=13)
    masks = masks_for(cfg)
    o = make_descriptor_oracle(cfg)
    t = make(cfg)
    for k, (l, r) in enumerate(frames):
        n0, g0, r0 = len(o.database.log), len(t.database.log), len(o.ransac_sizes)
        o.feed_new_camera(0.05 * k, [0, 1], [l, r], masks)
        t.feed_new_camera(track.CameraData(0.05 * k, [0, 1], [l, r], masks))
        assert not any(10 <= s <= 13 for s in o.ransac_sizes[r0:])
        assert _same_log(o.database.log[n0:], t.database.log[g0:]), f"frame {k}"
        for cam in (0, 1):
            ids, uv, d = t._last(cam)
            assert (ids == np.array(o.ids_last[cam], np.uint64)).all() and (uv == o.pts_last[cam]).all() and (d == o.desc_last[cam]).all()
    assert len(o.database.log) > 800
    t.close()


def test_two_cameras_of_a_non_stereo_tracker_host_simulation(sim_desc):
    _two_mono_cameras(sim_desc)


@pytest.mark.gpu
def test_two_cameras_of_a_non_stereo_tracker_gpu():
    from uvio_b200 import desc
    _two_mono_cameras(lambda cfg: desc.make_descriptor_tracker(cfg))
