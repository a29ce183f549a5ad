"""The device-side feature-track store (include/uvio_featdb.h) against the restatement of ov_core::FeatureDatabase
(oracle/featdb_oracle.py): a golden observation stream is replayed into both, and after every frame the query sequence the
estimator makes (ov_msckf/src/core/VioManager.cpp:369-375, :462-466, :579-591) is issued on both; feature sets, featids,
to_delete flags and every per-camera measurement list must be identical.  Runs on the host-simulation build (CPU) and on the
GPU through the same C ABI."""
import numpy as np
import pytest

from uvio_b200 import track
from uvio_b200.featdb import DeviceFeatureDatabase

from golden_util import load


def _same(dev_feats, ora_feats, what):
    d = {f.key if hasattr(f, "key") else None: f for f in dev_feats}
    assert len(d) == len(dev_feats), f"{what}: a feature was returned twice"
    o = {k: f for k, f in ora_feats}
    assert set(d) == set(o), f"{what}: feature sets differ: only device {sorted(set(d) - set(o))[:5]}, only oracle {sorted(set(o) - set(d))[:5]}"
    for k, f in d.items():
        g = o[k]
        assert f.featid == g.featid and f.to_delete == g.to_delete, f"{what}: feature {k}"
        cams = {c for c, t in g.timestamps.items() if len(t)}
        assert set(f.timestamps) == cams, f"{what}: feature {k} cameras {set(f.timestamps)} vs {cams}"
        for c in cams:
            assert np.array_equal(f.timestamps[c], np.array(g.timestamps[c], np.float64)), f"{what}: feature {k} cam {c} timestamps"
            assert np.array_equal(f.uvs[c], np.array(g.uvs[c], np.float32).reshape(-1, 2)), f"{what}: feature {k} cam {c} uvs"
            assert np.array_equal(f.uvs_norm[c], np.array(g.uvs_norm[c], np.float32).reshape(-1, 2)), f"{what}: feature {k} cam {c} uvs_norm"


def _keyed(ora, feats):
    inv = {id(f): k for k, f in ora.features_idlookup.items()}
    return [(inv.get(id(f), f.featid), f) for f in feats]


def _replay(name, n_frames, window=5):
    from oracle.featdb_oracle import FeatureDatabase
    g = load(name)
    n_frames = min(n_frames, int(g["n_frames"]))
    dev, ora = DeviceFeatureDatabase(max_observations=1 << 15, max_features=1 << 12), FeatureDatabase()
    rng = np.random.default_rng(3)
    renamed = False
    checked = 0
    for k in range(n_frames):
        t = 0.05 * k
        ids, cams, uv, uvn = g[f"ids_{k}"], g[f"cams_{k}"], g[f"uv_{k}"], g[f"uvn_{k}"]
        obs = np.zeros(len(ids), track.OBS_DTYPE)
        obs["id"], obs["cam_id"], obs["u"], obs["v"], obs["un"], obs["vn"] = ids, cams, uv[:, 0], uv[:, 1], uvn[:, 0], uvn[:, 1]
        dev.update(t, obs)
        for i in range(len(ids)):
            ora.update_feature(int(ids[i]), t, int(cams[i]), np.float32(uv[i, 0]), np.float32(uv[i, 1]), np.float32(uvn[i, 0]), np.float32(uvn[i, 1]))
        assert dev.size() == ora.size() and dev.num_measurements() == sum(len(x) for f in ora.features_idlookup.values() for x in f.timestamps.values())
        # VioManager.cpp:369: features lost in the newest frame, not yet used
        lost_d, lost_o = dev.features_not_containing_newer(t, False, True), ora.features_not_containing_newer(t, False, True)
        _same(lost_d, _keyed(ora, lost_o), f"frame {k} not_containing_newer")
        # the updaters mark what they used (UpdaterMSCKF.cpp: feat->to_delete = true): here a random half
        used = [f for f in lost_o if rng.random() < 0.5]
        for f in used:
            f.to_delete = True
        dev.mark_to_delete([kk for kk, f in _keyed(ora, used)])
        t_marg = 0.05 * (k - window)
        if k > window:                                    # :373: features seen at the marginalisation time
            _same(dev.features_containing(t_marg, False, True), _keyed(ora, ora.features_containing(t_marg, False, True)), f"frame {k} containing")
            _same(dev.features_containing_older(t_marg, False, False), _keyed(ora, ora.features_containing_older(t_marg, False, False)),
                  f"frame {k} containing_older")
        if len(ids) and k % 3 == 1:                       # :462-466 get_feature of a SLAM landmark's id
            fid = int(ids[len(ids) // 2])
            a, b = dev.get_feature(fid), ora.get_feature(fid)
            assert (a is None) == (b is None)
            if b is not None:
                _same([a], [(fid, b)], f"frame {k} get_feature")
        if k == 6 and len(ids) and not renamed:           # TrackBase::change_feat_id's database half: key stays, featid changes
            fid = int(ids[0])
            dev.change_feat_id(fid, fid + 500000)
            ora.change_feat_id_database_half(fid, fid + 500000)
            renamed = True
        dev.cleanup()                                     # :579
        ora.cleanup()
        if k > window:                                    # :589
            dev.cleanup_measurements(t_marg)
            ora.cleanup_measurements(t_marg)
        if k == 9:
            dev.cleanup_measurements_exact(0.05 * 8)
            ora.cleanup_measurements_exact(0.05 * 8)
        if k % 4 == 3:                                    # remove = true variant
            _same(dev.features_not_containing_newer(t, True, False), _keyed(ora, ora.features_not_containing_newer(t, True, False)), f"frame {k} remove")
        _same(list(dev.get_internal_data().values()), list(ora.features_idlookup.items()), f"frame {k} whole database")
        checked += 1
    return checked, dev, ora


def _edge_cases():
    dev = DeviceFeatureDatabase(max_observations=64, max_features=16)
    assert dev.size() == 0 and dev.features_containing(0.0) == [] and dev.get_feature(5) is None
    dev.cleanup()
    dev.cleanup_measurements(1.0)
    o = np.zeros(3, track.OBS_DTYPE)
    o["id"], o["cam_id"], o["u"] = [7, 7, 7], [0, 1, 0], [1.0, 2.0, 3.0]       # (7, cam 0) twice in one call: two update_feature calls
    dev.update(0.5, o)
    f = dev.get_feature(7)
    assert f.timestamps[0].tolist() == [0.5, 0.5] and f.uvs[0][:, 0].tolist() == [1.0, 3.0] and f.uvs[1][:, 0].tolist() == [2.0]
    with pytest.raises(track.UvioError) as e:                                   # more features than the store was sized for
        big = np.zeros(40, track.OBS_DTYPE)
        big["id"] = np.arange(100, 140)
        dev.update(0.6, big)
    assert e.value.code == track.UVIO_E_SIZE and dev.size() == 1 and dev.num_measurements() == 3      # unchanged
    bad = np.zeros(1, track.OBS_DTYPE)
    bad["cam_id"] = 9
    with pytest.raises(track.UvioError):
        dev.update(0.7, bad)
    dev.update(0.7, o[:1])
    assert dev.get_feature(7).timestamps[0].tolist() == [0.5, 0.5, 0.7]


def test_featdb_host_simulation(sim_tracker):
    sim_tracker                                            # the fixture points track.lib() at the host-simulation library
    assert _replay("euroc_mav", 24)[0] == 24
    assert _replay("tum_vi", 12)[0] == 12
    _edge_cases()


@pytest.mark.gpu
@pytest.mark.parametrize("name,n", [("euroc_mav", 24), ("rs_d455", 12), ("euroc_mono", 16), ("hd_1000", 12)])
def test_featdb_gpu(built_lib, name, n):
    assert _replay(name, n)[0] == n


@pytest.mark.gpu
def test_featdb_gpu_edge_cases_and_tracker_feed(built_lib):
    _edge_cases()
    from uvio_b200 import configs
    from uvio_b200.track import CameraData, make_tracker
    from util import frames_for, masks_for
    cfg = configs.EUROC_MAV
    trk, dev = make_tracker(cfg), DeviceFeatureDatabase()
    frames, masks = frames_for(cfg, 5), masks_for(cfg)
    for k, (l, r) in enumerate(frames):
        trk.feed_new_camera(CameraData(0.05 * k, [0, 1], [l, r], masks))
        dev.update_from_tracker(trk, 0.05 * k)                      # what TrackKLTGpu::replay_observations does on the host
    host = trk.get_feature_database()                               # the Python mirror filled by the binding
    got = dev.get_internal_data()
    assert set(got) == set(host.features_idlookup)
    for k, f in got.items():
        for c in f.timestamps:
            assert np.array_equal(f.uvs[c], np.array(host.features_idlookup[k].uvs[c], np.float32))


# ---- batched calls: n stores in one launch = the same stores driven one at a time ------------------------------------------------
def _batch_equals_single(n_frames=10):
    from uvio_b200 import featdb
    names = ["euroc_mav", "tum_vi", "euroc_mono", "rs_d455", "uzhfpv_indoor"]
    gs = [load(nm) for nm in names]
    a = [DeviceFeatureDatabase(max_observations=1 << 14, max_features=1 << 11) for _ in names]
    b = [DeviceFeatureDatabase(max_observations=1 << 14, max_features=1 << 11) for _ in names]
    for k in range(n_frames):
        ts = [0.05 * k + 0.001 * i for i in range(len(names))]
        lists = []
        for g in gs:
            ids, cams, uv, uvn = g[f"ids_{k}"], g[f"cams_{k}"], g[f"uv_{k}"], g[f"uvn_{k}"]
            o = np.zeros(len(ids), track.OBS_DTYPE)
            o["id"], o["cam_id"], o["u"], o["v"], o["un"], o["vn"] = ids, cams, uv[:, 0], uv[:, 1], uvn[:, 0], uvn[:, 1]
            lists.append(o)
        featdb.update_batch(a, ts, lists)
        for d, t, o in zip(b, ts, lists):
            d.update(t, o)
        lost_a = featdb.query_batch(a, featdb.NOT_CONTAINING_NEWER, ts, skip_deleted=True)
        for d, t, la in zip(b, ts, lost_a):
            lb = d.features_not_containing_newer(t, False, True)
            _same(la, [(f.key, f) for f in lb], f"frame {k} batch query")
        for da, db_, la in zip(a, b, lost_a):
            keys = [f.key for f in la][::2]
            da.mark_to_delete(keys)
            db_.mark_to_delete(keys)
        featdb.cleanup_batch(a, 0)
        for d in b:
            d.cleanup()
        if k > 4:
            tm = [t - 0.25 for t in ts]
            featdb.cleanup_batch(a, 1, tm)
            for d, t in zip(b, tm):
                d.cleanup_measurements(t)
        counts = featdb.query_batch(a, featdb.ALL, ts, fetch=False)
        for da, db_, c in zip(a, b, counts):
            assert da.size() == db_.size() == c[0] and da.num_measurements() == db_.num_measurements() == c[2]
            _same(list(da.get_internal_data().values()), [(f.key, f) for f in db_.get_internal_data().values()], f"frame {k} whole store")
    o = np.zeros(2, track.OBS_DTYPE)
    o["id"] = [5, 5]                                        # the same (id, camera) twice in one list: refused by the batched call
    with pytest.raises(track.UvioError):
        featdb.update_batch(a[:1], [9.0], [o])
    assert sum(d.size() for d in a) > 100


def test_featdb_batched_calls_host_simulation(sim_tracker):
    sim_tracker
    _batch_equals_single()


@pytest.mark.gpu
def test_featdb_batched_calls_gpu(built_lib):
    _batch_equals_single()


# ---- stores fed straight from the trackers' last batched feed (both tracker kinds) ------------------------------------------------
def _from_trackers(make_klt, make_desc):
    from uvio_b200 import configs, featdb
    from uvio_b200.track import CameraData
    from util import frames_for, masks_for
    cfg = configs.EUROC_MAV
    masks = masks_for(cfg)
    for make in (make_klt, make_desc):
        trks = [make(cfg) for _ in range(3)]
        dbs = [DeviceFeatureDatabase(max_observations=1 << 13, max_features=1 << 11) for _ in trks]
        seqs = [frames_for(cfg, 4, stream_id=30 + s) for s in range(3)]
        for k in range(4):
            for t, sq in zip(trks, seqs):
                t.feed_new_camera(CameraData(0.05 * k, [0, 1], list(sq[k]), masks))
            featdb.update_from_trackers(dbs, trks, [0.05 * k] * 3)
        for t, d in zip(trks, dbs):
            host = t.get_feature_database()                              # the Python mirror the binding fills with the same observations
            got = d.get_internal_data()
            assert set(got) == set(host.features_idlookup) and len(got) > 20
            for key, f in got.items():
                for c in f.timestamps:
                    assert np.array_equal(f.uvs[c], np.array(host.features_idlookup[key].uvs[c], np.float32))
        for t in trks:
            t.close()


def test_stores_fed_from_trackers_host_simulation(sim_tracker):
    from uvio_b200 import desc
    _from_trackers(lambda cfg: sim_tracker(cfg), lambda cfg: desc.make_descriptor_tracker(cfg))


@pytest.mark.gpu
def test_stores_fed_from_trackers_gpu(built_lib):
    from uvio_b200 import desc
    from uvio_b200.track import make_tracker
    _from_trackers(lambda cfg: make_tracker(cfg), lambda cfg: desc.make_descriptor_tracker(cfg))
