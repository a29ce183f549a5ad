"""Batched GPU triangulation (include/uvio_tri.h) against the restatement of ov_core::FeatureInitializer
(oracle/triangulation_oracle.py) on a synthetic stereo trajectory: per feature the success flags of single_triangulation(_1d)
and single_gaussnewton, the anchor, p_FinA and p_FinG.

Tolerance: the two sides solve the same 3x3 systems with different (backward-stable) solvers and both replicate the
reference's float casts, so positions are required to agree to 1e-6 relative (observed: ~1e-12); flags must be identical
except for features within 1e-6 relative of a threshold, of which the scene has none.
Runs on the host-simulation build (CPU) and on the GPU through the same C ABI, fed both from host CSR arrays and from the
device-resident result of a feature-store query."""
import numpy as np
import pytest

from uvio_b200 import track, tri
from uvio_b200.featdb import DeviceFeatureDatabase

RTOL = 1e-6


def _rot(ax, ang):
    ax = np.asarray(ax, float) / np.linalg.norm(ax)
    K = np.array([[0, -ax[2], ax[1]], [ax[2], 0, -ax[0]], [-ax[1], ax[0], 0]])
    return np.eye(3) + np.sin(ang) * K + (1 - np.cos(ang)) * K @ K


def _scene(seed=4, n_feat=160, n_clones=11):
    """clones[cam][t] = (R_GtoC, p_CinG) for a stereo rig moving sideways and forwards; features = list of (id, {cam: [(t, uvn)]})"""
    rng = np.random.default_rng(seed)
    times = [0.05 * k for k in range(n_clones)]
    clones = {0: {}, 1: {}}
    for k, t in enumerate(times):
        R_GtoI = _rot([0.2, 1.0, 0.1], 0.02 * k)
        p_IinG = np.array([0.12 * k, 0.01 * k, 0.05 * k])
        for cam, off in ((0, np.array([0.0, 0.0, 0.0])), (1, np.array([0.11, 0.0, 0.0]))):
            R_ItoC = _rot([0, 0, 1], 0.01 * cam)
            R_GtoC = R_ItoC @ R_GtoI
            clones[cam][t] = (R_GtoC, p_IinG + R_GtoI.T @ off)
    feats = []
    for i in range(n_feat):
        kind = i % 8
        depth = rng.uniform(2, 12) if kind < 5 else (rng.uniform(80, 200) if kind == 5 else (rng.uniform(0.02, 0.08) if kind == 6 else rng.uniform(25, 55)))
        P = np.array([rng.uniform(-1.5, 2.5), rng.uniform(-1, 1), depth])
        first = int(rng.integers(0, n_clones - 3))
        last = int(rng.integers(first + 2, n_clones))
        obs = {}
        for cam in ((0, 1) if i % 3 else (0,)):
            lst = []
            for k in range(first, last + 1 if cam == 0 else max(first + 1, last - (i % 2))):
                R, p = clones[cam][times[k]]
                pc = R @ (P - p)
                if pc[2] <= 0.01:
                    continue
                uvn = (pc[:2] / pc[2] + rng.normal(0, 0.6 / 460, 2)).astype(np.float32)
                lst.append((times[k], uvn))
            if lst:
                obs[cam] = lst
        if obs:
            feats.append((1000 + i, obs))
    return clones, feats


def _csr(feats):
    row0, row_cam, off, t, uvn = [0], [], [0], [], []
    for _fid, obs in feats:
        for cam in sorted(obs):
            row_cam.append(cam)
            for tt, u in obs[cam]:
                t.append(tt)
                uvn.append(u)
            off.append(len(t))
        row0.append(len(row_cam))
    return np.array(row0, np.int32), np.array(row_cam, np.int32), np.array(off, np.int32), np.array(t), np.array(uvn, np.float32).reshape(-1, 2)


def _oracle(feats, clones, **opt):
    from oracle.triangulation_oracle import Feat, FeatureInitializer
    init = FeatureInitializer(**opt)
    out = []
    for _fid, obs in feats:
        f = Feat({c: [t for t, _u in lst] for c, lst in obs.items()}, {c: [u for _t, u in lst] for c, lst in obs.items()})
        ok_t = init.single_triangulation_1d(f, clones) if init.o["triangulate_1d"] else init.single_triangulation(f, clones)
        ok_r = bool(ok_t) and init.o["refine_features"] and init.single_gaussnewton(f, clones)
        out.append((bool(ok_t), bool(ok_r), f))
    return out


def _compare(res, ref, what):
    n_ok = 0
    worst = 0.0
    for k, (r, (ok_t, ok_r, f)) in enumerate(zip(res, ref)):
        assert int(r["ok_triangulation"]) == int(ok_t), f"{what} feature {k}: triangulation flag {r['ok_triangulation']} vs {ok_t}"
        assert int(r["ok_refine"]) == int(ok_r), f"{what} feature {k}: refine flag {r['ok_refine']} vs {ok_r}"
        assert int(r["anchor_cam_id"]) == f.anchor_cam_id and r["anchor_clone_timestamp"] == f.anchor_clone_timestamp, f"{what} feature {k}: anchor"
        if ok_t:
            scale = max(1.0, float(np.abs(f.p_FinA).max()))
            d = float(np.abs(r["p_FinA"] - f.p_FinA).max()) / scale
            worst = max(worst, d)
            assert d <= RTOL, f"{what} feature {k}: p_FinA differs by {d} relative"
            if ok_r or not f_refine(ref):
                assert float(np.abs(r["p_FinG"] - f.p_FinG).max()) / max(1.0, float(np.abs(f.p_FinG).max())) <= RTOL, f"{what} feature {k}: p_FinG"
            n_ok += 1
    return n_ok, worst


def f_refine(ref):
    return any(r[1] for r in ref)


def _run_all():
    clones, feats = _scene()
    row0, row_cam, off, t, uvn = _csr(feats)
    total = 0
    for opt in ({}, {"refine_features": False}, {"triangulate_1d": True}, {"max_cond_number": 1500.0, "max_baseline": 25.0}):
        res = tri.triangulate(row0, row_cam, off, t, uvn, clones, tri.options(**{k: (int(v) if isinstance(v, bool) else v) for k, v in opt.items()}))
        ref = _oracle(feats, clones, **opt)
        n_ok, worst = _compare(res, ref, f"options {opt}")
        flags = [(int(r["ok_triangulation"]), int(r["ok_refine"])) for r in res]
        assert n_ok >= 10 and len(set(flags)) >= 2, f"options {opt}: the scene does not exercise both outcomes ({set(flags)}, {n_ok} ok)"
        print(f"options {opt}: {n_ok} of {len(res)} triangulated, flags {sorted(set(flags))}, worst relative difference {worst:.2e}")
        total += n_ok
    # the same features through the device-resident result of a feature-store query (no host copy of the measurements)
    db = DeviceFeatureDatabase(max_observations=1 << 14, max_features=1 << 10)
    by_time = {}
    for fid, obs in feats:
        for cam, lst in obs.items():
            for tt, u in lst:
                by_time.setdefault(tt, []).append((fid, cam, u))
    for tt in sorted(by_time):
        rows = sorted(by_time[tt], key=lambda r: (r[1], r[0]))            # all left, then all right, as the tracker reports them
        o = np.zeros(len(rows), track.OBS_DTYPE)
        for i, (fid, cam, u) in enumerate(rows):
            o[i] = (fid, cam, 0, 0, u[0], u[1])
        db.update(tt, o)
    got = db.features_containing_older(1e9)                                # every feature, insertion order
    order = [f.featid for f in got]
    res = tri.triangulate_last_query(db, clones, tri.options(), n_features=len(got))
    ref = dict((fid, r) for (fid, _o), r in zip(feats, _oracle(feats, clones)))
    _compare(res, [ref[i] for i in order], "from the feature store")
    # a measurement without a clone pose: clonesCAM.at() throws in the reference
    bad = {c: {tt: v for tt, v in d.items() if tt != 0.1} for c, d in clones.items()}
    with pytest.raises(track.UvioError):
        tri.triangulate(row0, row_cam, off, t, uvn, bad)
    assert len(tri.triangulate(np.zeros(1, np.int32), np.zeros(0, np.int32), np.zeros(1, np.int32), np.zeros(0), np.zeros((0, 2), np.float32), clones)) == 0
    return total


def test_triangulation_host_simulation(sim_tracker):
    sim_tracker
    assert _run_all() > 150


@pytest.mark.gpu
def test_triangulation_gpu(built_lib):
    assert _run_all() > 150
