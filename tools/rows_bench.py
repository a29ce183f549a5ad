"""Measurements of the rows after the tracker on one B200, with the CPU path beside them:

  n1  the per-frame feature-store calls of the estimator (update with the tracker's observations, features_not_containing_newer,
      mark, cleanup, cleanup_measurements: VioManager.cpp:369, :579-591) for S stores, batched (one launch per call for all
      stores) and one store at a time; CPU arm = the C++ mirror of ov_core::FeatureDatabase (tools/featdb_cpu_bench.cpp).
  n3  FeatureInitializer::single_triangulation + single_gaussnewton of F features in one launch; CPU arm = the numpy
      restatement (oracle/triangulation_oracle.py), one thread.

    python tools/rows_bench.py > profiles/r2_rows_n1_n3.json
"""
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np

from uvio_b200 import featdb, track, tri
from uvio_b200.featdb import DeviceFeatureDatabase


def frames_of(name="euroc_mav"):
    g = np.load(os.path.join(ROOT, "tests", "golden", f"seq_{name}.npz"))
    out = []
    for k in range(int(g["n_frames"])):
        ids, cams, uv, uvn = g[f"ids_{k}"], g[f"cams_{k}"], g[f"uv_{k}"], g[f"uvn_{k}"]
        o = np.zeros(len(ids), track.OBS_DTYPE)
        o["id"], o["cam_id"], o["u"], o["v"], o["un"], o["vn"] = ids, cams, uv[:, 0], uv[:, 1], uvn[:, 0], uvn[:, 1]
        out.append(o)
    return out


def n1(S=64, repeat=6):
    frames = frames_of()
    n_obs = sum(len(o) for o in frames)
    res = {"stores": S, "frames_per_pass": len(frames), "observations_per_frame": n_obs / len(frames)}
    for mode in ("batched", "single"):
        best = None
        for _ in range(repeat):
            dbs = [DeviceFeatureDatabase(max_observations=1 << 14, max_features=1 << 11) for _ in range(S)]
            t0 = time.perf_counter()
            for k, o in enumerate(frames):
                ts = [0.05 * k] * S
                if mode == "batched":
                    featdb.update_batch(dbs, ts, [o] * S)
                    counts = featdb.query_batch(dbs, featdb.NOT_CONTAINING_NEWER, ts, skip_deleted=True, fetch=False)
                    featdb.cleanup_batch(dbs, 0)
                    if k > 5:
                        featdb.cleanup_batch(dbs, 1, [t - 0.25 for t in ts])
                else:
                    for d in dbs:
                        d.update(0.05 * k, o)
                        track.lib().uvio_featdb_query(d._h, featdb.NOT_CONTAINING_NEWER, 0.05 * k, 0, 0, 1, None, None, None)
                        d.cleanup()
                        if k > 5:
                            d.cleanup_measurements(0.05 * k - 0.25)
            sec = time.perf_counter() - t0
            best = sec if best is None else min(best, sec)
            for d in dbs:
                d.close()
        res[mode] = {"seconds_per_pass": best, "store_frames_per_sec": S * len(frames) / best, "update_feature_calls_per_sec": S * n_obs / best,
                     "us_per_step_of_all_stores": 1e6 * best / len(frames)}
    # CPU arm: the C++ mirror, one store, one thread
    exe = "/tmp/featdb_cpu_bench"
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-I", os.path.join(ROOT, "tests", "cpp"), "-o", exe, os.path.join(ROOT, "tools", "featdb_cpu_bench.cpp")])
    raw = "/tmp/featdb_obs.bin"
    with open(raw, "wb") as f:
        for o in frames:
            f.write(np.int32(len(o)).tobytes())
            f.write(o.tobytes())
    out = subprocess.run([exe, raw, str(len(frames)), "200"], capture_output=True, text=True, check=True).stdout
    cpu = json.loads(out)
    res["cpu_baseline"] = {"value": cpu["frames_per_sec"], "unit": "store-frames/s", "cores": 1, "kind": "port",
                           "what": "tests/cpp/ov_core_mirror.h FeatureDatabase (the reference's unordered_map<shared_ptr<Feature>> statement by statement), "
                                   "same calls, one store, one thread, g++ -O2", "host_cores": os.cpu_count()}
    return res


def n3(copies=128, repeat=5):
    from test_tri import _csr, _oracle, _scene
    clones, feats = _scene()
    big = [(fid + 10000 * c, obs) for c in range(copies) for fid, obs in feats]
    row0, row_cam, off, t, uvn = _csr(big)
    tri.triangulate(row0, row_cam, off, t, uvn, clones)           # warm-up
    best = None
    for _ in range(repeat):
        t0 = time.perf_counter()
        res = tri.triangulate(row0, row_cam, off, t, uvn, clones)
        sec = time.perf_counter() - t0
        best = sec if best is None else min(best, sec)
    t0 = time.perf_counter()
    ref = _oracle(feats, clones)
    cs = time.perf_counter() - t0
    return {"features": len(big), "measurements": int(len(t)), "seconds_per_call_incl_transfers": best, "features_per_sec": len(big) / best,
            "triangulated": int(sum(int(r["ok_triangulation"]) for r in res)),
            "cpu_baseline": {"value": len(feats) / cs, "unit": "features/s", "cores": 1, "kind": "port",
                             "what": "oracle/triangulation_oracle.py (numpy restatement of FeatureInitializer.cpp:30-421), one thread", "features": len(feats)}}


if __name__ == "__main__":
    print(json.dumps({"n1_feature_store": n1(), "n3_triangulation": n3()}, indent=1))
