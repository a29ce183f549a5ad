"""bench.py's bookkeeping that runs after the timed legs (no GPU needed): whatever kernel dominates and whichever captures are
present under profiles/, the roofline entries are produced without raising -- a failure there would lose a whole bench run."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
from uvio_b200 import track


def _prof(**ms):
    return {k: (ms.get(k, 0.0), 10 if ms.get(k, 0.0) else 0) for k in track.KERNEL_NAMES}


def _stats():
    return {"lk_point_levels": 5_000_000.0, "lk_iterations": 30_000_000.0, "fast_pixels": 1e8, "launches": 1000, "h2d_bytes": 1e9, "d2h_bytes": 1e7}


def test_roofline_entries_for_every_dominant_kernel(monkeypatch):
    for name in ("euroc_mav", "rs_d455", "hd_1000"):
        cfg = bench.workload_config(name)
        for dominant in ("k_lk", "k_preprocess", "k_fast", "k_dv_logic", "k_ransac"):
            allk = _prof(k_lk=50.0, k_preprocess=20.0, k_pyrdown=10.0, k_fast=5.0, k_dv_logic=8.0, k_ransac=4.0, k_clahe_tiles=1.0, k_clahe_apply=1.0)
            allk[dominant] = (500.0, 10)
            alone = _prof(k_preprocess=2.0, k_pyrdown=3.0, k_clahe_tiles=0.5, k_clahe_apply=0.5)
            e = bench.roofline_entry((_prof(k_lk=40.0), allk, _stats(), 64, alone, 32), _stats(), 640, 64, cfg, {"sm_mhz": 1965.0})
            assert e["kernel"] == dominant and "bound" in e
            ip = e["image_pass"]
            assert ip is not None and ip["frac"] > 0 and ip["parts"]["equalise"]["GB_per_s"] > 0
            json.dumps(e)                                    # everything in the line must serialise
    # no capture under profiles/ at all: the entries still come out, with the fractions that need one set to None
    monkeypatch.setattr(bench, "_load_profile", lambda name: (None, None))
    e = bench.roofline_entry((_prof(k_lk=40.0), _prof(k_lk=50.0, k_preprocess=20.0), _stats(), 64, _prof(), 32), _stats(), 640, 64,
                             bench.workload_config("euroc_mav"), None)
    assert e["kernel"] == "k_lk" and e["frac"] is None and e["image_pass"] is None


def test_every_config_name_resolves():
    for name in bench.BENCH_CONFIGS:
        cfg = bench.workload_config(name)
        d = bench.config_dict(cfg, bench.BENCH_CONFIGS[name][1], 32)
        assert "workload" in d and cfg.width > 0
