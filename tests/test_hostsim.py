"""CPU: (1) the kernel bodies of uvio_b200/csrc/*_body.h, compiled with g++ and run one-lane (tests/hostsim/hostsim.cpp),
against the stage goldens; (2) the host logic of the tracker (engine.cpp + abi.cpp over a stand-in CUDA runtime,
tests/hostsim/launch_sim.cpp) against the golden sequences.  Test scaffolding only: the product library is never replaced;
the simulation build is loaded explicitly here and nowhere else."""
import ctypes
import os
import subprocess

import numpy as np
import pytest

from uvio_b200 import configs, track
from uvio_b200.synth import frames_checksum

from golden_util import GOLDEN, check_tracker_against_golden
from util import frames_for

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
SIM = os.path.join(HERE, "hostsim")
FLAGS = ["-O2", "-std=c++17", "-fPIC", "-shared", "-ffp-contract=off", "-Wno-unknown-pragmas"]


from conftest import sim_build as _build


@pytest.fixture(scope="module")
def bodies():
    return ctypes.CDLL(_build(os.path.join(SIM, "_sim", "libbodies.so"), [os.path.join(SIM, "hostsim.cpp")], ["-I", SIM]))


def _p(a):
    return ctypes.c_void_p(a.ctypes.data)


def test_cell_top_k_matches_real_std_sort(bodies):
    """The host's per-cell top-K (radix order by position + prefix-only replica of libstdc++'s introsort) against the
    oracle's real std::sort with the reference comparator, on cells with many response ties (Grider_GRID.h:125-141)."""
    from oracle.trackklt_cv2 import std_sort_perm
    rng = np.random.default_rng(5)
    L = bodies
    for trial in range(400):
        cw, ch = int(rng.integers(8, 120)), int(rng.integers(8, 100))
        n = int(rng.integers(1, min(cw * ch, 700)))
        if trial % 7 == 0:
            n = min(n, 40)                      # the comparison-sort path for small cells
        K = int(rng.integers(1, 24))
        pos = rng.choice(cw * ch, size=n, replace=False)
        nresp = int(rng.integers(1, 40))        # few distinct scores -> many ties
        score = rng.integers(20, 20 + nresp, size=n).astype(np.uint32)
        words = (score << 24) | ((pos // cw).astype(np.uint32) << 12) | (pos % cw).astype(np.uint32)
        words = np.ascontiguousarray(words[rng.permutation(n)], dtype=np.uint32)   # the GPU's append order is arbitrary
        out = np.zeros((K, 3), np.float32)
        got = L.hs_cell_top_k(_p(words), n, cw, ch, K, _p(out))
        order = np.argsort(words & 0xFFFFFF, kind="stable")                        # cv::FAST emits row-major
        rm = words[order]
        perm = std_sort_perm((rm >> 24).astype(np.float32))
        want = rm[perm][:K]
        assert got == min(K, n)
        exp = np.stack([(want & 4095), (want >> 12) & 4095, want >> 24], axis=1).astype(np.float32)
        np.testing.assert_array_equal(out[:got], exp, err_msg=f"trial {trial}: n={n} K={K} cell {cw}x{ch}")


def test_device_top_k_matches_real_std_sort(bodies):
    """The device-side pipeline's per-cell top-K (track_body.h: bitonic order by position + dv_sort_prefix, libstdc++'s
    introsort restated without the standard library, heap-sort fallback included) against the real std::sort."""
    from oracle.trackklt_cv2 import std_sort_perm
    import subprocess, os, ctypes
    rng = np.random.default_rng(9)
    L = bodies
    # the real std::make_heap + std::sort_heap, for the fallback introsort takes when its depth limit is hit
    src = os.path.join(SIM, "_sim", "heapsort_ref.cpp")
    lib = os.path.join(SIM, "_sim", "libheapsort_ref.so")
    os.makedirs(os.path.dirname(src), exist_ok=True)
    open(src, "w").write('#include <algorithm>\n#include <cstdint>\nextern "C" void ref_heap_sort(uint32_t *a, int n) {\n'
                         '  auto c = [](uint32_t x, uint32_t y) { return (x >> 24) > (y >> 24); };\n'
                         '  std::make_heap(a, a + n, c); std::sort_heap(a, a + n, c); }\n')
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-o", lib, src])
    R = ctypes.CDLL(lib)
    for trial in range(400):
        cw, ch = int(rng.integers(8, 160)), int(rng.integers(8, 110))
        n = int(rng.integers(1, min(cw * ch, 2500 if trial % 9 == 0 else 700)))
        K = int(rng.integers(1, 45))
        pos = rng.choice(cw * ch, size=n, replace=False)
        nresp = int(rng.integers(1, 40))
        score = rng.integers(20, 20 + nresp, size=n).astype(np.uint32)
        words = (score << 24) | ((pos // cw).astype(np.uint32) << 12) | (pos % cw).astype(np.uint32)
        words = np.ascontiguousarray(words[rng.permutation(n)], dtype=np.uint32)
        rm = words[np.argsort(words & 0xFFFFFF, kind="stable")]
        want = rm[std_sort_perm((rm >> 24).astype(np.float32))][:K]
        out = np.zeros(K, np.uint32)
        got = L.hs_dv_top_k(_p(words), n, K, 0, _p(out))
        assert got == min(K, n)
        np.testing.assert_array_equal(out[:got], want, err_msg=f"trial {trial}: n={n} K={K}")
        if trial % 4 == 0:
            full = np.zeros(n, np.uint32)
            L.hs_dv_top_k(_p(words), n, n, 1, _p(full))
            ref = rm.copy()
            R.ref_heap_sort(_p(ref), n)
            np.testing.assert_array_equal(full, ref, err_msg=f"trial {trial}: heap-sort fallback n={n}")


def test_ransac_root_order_on_tie_cases(bodies):
    """Two models of one 7-point sample with the same best inlier count: the winner is the one cv::run7Point lists first
    (tests/golden/ransac_ties.npz, masks from cv2; see tests/golden/make_ransac_ties.py)."""
    L = bodies
    g = np.load(os.path.join(GOLDEN, "ransac_ties.npz"))
    K = np.ascontiguousarray(g["intrinsics"], np.float64)
    D = np.ascontiguousarray(g["dist"], np.float64)
    for k in ("a", "b"):
        p0 = np.ascontiguousarray(g[f"p0_{k}"], np.float32)
        p1 = np.ascontiguousarray(g[f"p1_{k}"], np.float32)
        a_n, b_n = np.zeros_like(p0), np.zeros_like(p1)
        L.hs_undistort(_p(p0), len(p0), int(g["model"]), _p(K), _p(D), _p(a_n))
        L.hs_undistort(_p(p1), len(p1), int(g["model"]), _p(K), _p(D), _p(b_n))
        mask = np.zeros(len(p0), np.uint8)
        L.hs_ransac(_p(a_n), _p(b_n), len(p0), ctypes.c_double(2.0 / max(K[0], K[1])), ctypes.c_double(0.999), 1000, _p(mask))
        assert np.array_equal(mask, g[f"mask_{k}"]), f"case {k}: {np.nonzero(mask != g[f'mask_{k}'])[0]}"


def test_kernel_bodies_against_stage_goldens(bodies):
    L = bodies
    g = np.load(os.path.join(GOLDEN, "stages_euroc.npz"))
    cfg = configs.EUROC_MAV
    frames = frames_for(cfg, 2, events=False, start=7)
    w, h = cfg.width, cfg.height
    eq = []
    for fr in frames:
        o = np.empty_like(fr[0])
        L.hs_equalize(_p(fr[0]), w, h, w, _p(o), w)
        eq.append(o)
    cl = np.empty_like(frames[0][0])
    L.hs_clahe(_p(frames[0][0]), w, h, w, _p(cl), w, ctypes.c_double(10.0), 8, 8)
    assert frames_checksum([cl]) == int(g["clahe_sum"])
    pyr = [[e] for e in eq]
    for k in range(2):
        for _ in range(4):
            a = pyr[k][-1]
            d = np.empty(((a.shape[0] + 1) // 2, (a.shape[1] + 1) // 2), np.uint8)
            L.hs_pyrdown(_p(a), a.shape[1], a.shape[0], a.shape[1], _p(d), d.shape[1])
            pyr[k].append(d)
    for lv in range(5):
        assert frames_checksum([pyr[0][lv]]) == int(g[f"pyr_sum_{lv}"])
    sx, sy = w // cfg.grid_x, h // cfg.grid_y
    for (gx, gy) in ((0, 0), (2, 3), (4, 4)):
        xy = np.empty((4096, 2), np.int32)
        r = np.empty(4096, np.float32)
        n = L.hs_fast(_p(eq[0]), w, h, w, gx * sx, gy * sy, sx, sy, cfg.fast_threshold, _p(xy), _p(r), 4096)
        assert np.array_equal(xy[:n], g[f"fast_xy_{gx}_{gy}"].astype(np.int32)) and np.array_equal(r[:n], g[f"fast_resp_{gx}_{gy}"].astype(np.float32))
    sp = g["subpix_in"].copy()
    L.hs_subpix(_p(eq[0]), w, h, w, _p(sp), len(sp), 20, ctypes.c_double(0.001))
    assert np.array_equal(sp, g["subpix_out"])
    vp = ctypes.c_void_p
    PP = (vp * 5)(*[a.ctypes.data for a in pyr[0]])
    NN = (vp * 5)(*[a.ctypes.data for a in pyr[1]])
    lw = np.array([a.shape[1] for a in pyr[0]], np.int32)
    lh = np.array([a.shape[0] for a in pyr[0]], np.int32)
    p0 = g["lk_in"].copy()
    p1 = p0.copy()
    st = np.zeros(len(p0), np.uint8)
    L.hs_lk(PP, NN, _p(lw), _p(lh), _p(lw), 4, _p(p0), _p(p1), _p(st), len(p0), 30, ctypes.c_double(0.01))
    assert np.array_equal(st, g["lk_status"]) and np.array_equal(p1[st == 1], g["lk_out"][st == 1])
    c = cfg.cams[0]
    K, D = np.array(c.intrinsics, np.float64), np.array(c.dist, np.float64)
    uin = np.ascontiguousarray(g["undist_in_radtan"])           # keep the arrays alive while C reads them
    un = np.empty_like(uin)
    L.hs_undistort(_p(uin), len(un), 0, _p(K), _p(D), _p(un))
    assert np.array_equal(un, g["undist_out_radtan"])
    rp0, rp1 = np.ascontiguousarray(g["ransac_p0"]), np.ascontiguousarray(g["ransac_p1"])
    a_n, b_n = np.empty_like(rp0), np.empty_like(rp1)
    L.hs_undistort(_p(rp0), len(a_n), 0, _p(K), _p(D), _p(a_n))
    L.hs_undistort(_p(rp1), len(b_n), 0, _p(K), _p(D), _p(b_n))
    mask = np.zeros(len(a_n), np.uint8)
    L.hs_ransac(_p(a_n), _p(b_n), len(a_n), ctypes.c_double(2.0 / max(c.intrinsics[0], c.intrinsics[1])), ctypes.c_double(0.999), 1000, _p(mask))
    assert np.array_equal(mask, g["ransac_mask"])


@pytest.mark.parametrize("name,n", [("euroc_mav", 8), ("tum_vi", 6), ("rs_d455", 4), ("euroc_mono", 6), ("hd_1000", 3)])
def test_host_logic_against_golden_sequences(sim_tracker, name, n):
    assert check_tracker_against_golden(name, sim_tracker, n=n) > 300
