"""ORACLE — TEST INFRASTRUCTURE ONLY. ctypes access to oracle/stages.c (plain-C restatement of the OpenCV routines)."""
from __future__ import annotations

import ctypes
import os

import numpy as np

from .trackklt_cv2 import _BUILD, build_helpers

_lib = None
_vp, _i, _d = ctypes.c_void_p, ctypes.c_int, ctypes.c_double


def lib():
    global _lib
    if _lib is None:
        build_helpers()
        _lib = ctypes.CDLL(os.path.join(_BUILD, "libcstages.so"))
    return _lib


def _u8(a):
    a = np.ascontiguousarray(a, np.uint8)
    assert a.ndim == 2
    return a


def equalize_hist(img):
    img = _u8(img)
    out = np.empty_like(img)
    lib().oracle_equalize_hist(_vp(img.ctypes.data), _i(img.shape[1]), _i(img.shape[0]), _i(img.strides[0]),
                               _vp(out.ctypes.data), _i(out.strides[0]))
    return out


def clahe(img, clip=10.0, tiles=(8, 8)):
    img = _u8(img)
    out = np.empty_like(img)
    lib().oracle_clahe(_vp(img.ctypes.data), _i(img.shape[1]), _i(img.shape[0]), _i(img.strides[0]),
                       _vp(out.ctypes.data), _i(out.strides[0]), _d(clip), _i(tiles[0]), _i(tiles[1]))
    return out


def pyr_down(img):
    img = _u8(img)
    h, w = img.shape
    out = np.empty(((h + 1) // 2, (w + 1) // 2), np.uint8)
    lib().oracle_pyr_down(_vp(img.ctypes.data), _i(w), _i(h), _i(img.strides[0]), _vp(out.ctypes.data), _i(out.strides[0]))
    return out


def pyr_max_level(w, h, requested, win=15):
    return int(lib().oracle_pyr_max_level(_i(w), _i(h), _i(requested), _i(win)))


def build_pyramid(img, pyr_levels, win=15):
    levels = [_u8(img)]
    for _ in range(pyr_max_level(img.shape[1], img.shape[0], pyr_levels, win)):
        levels.append(pyr_down(levels[-1]))
    return levels


def scharr(img):
    img = _u8(img)
    h, w = img.shape
    out = np.empty((h, w, 2), np.int16)
    lib().oracle_scharr(_vp(img.ctypes.data), _i(w), _i(h), _i(img.strides[0]), _vp(out.ctypes.data))
    return out


def fast(roi, threshold, nonmax=True, max_out=65536):
    roi = _u8(roi)
    xy = np.empty((max_out, 2), np.int32)
    resp = np.empty(max_out, np.float32)
    n = lib().oracle_fast(_vp(roi.ctypes.data), _i(roi.shape[1]), _i(roi.shape[0]), _i(roi.strides[0]), _i(threshold),
                          _i(1 if nonmax else 0), _vp(xy.ctypes.data), _vp(resp.ctypes.data), _i(max_out))
    assert n <= max_out
    return xy[:n].copy(), resp[:n].copy()


def undistort(uv, model, intrinsics, dist):
    uv = np.ascontiguousarray(uv, np.float32).reshape(-1, 2)
    out = np.empty_like(uv)
    K = np.array(intrinsics, np.float64)
    D = np.array(dist, np.float64)
    fn = lib().oracle_undistort_radtan if model == 0 else lib().oracle_undistort_equi
    fn(_vp(uv.ctypes.data), _i(uv.shape[0]), _vp(K.ctypes.data), _vp(D.ctypes.data), _vp(out.ctypes.data))
    return out


def corner_subpix(img, pts, win=5, max_iters=20, eps=0.001):
    img = _u8(img)
    p = np.ascontiguousarray(pts, np.float32).reshape(-1, 2).copy()
    lib().oracle_corner_subpix(_vp(img.ctypes.data), _i(img.shape[1]), _i(img.shape[0]), _i(img.strides[0]),
                               _vp(p.ctypes.data), _i(p.shape[0]), _i(win), _i(max_iters), _d(eps))
    return p


def lk(prev_levels, next_levels, prev_pts, next_pts_init, win=15, max_count=30, eps=0.01, accum_mode=1):
    """prev_levels / next_levels: lists of uint8 pyramid levels (level 0 first). Returns (next_pts, status)."""
    L = len(prev_levels)
    pl = [_u8(a) for a in prev_levels]
    nl = [_u8(a) for a in next_levels]
    PP = (ctypes.c_void_p * L)(*[a.ctypes.data for a in pl])
    NP = (ctypes.c_void_p * L)(*[a.ctypes.data for a in nl])
    lw = np.array([a.shape[1] for a in pl], np.int32)
    lh = np.array([a.shape[0] for a in pl], np.int32)
    ls = np.array([a.strides[0] for a in pl], np.int32)
    assert all(a.strides[0] == b.strides[0] and a.shape == b.shape for a, b in zip(pl, nl))
    p0 = np.ascontiguousarray(prev_pts, np.float32).reshape(-1, 2)
    p1 = np.ascontiguousarray(next_pts_init, np.float32).reshape(-1, 2).copy()
    st = np.zeros(p0.shape[0], np.uint8)
    lib().oracle_lk(PP, NP, _vp(lw.ctypes.data), _vp(lh.ctypes.data), _vp(ls.ctypes.data), _i(L - 1),
                    _vp(p0.ctypes.data), _vp(p1.ctypes.data), _vp(st.ctypes.data), _i(p0.shape[0]), _i(win),
                    _i(max_count), _d(eps), _i(accum_mode))
    return p1, st
