"""ctypes binding of include/uvio_tri.h: batched triangulation of the features a DeviceFeatureDatabase query returned
(ov_core::FeatureInitializer::single_triangulation / single_triangulation_1d / single_gaussnewton, one GPU thread per feature)."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import track
from .featdb import _lib as _fdb_lib

CLONE_DTYPE = np.dtype([("cam_id", np.uint64), ("timestamp", np.float64), ("R_GtoC", np.float64, (9,)), ("p_CinG", np.float64, (3,))])
RESULT_DTYPE = np.dtype([("ok_triangulation", np.int32), ("ok_refine", np.int32), ("anchor_cam_id", np.int32), ("n_meas", np.int32),
                         ("anchor_clone_timestamp", np.float64), ("p_FinA", np.float64, (3,)), ("p_FinG", np.float64, (3,))])


class UvioTriOptions(C.Structure):
    _fields_ = [("triangulate_1d", C.c_int), ("refine_features", C.c_int), ("max_runs", C.c_int), ("init_lamda", C.c_double),
                ("max_lamda", C.c_double), ("min_dx", C.c_double), ("min_dcost", C.c_double), ("lam_mult", C.c_double),
                ("min_dist", C.c_double), ("max_dist", C.c_double), ("max_baseline", C.c_double), ("max_cond_number", C.c_double)]


def _lib():
    L = _fdb_lib()
    L.uvio_tri_default_options.argtypes = [C.POINTER(UvioTriOptions)]
    L.uvio_tri_default_options.restype = None
    L.uvio_tri_triangulate_featdb.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.POINTER(UvioTriOptions), C.c_void_p]
    L.uvio_tri_triangulate.argtypes = [C.c_int, C.c_int] + [C.c_void_p] * 6 + [C.c_int, C.POINTER(UvioTriOptions), C.c_void_p]
    return L


def options(**over) -> UvioTriOptions:
    o = UvioTriOptions()
    _lib().uvio_tri_default_options(C.byref(o))
    for k, v in over.items():
        setattr(o, k, v)
    return o


def pack_clones(clones) -> np.ndarray:
    """clones[cam][timestamp] = (R_GtoC 3x3, p_CinG 3) -> array of CLONE_DTYPE"""
    rows = [(cam, t, np.asarray(R, np.float64).reshape(9), np.asarray(p, np.float64)) for cam, d in clones.items() for t, (R, p) in d.items()]
    a = np.zeros(len(rows), CLONE_DTYPE)
    for i, r in enumerate(rows):
        a[i] = r
    return a


def triangulate_last_query(db, clones, opt=None, n_features=None) -> np.ndarray:
    """Triangulates every feature of db's last query result, in result order."""
    L = _lib()
    opt = opt or options()
    c = pack_clones(clones)
    nf = C.c_int()
    if n_features is None:
        raise ValueError("pass the number of features the query returned")
    out = np.zeros(max(n_features, 1), RESULT_DTYPE)
    rc = L.uvio_tri_triangulate_featdb(db._h, c.ctypes.data, len(c), C.byref(opt), out.ctypes.data)
    if rc != track.UVIO_OK:
        raise track.UvioError(rc, "uvio_tri_triangulate_featdb")
    return out[:n_features]


def triangulate(row0, row_cam, off, t, uvn, clones, opt=None, device=0) -> np.ndarray:
    L = _lib()
    opt = opt or options()
    c = pack_clones(clones)
    row0, row_cam, off = (np.ascontiguousarray(a, np.int32) for a in (row0, row_cam, off))
    t, uvn = np.ascontiguousarray(t, np.float64), np.ascontiguousarray(uvn, np.float32)
    n = len(row0) - 1
    out = np.zeros(max(n, 1), RESULT_DTYPE)
    rc = L.uvio_tri_triangulate(int(device), n, row0.ctypes.data, row_cam.ctypes.data, off.ctypes.data, t.ctypes.data, uvn.ctypes.data,
                                c.ctypes.data, len(c), C.byref(opt), out.ctypes.data)
    if rc != track.UVIO_OK:
        raise track.UvioError(rc, "uvio_tri_triangulate")
    return out[:n]
