"""GPU: uvio_klt_clahe_noalloc (the allocation-free CLAHE entry point, static __device__ scratch) against the tracker's own
CLAHE path (level 0 of uvio_klt_stage_get_level after a feed), through the C ABI, bit-exact, on two calls in a row.

Collected last on purpose; passes alone in a fresh process and after the other tests.
torch is used here only to hold the caller's device buffers -- the entry point itself allocates nothing."""
import ctypes as C

import numpy as np
import pytest

from uvio_b200 import configs
from uvio_b200.track import UVIO_E_ARG, UVIO_OK, CameraData, lib, make_tracker

from util import frames_for, masks_for

pytestmark = pytest.mark.gpu


def _bind():
    L = lib()
    f = L.uvio_klt_clahe_noalloc
    f.argtypes = [C.c_void_p, C.c_size_t, C.c_size_t, C.c_void_p, C.c_size_t, C.c_size_t, C.c_int, C.c_int, C.c_int, C.c_void_p,
                  C.c_char_p, C.c_size_t]
    f.restype = C.c_int
    return f


def _tracker_clahe(cfg, frame):
    """Level 0 of both cameras through the existing entry points (feed -> stage_get_level)."""
    trk = make_tracker(cfg)
    trk.feed_new_camera(CameraData(0.0, [0, 1], [frame[0], frame[1]], masks_for(cfg)))
    out = [trk.stage_level(0, 0, 0), trk.stage_level(1, 0, 0)]
    trk.close()
    return out


def test_clahe_noalloc_matches_the_tracker_twice_in_a_row(built_lib):
    import torch
    f = _bind()
    cfg = configs.RS_D455                      # the CLAHE configuration (640x480)
    frames = frames_for(cfg, 2, stream_id=3)
    err = C.create_string_buffer(256)
    stream = torch.cuda.current_stream().cuda_stream
    # call 1: a batch of two tightly packed images; call 2: other content, padded pitch, into the same (dirty) scratch
    for call, (frame, pitch) in enumerate(((frames[0], cfg.width), (frames[1], cfg.width + 64))):
        ref = _tracker_clahe(cfg, frame)
        host = np.zeros((2, cfg.height, pitch), np.uint8)
        host[0, :, :cfg.width], host[1, :, :cfg.width] = frame[0], frame[1]
        src = torch.from_numpy(host).cuda()
        dst = torch.full_like(src, 0xAB)
        rc = f(src.data_ptr(), pitch, pitch * cfg.height, dst.data_ptr(), pitch, pitch * cfg.height, cfg.width, cfg.height, 2,
               stream, err, len(err))
        assert rc == UVIO_OK, err.value
        torch.cuda.synchronize()
        got = dst.cpu().numpy()
        for cam in (0, 1):
            assert np.array_equal(got[cam, :, :cfg.width], ref[cam]), f"call {call} cam {cam}: differs from the tracker's CLAHE"
        assert (got[:, :, cfg.width:] == 0xAB).all()      # the padding of the output rows is not touched


def test_clahe_noalloc_refuses_sizes_beyond_the_scratch(built_lib):
    import torch
    f = _bind()
    err = C.create_string_buffer(256)
    buf = torch.zeros(64 * 64, dtype=torch.uint8, device="cuda")
    rc = f(buf.data_ptr(), 64, 64 * 64, buf.data_ptr(), 64, 64 * 64, 64, 64, 33, None, err, len(err))   # > 32 images
    assert rc == UVIO_E_ARG and b"at most 32" in err.value
    rc = f(buf.data_ptr(), 9000, 0, buf.data_ptr(), 9000, 0, 9000, 8, 1, None, err, len(err))            # > 8192 wide
    assert rc == UVIO_E_ARG and b"8192" in err.value
    assert f(buf.data_ptr(), 64, 0, buf.data_ptr(), 64, 0, 64, 64, 0, None, err, len(err)) == UVIO_OK     # empty batch
