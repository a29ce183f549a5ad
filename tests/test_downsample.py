"""downsample_cameras: the step VioManager::track_image_and_update runs in front of feed_new_camera
(ov_msckf/src/core/VioManager.cpp:268-278) -- cv::pyrDown of every image and mask to Size(cols / 2.0, rows / 2.0) -- done
inside the tracker.  Oracle: cv2.pyrDown in front of the cv2-driven TrackKLT restatement, intrinsics halved as
VioManagerOptions.h:251-260 does."""
import numpy as np
import pytest

from uvio_b200 import configs
from uvio_b200.synth import SynthStereoStream, circular_mask
from uvio_b200.track import CameraData, DeviceImage, make_tracker

from util import log_arrays


def halved(cfg, w, h):
    cams = tuple(configs.CamCalib(c.model, tuple(v / 2.0 for v in c.intrinsics), c.dist) for c in cfg.cams)
    return cfg.with_(cams=cams, width=w // 2, height=h // 2)


def run_case(make, cfg, w, h, n, masked, mono=False):
    from oracle.trackklt_cv2 import downsample_message, make_oracle
    s = SynthStereoStream(w, h, stream_id=77, events=True)
    frames = [s.frame(k) for k in range(n)]
    m = circular_mask(w, h) if masked else np.zeros((h, w), np.uint8)
    masks = [m, m.copy()]
    small = halved(cfg, w, h)
    ora = make_oracle(small)
    trk = make(small, downsample_cameras=True)
    total = 0
    for k, (l, r) in enumerate(frames):
        imgs, ids, mk = ([l], [0], masks[:1]) if mono else ([l, r], [0, 1], masks)
        n0 = len(ora.database.log)
        di, dm = downsample_message(imgs, mk)
        ora.feed_new_camera(0.05 * k, ids, di, dm)
        trk.feed_new_camera(CameraData(0.05 * k, ids, imgs, mk))
        oi, oc, ouv, ouvn = log_arrays(ora.database.log[n0:])
        obs = trk.new_observations()
        assert np.array_equal(obs["id"].astype(np.int64), oi), f"frame {k}: id stream differs"
        assert np.array_equal(obs["cam_id"].astype(np.int64), oc), f"frame {k}: camera stream differs"
        if len(oi):
            assert np.array_equal(np.c_[obs["u"], obs["v"]], ouv), f"frame {k}: positions not bit-exact"
            assert np.abs(np.c_[obs["un"], obs["vn"]] - ouvn).max() <= 1e-5
        last = trk.get_last_ids()
        for c in ids:
            assert np.array_equal(last[c].astype(np.int64), np.array(ora.ids_last[c], np.int64)), f"frame {k} cam {c}: ids_last"
        total += len(oi)
    return total


CASES = [("even", configs.EUROC_MAV, 752, 480, False, False),
         ("odd_masked", configs.TUM_VI.with_(num_features=120), 753, 481, True, False),   # odd sizes lose the last row / column
         ("mono_masked", configs.EUROC_MAV.with_(num_features=100), 640, 482, True, True)]


@pytest.mark.parametrize("name,cfg,w,h,masked,mono", CASES, ids=[c[0] for c in CASES])
def test_downsample_host_logic_sim(sim_tracker, name, cfg, w, h, masked, mono):
    """engine + ABI over the host-simulated kernels (CPU): the halving, the mask path and the size bookkeeping."""
    assert run_case(sim_tracker, cfg, w, h, 5, masked, mono) > 100


@pytest.mark.gpu
@pytest.mark.parametrize("name,cfg,w,h,masked,mono", CASES, ids=[c[0] for c in CASES])
def test_downsample_gpu(name, cfg, w, h, masked, mono):
    assert run_case(lambda c, **kw: make_tracker(c, device=0, **kw), cfg, w, h, 20, masked, mono) > 500


@pytest.mark.gpu
def test_downsample_device_frames_and_size_errors():
    """Frames already in device memory are halved where they lie; a message of the wrong (full) size is refused."""
    import torch
    from uvio_b200.track import UvioError
    from oracle.trackklt_cv2 import downsample_message, make_oracle
    cfg, w, h = configs.EUROC_MAV, 752, 480
    small = halved(cfg, w, h)
    s = SynthStereoStream(w, h, stream_id=5, events=False)
    z = np.zeros((h, w), np.uint8)
    ora, trk = make_oracle(small), make_tracker(small, device=0, downsample_cameras=True)
    for k in range(6):
        l, r = s.frame(k)
        n0 = len(ora.database.log)
        di, dm = downsample_message([l, r], [z, z])
        ora.feed_new_camera(0.05 * k, [0, 1], di, dm)
        dl, dr = torch.from_numpy(l).cuda(), torch.from_numpy(r).cuda()
        trk.feed_new_camera(CameraData(0.05 * k, [0, 1], [DeviceImage.of(dl), DeviceImage.of(dr)], [z, z]))
        oi, oc, ouv, _ = log_arrays(ora.database.log[n0:])
        obs = trk.new_observations()
        assert np.array_equal(obs["id"].astype(np.int64), oi) and np.array_equal(np.c_[obs["u"], obs["v"]], ouv)
    with pytest.raises(UvioError):
        l, r = s.frame(0)
        trk.feed_new_camera(CameraData(1.0, [0, 1], [l[:240, :376].copy(), r[:240, :376].copy()], [z[:240, :376].copy()] * 2))
