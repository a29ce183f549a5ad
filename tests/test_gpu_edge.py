"""GPU edge cases through the C ABI: input layouts (strided ROI views, frames already in device memory), small and odd image
sizes, blank and fully masked frames, few features (the LMedS / no-RANSAC branches of TrackKLT.cpp:848-852)."""
import numpy as np
import pytest

from uvio_b200 import configs
from uvio_b200.track import CameraData, DeviceImage, feed_batch, make_tracker

from test_gpu_tracker import _compare_frame
from util import frames_for, log_arrays, masks_for, run_oracle

pytestmark = pytest.mark.gpu


def _roi_view(img, pad_x, pad_y, fill):
    """img embedded in a larger array; the returned view has stride > width (a cv::Mat ROI)."""
    big = np.full((img.shape[0] + 2 * pad_y, img.shape[1] + 2 * pad_x), fill, np.uint8)
    big[pad_y:pad_y + img.shape[0], pad_x:pad_x + img.shape[1]] = img
    v = big[pad_y:pad_y + img.shape[0], pad_x:pad_x + img.shape[1]]
    assert v.strides[0] > img.shape[1]
    return v


def test_input_layouts_give_identical_tracks(built_lib):
    """Contiguous host frames, ROI views with a stride (image and mask), and frames resident in device memory."""
    import torch
    cfg = configs.TUM_VI                      # non-zero masks
    n = 8
    frames = frames_for(cfg, n, stream_id=21)
    masks = masks_for(cfg)
    a, b, c = make_tracker(cfg), make_tracker(cfg), make_tracker(cfg)
    for k, (l, r) in enumerate(frames):
        a.feed_new_camera(CameraData(0.05 * k, [0, 1], [l, r], masks))
        b.feed_new_camera(CameraData(0.05 * k, [0, 1], [_roi_view(l, 9, 3, 200), _roi_view(r, 5, 2, 7)],
                                     [_roi_view(masks[0], 4, 1, 255), _roi_view(masks[1], 11, 6, 255)]))
        dl, dr = torch.from_numpy(l).cuda(), torch.from_numpy(r).cuda()
        c.feed_new_camera(CameraData(0.05 * k, [0, 1], [DeviceImage.of(dl), DeviceImage.of(dr)], masks))
        oa, ob, oc = a.new_observations(), b.new_observations(), c.new_observations()
        assert oa.tobytes() == ob.tobytes(), f"frame {k}: strided input differs"
        assert oa.tobytes() == oc.tobytes(), f"frame {k}: device-resident input differs"
    assert a.get_feature_database().size() > 50


@pytest.mark.parametrize("w,h", [(161, 127), (96, 80), (333, 241)])
def test_small_and_odd_sizes_match_oracle(built_lib, w, h):
    """Odd sizes: pyramid levels with odd widths, fewer levels than requested, cells narrower than a FAST tile."""
    cams = tuple(configs.CamCalib(c.model, (c.intrinsics[0] * w / 752, c.intrinsics[1] * h / 480, w / 2.0 + 1.5, h / 2.0 - 2.0), c.dist)
                 for c in configs.EUROC_MAV.cams)
    cfg = configs.EUROC_MAV.with_(name=f"odd_{w}x{h}", width=w, height=h, num_features=60, grid_x=3, grid_y=3, min_px_dist=6, cams=cams)
    frames = frames_for(cfg, 10, stream_id=31)
    masks = masks_for(cfg)
    _o, ref = run_oracle(cfg, frames, masks)
    trk = make_tracker(cfg)
    total = 0
    for k, (l, r) in enumerate(frames):
        trk.feed_new_camera(CameraData(0.05 * k, [0, 1], [l, r], masks))
        obs = trk.new_observations()
        _compare_frame(k, trk, obs, ref[k], cfg.name)
        total += len(obs)
    assert total > 0


def test_blank_then_textured_and_fully_masked(built_lib):
    """Blank frames (no corners at all), then texture; then a mask of 255 everywhere on one camera."""
    cfg = configs.EUROC_MAV.with_(name="blank", num_features=80)
    tex = frames_for(cfg, 6, stream_id=41, events=False)
    blank = np.full((cfg.height, cfg.width), 90, np.uint8)
    frames = [(blank, blank.copy())] * 2 + tex[:3] + [(blank, blank.copy())] + tex[3:]
    zero = np.zeros((cfg.height, cfg.width), np.uint8)
    full = np.full((cfg.height, cfg.width), 255, np.uint8)
    mask_seq = [[zero, zero]] * 6 + [[full, zero]] * 2 + [[zero, zero]]
    from oracle.trackklt_cv2 import make_oracle
    o = make_oracle(cfg)
    trk = make_tracker(cfg)
    for k, ((l, r), m) in enumerate(zip(frames, mask_seq)):
        n0 = len(o.database.log)
        o.feed_new_camera(0.05 * k, [0, 1], [l, r], m)
        ref = (list(o.database.log[n0:]), {c: list(p) for c, p in o.pts_last.items()}, {c: list(i) for c, i in o.ids_last.items()})
        trk.feed_new_camera(CameraData(0.05 * k, [0, 1], [l, r], m))
        _compare_frame(k, trk, trk.new_observations(), ref, "blank")


@pytest.mark.parametrize("nf", [4, 12, 16, 24])
def test_few_features_match_oracle(built_lib, nf):
    """num_features 4: a camera never holds 10 points, so perform_matching returns without tracking every frame
    (TrackKLT.cpp:848-852) and detection starts over.  12 / 16 / 24: RANSAC around its 15-point minimum; a camera holding
    10..13 points sends cv::findFundamentalMat into LMedS, which is decided by rounding noise (15 % / 7 % /
    1.3 % of the matching calls at these budgets, profiles/r2_lmeds_census.json).  Up to the first such feed everything equals
    the oracle.  At that feed the state going in is still identical, so: every observed id is one the oracle tracked into
    this feed or assigned in it, and the positions of ids both report are bit-equal (LK does not depend on the RANSAC mask);
    from then on the streams are different random draws of the same procedure and only the invariants are checked."""
    cfg = configs.EUROC_MAV.with_(name=f"few{nf}", num_features=nf, grid_x=2, grid_y=2)
    frames = frames_for(cfg, 14, stream_id=51, events=False)
    masks = masks_for(cfg)
    _o, ref = run_oracle(cfg, frames, masks)
    trk = make_tracker(cfg)
    compared, diverged_at = 0, None
    for k, (l, r) in enumerate(frames):
        trk.feed_new_camera(CameraData(0.05 * k, [0, 1], [l, r], masks))
        obs = trk.new_observations()
        if diverged_at is None and k and any(10 <= len(v) <= 13 for v in ref[k - 1][2].values()):
            diverged_at = k
            log, _pts, ids_last = ref[k]
            o_ids, o_cams, o_uv, _ = log_arrays(log)
            candidates = set(i for v in ref[k - 1][2].values() for i in v) | set(o_ids.tolist()) | set(i for v in ids_last.values() for i in v)
            assert set(obs["id"].astype(np.int64).tolist()) <= candidates, f"frame {k}: an id the oracle never held"
            want = {(int(i), int(c)): tuple(uv) for i, c, uv in zip(o_ids, o_cams, o_uv)}
            for ob in obs:
                key = (int(ob["id"]), int(ob["cam_id"]))
                if key in want:
                    assert (ob["u"], ob["v"]) == want[key], f"frame {k} id {key}: LK position differs"
        if diverged_at is None:
            _compare_frame(k, trk, obs, ref[k], cfg.name)
            compared += 1
        ids = trk.get_last_ids()
        for c in (0, 1):                                                   # invariants, every frame
            assert len(set(ids[c].tolist())) == len(ids[c])
            assert len(ids[c]) <= 2 * (nf + 4 * (nf // 4 + 1))      # left top-up + stereo matches + right top-up
    assert compared >= (4 if nf != 12 else 2), f"only {compared} frames could be compared"


def test_identical_frames_where_the_reference_would_terminate(built_lib):
    """The same frame delivered twice: every correspondence is identical, cv::findFundamentalMat asserts (matrix.cpp:764, any
    N) and the reference, which does not catch it (TrackKLT.cpp:873), terminates.  Documented difference (include/uvio_klt.h):
    the GPU tracker keeps every tracked point (a zero-motion model fits all of them) and carries on."""
    import cv2
    cfg = configs.EUROC_MAV
    frames = frames_for(cfg, 2, stream_id=9, events=False)
    masks = masks_for(cfg)
    from oracle.trackklt_cv2 import make_oracle
    o = make_oracle(cfg)
    trk = make_tracker(cfg)
    for k, fr in enumerate([frames[0], frames[1], frames[1]]):
        trk.feed_new_camera(CameraData(0.05 * k, [0, 1], [fr[0], fr[1]], masks))
        if k < 2:
            o.feed_new_camera(0.05 * k, [0, 1], [fr[0], fr[1]], masks)
        else:
            with pytest.raises(cv2.error):
                o.feed_new_camera(0.05 * k, [0, 1], [fr[0], fr[1]], masks)
    obs = trk.new_observations()
    assert len(obs) > 300 and trk.get_feature_database().size() > 0


def test_batch_mixes_sizes_and_modes(built_lib):
    """One feed_batch call with trackers of different image sizes, camera models and stereo/mono modes."""
    cfgs = [configs.EUROC_MAV, configs.TUM_VI, configs.UZHFPV_INDOOR]
    n = 6
    seqs = [frames_for(c, n, stream_id=61 + i) for i, c in enumerate(cfgs)]
    masks = [masks_for(c) for c in cfgs]
    single = [make_tracker(c) for c in cfgs] + [make_tracker(configs.EUROC_MAV)]
    batched = [make_tracker(c) for c in cfgs] + [make_tracker(configs.EUROC_MAV)]
    for k in range(n):
        msgs = [CameraData(0.05 * k, [0, 1], [seqs[i][k][0], seqs[i][k][1]], masks[i]) for i in range(len(cfgs))]
        msgs.append(CameraData(0.05 * k, [0], [seqs[0][k][1]], [masks[0][0]]))       # a mono feed of a stereo-capable rig's camera 0
        for t, m in zip(single, msgs):
            t.feed_new_camera(m)
        feed_batch(batched, msgs)
        for i, (a, b) in enumerate(zip(single, batched)):
            assert a.new_observations().tobytes() == b.new_observations().tobytes(), f"tracker {i} frame {k}"


def test_ransac_tie_cases_match_cv2(built_lib):
    """tests/golden/ransac_ties.npz through uvio_klt_stage_ransac (root order of the 7-point solver decides these)."""
    import os
    from golden_util import GOLDEN
    g = np.load(os.path.join(GOLDEN, "ransac_ties.npz"))
    w, h = 161, 127
    cam = configs.CamCalib(int(g["model"]), tuple(float(v) for v in g["intrinsics"]), tuple(float(v) for v in g["dist"]))
    cfg = configs.EUROC_MAV.with_(name="ties", width=w, height=h, cams=(cam, cam))
    trk = make_tracker(cfg)
    img = np.zeros((h, w), np.uint8)
    trk.feed_new_camera(CameraData(0.0, [0, 1], [img, img], [img, img]))
    for k in ("a", "b"):
        mask, _p0n, _p1n = trk.stage_ransac(0, 0, g[f"p0_{k}"], g[f"p1_{k}"])
        assert np.array_equal(mask, g[f"mask_{k}"]), f"case {k}: {np.nonzero(mask != g[f'mask_{k}'])[0]}"


def test_destroyed_trackers_give_their_device_memory_back(built_lib):
    """uvio_klt_destroy (TrackBase::~TrackBase): pyramids are freed, upload buffers recycled -- creating and destroying
    trackers in a loop must not grow the device footprint, and a recycled buffer must not change results."""
    import torch
    cfg = configs.HD_1000.with_(num_features=150)
    frames = frames_for(cfg, 3, stream_id=3)
    masks = masks_for(cfg)

    def one_life():
        t = make_tracker(cfg)
        for k, (l, r) in enumerate(frames):
            t.feed_new_camera(CameraData(0.05 * k, [0, 1], [l, r], masks))
        out = t.new_observations().tobytes()
        t.close()
        return out

    first = one_life()
    one_life()
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    for _ in range(12):
        assert one_life() == first
    torch.cuda.synchronize()
    free1 = torch.cuda.mem_get_info()[0]
    assert free0 - free1 < 8 << 20, f"device memory grew by {(free0 - free1) >> 20} MiB over 12 tracker lifetimes"
