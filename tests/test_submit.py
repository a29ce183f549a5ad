"""uvio_klt_submit / uvio_klt_submit_batch: frames announced ahead of their feed are uploaded and pre-processed into the
third pyramid slot while the previous feed runs; results must be identical to plain feeds (and to the golden sequences)."""
import numpy as np
import pytest

from uvio_b200 import configs
from uvio_b200.track import CameraData, UvioError, feed_batch, make_tracker, submit_batch

from golden_util import check_tracker_against_golden
from util import frames_for, masks_for


class _Announcing:
    """Tracker wrapper that announces frame k+1 before feeding frame k (the pipelined calling pattern)."""

    def __init__(self, trk, depth=1):
        self.t, self.queue, self.depth = trk, [], depth

    def __getattr__(self, name):
        return getattr(self.t, name)


def _pipelined_run(make, cfg, frames, masks, mono=False, depth=1):
    """feed every frame, announcing `depth` frames ahead; returns the observation arrays per frame."""
    trk = make(cfg)
    msgs = [CameraData(0.05 * k, [0] if mono else [0, 1], [l] if mono else [l, r], masks[:1] if mono else masks)
            for k, (l, r) in enumerate(frames)]
    out, announced = [], 0
    for k, m in enumerate(msgs):
        while announced < min(len(msgs), k + 1 + depth):
            if announced >= k:                       # frame k itself is announced only when depth allows it ahead of time
                trk.submit(msgs[announced])
            announced += 1
        trk.feed_new_camera(m)
        out.append((trk.new_observations(), {c: v.copy() for c, v in trk.get_last_ids().items()}))
    return out


def _plain_run(make, cfg, frames, masks, mono=False):
    trk = make(cfg)
    out = []
    for k, (l, r) in enumerate(frames):
        trk.feed_new_camera(CameraData(0.05 * k, [0] if mono else [0, 1], [l] if mono else [l, r], masks[:1] if mono else masks))
        out.append((trk.new_observations(), {c: v.copy() for c, v in trk.get_last_ids().items()}))
    return out


def _same(a, b):
    assert len(a) == len(b)
    for k, ((oa, la), (ob, lb)) in enumerate(zip(a, b)):
        assert oa.tobytes() == ob.tobytes(), f"frame {k}: observations differ"
        for c in la:
            assert np.array_equal(la[c], lb[c]), f"frame {k} cam {c}: ids_last differ"


CASES = [("euroc", configs.EUROC_MAV, False, {}), ("tum_vi_masked", configs.TUM_VI, False, {}),
         ("d455_clahe", configs.RS_D455, False, {}), ("mono", configs.EUROC_MAV.with_(num_features=120), True, {}),
         ("downsample", configs.EUROC_MAV.with_(num_features=100), False, {"downsample_cameras": True})]


def _case(make, cfg, mono, kw, n, depth):
    frames = frames_for(cfg, n, stream_id=13)
    masks = masks_for(cfg)
    mk = lambda c: make(c, **kw)
    _same(_plain_run(mk, cfg, frames, masks, mono), _pipelined_run(mk, cfg, frames, masks, mono, depth))


@pytest.mark.parametrize("name,cfg,mono,kw", CASES[:2] + CASES[4:], ids=[c[0] for c in CASES[:2] + CASES[4:]])
def test_submit_host_logic_sim(sim_tracker, name, cfg, mono, kw):
    _case(sim_tracker, cfg, mono, kw, 5, 1)


def test_submit_contract_sim(sim_tracker):
    """Two frames may be announced ahead, a third is refused; feeding a frame other than the announced one is refused."""
    cfg = configs.EUROC_MAV.with_(num_features=60)
    frames = frames_for(cfg, 4, stream_id=3)
    masks = masks_for(cfg)
    msgs = [CameraData(0.05 * k, [0, 1], [l, r], masks) for k, (l, r) in enumerate(frames)]
    t = sim_tracker(cfg)
    t.submit(msgs[0])
    t.submit(msgs[1])
    with pytest.raises(UvioError):
        t.submit(msgs[2])
    with pytest.raises(UvioError):
        t.feed_new_camera(msgs[3])
    t.feed_new_camera(msgs[0])
    t.feed_new_camera(msgs[1])
    t.feed_new_camera(msgs[2])                       # nothing announced: uploaded by the feed itself
    assert len(t.new_observations()) > 50


@pytest.mark.gpu
@pytest.mark.parametrize("name,cfg,mono,kw", CASES, ids=[c[0] for c in CASES])
@pytest.mark.parametrize("depth", [0, 1])
def test_submit_gpu_identical_to_plain_feeds(name, cfg, mono, kw, depth):
    _case(lambda c, **k: make_tracker(c, device=0, **k), cfg, mono, kw, 14, depth)


@pytest.mark.gpu
def test_submit_batch_gpu_against_golden():
    """Batched announce + batched feed of 6 trackers on staggered sequences, checked against per-tracker plain feeds."""
    cfg = configs.EUROC_MAV
    n_trk, n = 6, 12
    seqs = [frames_for(cfg, n, stream_id=40 + s) for s in range(n_trk)]
    masks = masks_for(cfg)
    ref = [_plain_run(lambda c: make_tracker(c, device=0), cfg, seqs[s], masks) for s in range(n_trk)]
    trks = [make_tracker(cfg, device=0) for _ in range(n_trk)]
    msgs = [[CameraData(0.05 * k, [0, 1], list(seqs[s][k]), masks) for s in range(n_trk)] for k in range(n)]
    submit_batch(trks, msgs[0])
    for k in range(n):
        if k + 1 < n:
            submit_batch(trks, msgs[k + 1])
        feed_batch(trks, msgs[k])
        for s, t in enumerate(trks):
            assert t.new_observations().tobytes() == ref[s][k][0].tobytes(), f"frame {k} tracker {s}"


@pytest.mark.gpu
def test_submit_then_golden_sequence():
    class Wrap:
        def __init__(self, t):
            self.t, self.pending = t, None

        def __getattr__(self, n):
            return getattr(self.t, n)

    # the golden check feeds frame by frame; announce each frame right before its feed (depth 0 pipelining: same path, no overlap)
    def make(cfg):
        t = make_tracker(cfg, device=0)
        orig = t.feed_new_camera

        def feed(m):
            t.submit(m)
            orig(m)
        t.feed_new_camera = feed
        return t
    assert check_tracker_against_golden("euroc_mav", make, n=30) > 5000
