"""Shared helpers of the parity tests (test infrastructure; the only place besides bench/smoke that touches oracle/)."""
import numpy as np

from uvio_b200 import configs
from uvio_b200.synth import SynthStereoStream, circular_mask


def frames_for(cfg, n, stream_id=0, events=True, start=0):
    s = SynthStereoStream(cfg.width, cfg.height, stream_id, events=events)
    return [s.frame(start + k) for k in range(n)]


def masks_for(cfg):
    if cfg.use_mask:
        m = circular_mask(cfg.width, cfg.height)
        return [m, m.copy()]
    z = np.zeros((cfg.height, cfg.width), np.uint8)
    return [z, z.copy()]


def run_oracle(cfg, frames, masks, mono=False, num_features=None, set_nf_at=None):
    """Feeds the cv2-driven oracle; returns (oracle, per-frame list of (log slice, pts_last, ids_last))."""
    from oracle.trackklt_cv2 import make_oracle
    o = make_oracle(cfg) if num_features is None else make_oracle(cfg, numfeats=num_features)
    out = []
    for k, (l, r) in enumerate(frames):
        if set_nf_at and k in set_nf_at:
            o.set_num_features(set_nf_at[k])
        n0 = len(o.database.log)
        if mono:
            o.feed_new_camera(float(k) * 0.05, [0], [l], [masks[0]])
        else:
            o.feed_new_camera(float(k) * 0.05, [0, 1], [l, r], masks)
        out.append((list(o.database.log[n0:]), {c: list(p) for c, p in o.pts_last.items()},
                    {c: list(i) for c, i in o.ids_last.items()}))
    return o, out


def log_arrays(log):
    """update_feature log -> (ids, cams, uv, uvn)"""
    if not log:
        return (np.zeros(0, np.int64), np.zeros(0, np.int64), np.zeros((0, 2), np.float32), np.zeros((0, 2), np.float32))
    ids = np.array([e[0] for e in log], np.int64)
    cams = np.array([e[2] for e in log], np.int64)
    uv = np.array([[e[3], e[4]] for e in log], np.float32)
    uvn = np.array([[e[5], e[6]] for e in log], np.float32)
    return ids, cams, uv, uvn
