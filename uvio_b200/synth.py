"""Deterministic synthetic stereo frame generator (EuRoC-shaped and friends).

Test / bench infrastructure only (SURVEY.md section 8(d) "Synthetic inputs"): a static random world texture is
rendered through a smooth camera motion (translation <= 8 px/frame, rotation <= 1 deg/frame, scale <= 1 %/frame); the
right camera is the left one shifted by a row-dependent disparity of 5..40 px.  Additive integer noise (sigma = 2), a
slow gain/bias drift, 5 % jump frames (40 px) and 2 % blank frames exercise the LK failure / reset paths of the tracker
(TrackKLT.cpp:143-152,285-299 in the reference).

Everything is elementwise IEEE float64 or integer arithmetic on seeded PCG64 draws, so the same frames come out on every
machine; the oracle and the GPU tracker are always fed the very same arrays.
"""
from __future__ import annotations

import numpy as np

SEED_BASE = 0xC0FFEE


def _upsample_bilinear(grid: np.ndarray, h: int, w: int) -> np.ndarray:
    gh, gw = grid.shape
    ys = np.linspace(0.0, gh - 1.0, h)
    xs = np.linspace(0.0, gw - 1.0, w)
    y0 = np.minimum(np.floor(ys).astype(np.int64), gh - 2)
    x0 = np.minimum(np.floor(xs).astype(np.int64), gw - 2)
    fy = (ys - y0)[:, None]
    fx = (xs - x0)[None, :]
    g00 = grid[y0][:, x0]
    g01 = grid[y0][:, x0 + 1]
    g10 = grid[y0 + 1][:, x0]
    g11 = grid[y0 + 1][:, x0 + 1]
    return (g00 * (1 - fx) + g01 * fx) * (1 - fy) + (g10 * (1 - fx) + g11 * fx) * fy


def make_world(h: int, w: int, rng: np.random.Generator) -> np.ndarray:
    """Multi-octave smooth noise + many small high-contrast rectangles (FAST fires at thresholds 20..50)."""
    world = np.zeros((h, w), np.float64)
    amp = 1.0
    for cell in (128, 64, 32, 16, 8):
        gh, gw = h // cell + 2, w // cell + 2
        g = rng.integers(0, 1 << 16, size=(gh, gw)).astype(np.float64) / 65535.0
        world += amp * _upsample_bilinear(g, h, w)
        amp *= 0.6
    world -= world.min()
    world /= world.max()
    world = 40.0 + 110.0 * world
    # rectangles: corners for the detector
    n_rect = (h * w) // 220
    ys = rng.integers(0, h, n_rect)
    xs = rng.integers(0, w, n_rect)
    hs = rng.integers(3, 14, n_rect)
    ws = rng.integers(3, 14, n_rect)
    dv = rng.integers(25, 90, n_rect) * (rng.integers(0, 2, n_rect) * 2 - 1)
    for i in range(n_rect):
        world[ys[i]:ys[i] + hs[i], xs[i]:xs[i] + ws[i]] += float(dv[i])
    # light 3x3 box blur (exact in float64 order: fixed sequence of adds)
    p = np.pad(world, 1, mode="edge")
    blur = (p[:-2, :-2] + p[:-2, 1:-1] + p[:-2, 2:] + p[1:-1, :-2] + p[1:-1, 1:-1] + p[1:-1, 2:] +
            p[2:, :-2] + p[2:, 1:-1] + p[2:, 2:]) / 9.0
    world = 0.5 * world + 0.5 * blur
    return np.clip(world, 30.0, 180.0)


class SynthStereoStream:
    """One synthetic stereo rig. ``frame(k)`` is a pure function of (seed, k): random access, repeatable."""

    def __init__(self, width: int, height: int, stream_id: int = 0, events: bool = True,
                 max_disp: float = 40.0, speed: float = 1.0):
        self.w, self.h = int(width), int(height)
        self.stream_id = int(stream_id)
        self.seed = SEED_BASE + self.stream_id
        self.events = events
        self.speed = float(speed)
        self.max_disp = float(max_disp)
        rng = np.random.default_rng(self.seed)
        self.ax, self.ay = 250.0, 180.0
        margin_x = int(self.ax + 0.75 * self.w * 0.25 + 160)
        margin_y = int(self.ay + 0.75 * self.h * 0.25 + 160)
        self.wh, self.ww = self.h + 2 * margin_y, self.w + 2 * margin_x + int(max_disp) + 8
        self.world = make_world(self.wh, self.ww, rng)
        self.phase = rng.random(6) * 2.0 * np.pi
        yy, xx = np.mgrid[0:self.h, 0:self.w]
        self.xc = xx.astype(np.float64) - 0.5 * self.w
        self.yc = yy.astype(np.float64) - 0.5 * self.h
        self.yrel = yy.astype(np.float64) / float(self.h)

    # -- motion model ------------------------------------------------------------------------------------------------
    def _pose(self, k: int):
        t = float(k) * self.speed
        p = self.phase
        cx = 0.5 * self.ww + self.ax * np.sin(2.0 * np.pi * t / 200.0 + p[0])       # <= 7.9 px/frame
        cy = 0.5 * self.wh + self.ay * np.sin(2.0 * np.pi * t / 160.0 + p[1])       # <= 7.1 px/frame
        th = 0.15 * np.sin(2.0 * np.pi * t / 170.0 + p[2])                          # <= 0.32 deg/frame
        sc = 1.0 + 0.08 * np.sin(2.0 * np.pi * t / 230.0 + p[3])                    # <= 0.22 %/frame
        return cx, cy, th, sc

    def _event(self, k: int):
        """(jump_px, blank) for frame k, a pure function of (seed, k)."""
        if not self.events or k < 5:
            return 0.0, False
        r = np.random.default_rng([self.seed, 7919, int(k)]).random(2)
        jump = 40.0 if r[0] < 0.05 else 0.0
        blank = bool(r[1] < 0.02)
        return jump, blank

    def _sample(self, wx: np.ndarray, wy: np.ndarray) -> np.ndarray:
        wx = np.clip(wx, 0.0, self.ww - 1.001)
        wy = np.clip(wy, 0.0, self.wh - 1.001)
        x0 = np.floor(wx).astype(np.int64)
        y0 = np.floor(wy).astype(np.int64)
        fx = wx - x0
        fy = wy - y0
        w = self.world
        return ((w[y0, x0] * (1.0 - fx) + w[y0, x0 + 1] * fx) * (1.0 - fy) +
                (w[y0 + 1, x0] * (1.0 - fx) + w[y0 + 1, x0 + 1] * fx) * fy)

    def frame(self, k: int):
        """Returns (left, right) uint8 images of frame k."""
        cx, cy, th, sc = self._pose(k)
        jump, blank = self._event(k)
        if blank:
            v = np.uint8(96)
            return np.full((self.h, self.w), v, np.uint8), np.full((self.h, self.w), v, np.uint8)
        c, s = np.cos(th) * sc, np.sin(th) * sc
        wx = cx + jump + c * self.xc - s * self.yc
        wy = cy + s * self.xc + c * self.yc
        disp = 5.0 + (self.max_disp - 5.0) * self.yrel
        gain = 1.0 + 0.10 * np.sin(2.0 * np.pi * k / 90.0 + self.phase[4])
        bias = 8.0 * np.sin(2.0 * np.pi * k / 130.0 + self.phase[5])
        out = []
        for cam, off in ((0, 0.0), (1, 1.0)):
            img = self._sample(wx + off * disp, wy) * gain + bias
            nrng = np.random.default_rng([self.seed, 104729, int(k), cam])
            noise = nrng.integers(-2, 3, size=(2, self.h, self.w)).sum(axis=0)
            img = np.clip(np.rint(img) + noise, 0, 255).astype(np.uint8)
            out.append(img)
        return out[0], out[1]

    def frames(self, n: int, start: int = 0):
        return [self.frame(start + k) for k in range(n)]


def circular_mask(width: int, height: int, radius_frac: float = 0.52) -> np.ndarray:
    """Fisheye-vignette style mask (255 = ignore), the synthetic equivalent of config/tum_vi/mask_tumvi*.png."""
    yy, xx = np.mgrid[0:height, 0:width]
    r2 = (xx - 0.5 * (width - 1)) ** 2 + (yy - 0.5 * (height - 1)) ** 2
    rad = radius_frac * min(width, height)
    return np.where(r2 > rad * rad, 255, 0).astype(np.uint8)


def frames_checksum(frames) -> int:
    """Order-sensitive 64-bit checksum of a list of uint8 arrays (fixtures record it to detect generator drift)."""
    acc = np.uint64(1469598103934665603)
    for f in frames:
        a = np.ascontiguousarray(f).view(np.uint8).ravel().astype(np.uint64)
        idx = (np.arange(a.size, dtype=np.uint64) % np.uint64(65521)) + np.uint64(1)
        with np.errstate(over="ignore"):
            acc = acc * np.uint64(1099511628211) + np.uint64((a * idx).sum(dtype=np.uint64))
    return int(acc)
