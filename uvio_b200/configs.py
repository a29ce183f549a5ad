"""The five BASELINE.json tracker configurations, concretised as in SURVEY.md section 8(d).

Values come from the reference's config files (config/<dataset>/estimator_config.yaml and kalibr_imucam_chain.yaml);
where BASELINE.json names a different shape / feature count than the file, BASELINE wins and the difference is noted.
"""
from __future__ import annotations

from dataclasses import dataclass, field, replace
from typing import Tuple

HIST_NONE, HIST_HISTOGRAM, HIST_CLAHE = 0, 1, 2     # TrackBase::HistogramMethod (TrackBase.h:78)
CAM_RADTAN, CAM_EQUI = 0, 1                          # CamRadtan.h / CamEqui.h


@dataclass(frozen=True)
class CamCalib:
    model: int
    intrinsics: Tuple[float, float, float, float]     # fx, fy, cx, cy
    dist: Tuple[float, float, float, float]           # k1 k2 p1 p2 (radtan) or k1..k4 (equidistant)


@dataclass(frozen=True)
class TrackerConfig:
    name: str
    width: int
    height: int
    num_features: int            # per camera (the numfeats ctor argument, test_tracking.cpp:179)
    fast_threshold: int
    grid_x: int = 5
    grid_y: int = 5
    min_px_dist: int = 10
    histogram: int = HIST_HISTOGRAM
    pyr_levels: int = 5          # the reference hard-codes 5 (TrackKLT.h:143)
    use_stereo: bool = True
    num_aruco: int = 0
    use_mask: bool = False
    cams: Tuple[CamCalib, CamCalib] = field(default=None)

    def with_(self, **kw) -> "TrackerConfig":
        return replace(self, **kw)


_EUROC = (
    CamCalib(CAM_RADTAN, (458.654, 457.296, 367.215, 248.375), (-0.28340811, 0.07395907, 0.00019359, 1.76187114e-05)),
    CamCalib(CAM_RADTAN, (457.587, 456.134, 379.999, 255.238), (-0.28368365, 0.07451284, -0.00010473, -3.55590700e-05)),
)
_TUMVI = (
    CamCalib(CAM_EQUI, (190.97847715128717, 190.9733070521226, 254.93170605935475, 256.8974428996504),
             (0.0034823894022493434, 0.0007150348452162257, -0.0020532361418706202, 0.00020293673591811182)),
    CamCalib(CAM_EQUI, (190.44236969414825, 190.4344384721956, 252.59949716835982, 254.91723064636983),
             (0.0034003170790442797, 0.001766278153469831, -0.00266312569781606, 0.0003299517423931039)),
)
# rs_d455: the reference file is 848x480 mono; BASELINE.json asks for 640x480 stereo / 400 features / CLAHE.
_D455 = (
    CamCalib(CAM_RADTAN, (416.85223429743274 * 640.0 / 848.0, 414.92069080087543, 421.02459311003213 * 640.0 / 848.0,
                          237.76180565241077),
             (-0.045761895748285604, 0.03423951132164367, -0.00040139057556727315, 0.000431371425853453)),
    CamCalib(CAM_RADTAN, (416.85223429743274 * 640.0 / 848.0, 414.92069080087543, 421.02459311003213 * 640.0 / 848.0,
                          237.76180565241077),
             (-0.045761895748285604, 0.03423951132164367, -0.00040139057556727315, 0.000431371425853453)),
)
_HD = (
    CamCalib(CAM_RADTAN, (780.0, 778.0, 640.5, 361.2), (-0.21, 0.05, 0.0002, -0.0001)),
    CamCalib(CAM_RADTAN, (779.0, 777.5, 645.1, 358.7), (-0.20, 0.048, -0.0001, 0.0002)),
)
_UZH = (
    CamCalib(CAM_EQUI, (278.66723066149086, 278.48991409740296, 319.75221200593535, 241.96858910358173),
             (-0.013721808247486035, 0.020727425669427896, -0.012786476702685545, 0.0025242267320687625)),
    CamCalib(CAM_EQUI, (277.61640629770613, 277.63749695723294, 314.8944703346039, 236.04310050462587),
             (-0.008456929295619607, 0.011407590938612062, -0.006951788325762078, 0.0015368127092821786)),
)

EUROC_MAV = TrackerConfig("euroc_mav", 752, 480, 200, 20, 5, 5, 10, HIST_HISTOGRAM, cams=_EUROC)
TUM_VI = TrackerConfig("tum_vi", 512, 512, 250, 20, 5, 5, 15, HIST_HISTOGRAM, use_mask=True, cams=_TUMVI)
RS_D455 = TrackerConfig("rs_d455", 640, 480, 400, 30, 5, 5, 15, HIST_CLAHE, cams=_D455)
# BASELINE.json configs[3]: "1280x720 stereo high-density tracking, 1000 features, 4 pyramid levels" (no reference file).  The
# reference itself cannot express 4 (TrackKLT.h:143 hard-codes 5): HD_1000_PYR5 is that variant (5 -> effective L = 5 at 1280x720).
HD_1000 = TrackerConfig("hd_1000", 1280, 720, 1000, 20, 5, 5, 10, HIST_HISTOGRAM, pyr_levels=4, cams=_HD)
HD_1000_PYR5 = HD_1000.with_(name="hd_1000_pyr5", pyr_levels=5)
UZHFPV_INDOOR = TrackerConfig("uzhfpv_indoor", 640, 480, 200, 50, 5, 5, 15, HIST_HISTOGRAM, cams=_UZH)

CONFIGS = {c.name: c for c in (EUROC_MAV, TUM_VI, RS_D455, HD_1000, UZHFPV_INDOOR)}
