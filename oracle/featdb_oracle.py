"""CPU restatement of ov_core::FeatureDatabase / ov_core::Feature (TEST INFRASTRUCTURE: the checker of the device-side
feature-track store; never imported by the product).  Follows ov_core/src/feat/FeatureDatabase.cpp and Feature.cpp statement by
statement; Python dicts keep insertion order where std::unordered_map's order is unspecified, so results are compared as
sets of features with exact per-camera measurement lists.

Parity unpinned by the reference: it ships no tests or vectors for this class; the restatement is pinned
to the source text only.
"""
from __future__ import annotations

from typing import Dict, List


class Feature:
    """feat/Feature.h:40-97"""

    def __init__(self, featid: int):
        self.featid = featid
        self.to_delete = False
        self.uvs: Dict[int, List[tuple]] = {}
        self.uvs_norm: Dict[int, List[tuple]] = {}
        self.timestamps: Dict[int, List[float]] = {}

    def clean_older_measurements(self, timestamp: float):
        """Feature.cpp:84-106: drop measurements with time <= timestamp"""
        for cam in self.timestamps:
            keep = [i for i, t in enumerate(self.timestamps[cam]) if not t <= timestamp]
            self._keep(cam, keep)

    def clean_invalid_measurements(self, invalid_times):
        """Feature.cpp:60-82: drop measurements whose time is listed"""
        for cam in self.timestamps:
            keep = [i for i, t in enumerate(self.timestamps[cam]) if t not in invalid_times]
            self._keep(cam, keep)

    def _keep(self, cam, keep):
        self.timestamps[cam] = [self.timestamps[cam][i] for i in keep]
        self.uvs[cam] = [self.uvs[cam][i] for i in keep]
        self.uvs_norm[cam] = [self.uvs_norm[cam][i] for i in keep]


class FeatureDatabase:
    def __init__(self):
        self.features_idlookup: Dict[int, Feature] = {}

    def update_feature(self, fid, timestamp, cam_id, u, v, u_n, v_n):
        """FeatureDatabase.cpp:59-85"""
        f = self.features_idlookup.get(fid)
        if f is None:
            f = Feature(fid)
            self.features_idlookup[fid] = f
        f.uvs.setdefault(cam_id, []).append((u, v))
        f.uvs_norm.setdefault(cam_id, []).append((u_n, v_n))
        f.timestamps.setdefault(cam_id, []).append(timestamp)

    def get_feature(self, fid, remove=False):
        """FeatureDatabase.cpp:36-57"""
        f = self.features_idlookup.get(fid)
        if f is not None and remove:
            del self.features_idlookup[fid]
        return f

    def _select(self, pred, remove, skip_deleted):
        out = []
        for key in list(self.features_idlookup):
            f = self.features_idlookup[key]
            if skip_deleted and f.to_delete:
                continue
            if pred(f):
                out.append(f)
                if remove:
                    del self.features_idlookup[key]
        return out

    def features_not_containing_newer(self, timestamp, remove=False, skip_deleted=False):
        """FeatureDatabase.cpp:87-126 (:104: a camera whose newest measurement is >= timestamp keeps the feature out)"""
        return self._select(lambda f: not any(len(t) > 0 and t[-1] >= timestamp for t in f.timestamps.values()), remove, skip_deleted)

    def features_containing_older(self, timestamp, remove=False, skip_deleted=False):
        """FeatureDatabase.cpp:128-167 (:145: a camera whose oldest measurement is < timestamp)"""
        return self._select(lambda f: any(len(t) > 0 and t[0] < timestamp for t in f.timestamps.values()), remove, skip_deleted)

    def features_containing(self, timestamp, remove=False, skip_deleted=False):
        """FeatureDatabase.cpp:169-209 (:189: std::find of the exact timestamp)"""
        return self._select(lambda f: any(timestamp in t for t in f.timestamps.values()), remove, skip_deleted)

    def cleanup(self):
        """FeatureDatabase.cpp:211-224"""
        for key in [k for k, f in self.features_idlookup.items() if f.to_delete]:
            del self.features_idlookup[key]

    def cleanup_measurements(self, timestamp):
        """FeatureDatabase.cpp:226-243"""
        for key in list(self.features_idlookup):
            f = self.features_idlookup[key]
            f.clean_older_measurements(timestamp)
            if sum(len(t) for t in f.timestamps.values()) < 1:
                del self.features_idlookup[key]

    def cleanup_measurements_exact(self, timestamp):
        """FeatureDatabase.cpp:245-263"""
        for key in list(self.features_idlookup):
            f = self.features_idlookup[key]
            f.clean_invalid_measurements([timestamp])
            if sum(len(t) for t in f.timestamps.values()) < 1:
                del self.features_idlookup[key]

    def size(self):
        return len(self.features_idlookup)

    def change_feat_id_database_half(self, id_old, id_new):
        """TrackBase.cpp:233-238: get_internal_data() returns the map by value, so only feat->featid changes"""
        f = self.features_idlookup.get(id_old)
        if f is not None:
            f.featid = id_new
