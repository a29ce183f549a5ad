"""CPU restatement of ov_core::FeatureInitializer (TEST INFRASTRUCTURE: the checker of the batched GPU triangulation; never
imported by the product).  Follows ov_core/src/feat/FeatureInitializer.cpp statement by statement in numpy float64, with the
reference's float casts (z and the residual are Eigen float vectors, :283-285, :413-415); Eigen's colPivHouseholderQr().solve,
JacobiSVD and HouseholderQR become numpy.linalg.solve / svd / qr (same quantities; Eigen itself is not available here).

Parity unpinned by the reference: no tests or vectors for this class exist in it and it cannot be built
here (Eigen3 missing), so this restatement is pinned to the source text only; the GPU kernel is compared with it within the
tolerance tests/test_tri.py states.
"""
from __future__ import annotations

import numpy as np

DEFAULTS = dict(triangulate_1d=False, refine_features=True, max_runs=5, init_lamda=1e-3, max_lamda=1e10, min_dx=1e-6, min_dcost=1e-6,
                lam_mult=10.0, min_dist=0.10, max_dist=60.0, max_baseline=40.0, max_cond_number=10000.0)   # FeatureInitializerOptions.h:35-69


def skew_x(w):
    return np.array([[0, -w[2], w[1]], [w[2], 0, -w[0]], [-w[1], w[0], 0]], np.float64)


class Feat:
    """the fields of ov_core::Feature the initializer touches; cams are visited from the highest id down (the iteration order
    of libstdc++'s unordered_map for a feature first seen in camera 0 -- the reference's order is unspecified)"""

    def __init__(self, timestamps, uvs_norm):
        self.timestamps = {c: list(t) for c, t in sorted(timestamps.items(), reverse=True)}
        self.uvs_norm = {c: [np.asarray(x, np.float32) for x in uvs_norm[c]] for c in self.timestamps}
        self.anchor_cam_id = -1
        self.anchor_clone_timestamp = 0.0
        self.p_FinA = np.zeros(3)
        self.p_FinG = np.zeros(3)


class FeatureInitializer:
    def __init__(self, **opt):
        self.o = dict(DEFAULTS)
        self.o.update(opt)

    def _anchor(self, feat):                                    # :36-47
        most = 0
        for cam, ts in feat.timestamps.items():
            if len(ts) > most:
                feat.anchor_cam_id, most = cam, len(ts)
        feat.anchor_clone_timestamp = feat.timestamps[feat.anchor_cam_id][-1]

    def single_triangulation(self, feat, clones):               # :30-112
        self._anchor(feat)
        A, b = np.zeros((3, 3)), np.zeros(3)
        R_GtoA, p_AinG = clones[feat.anchor_cam_id][feat.anchor_clone_timestamp]
        for cam, ts in feat.timestamps.items():
            for m, t in enumerate(ts):
                R_GtoCi, p_CiinG = clones[cam][t]
                R_AtoCi = R_GtoCi @ R_GtoA.T
                p_CiinA = R_GtoA @ (p_CiinG - p_AinG)
                b_i = np.array([feat.uvs_norm[cam][m][0], feat.uvs_norm[cam][m][1], 1.0], np.float64)
                b_i = R_AtoCi.T @ b_i
                b_i = b_i / np.linalg.norm(b_i)
                Bperp = skew_x(b_i)
                Ai = Bperp.T @ Bperp
                A += Ai
                b += Ai @ p_CiinA
        try:
            p_f = np.linalg.solve(A, b)
        except np.linalg.LinAlgError:
            return False
        sv = np.linalg.svd(A, compute_uv=False)
        condA = sv[0] / sv[-1]
        if abs(condA) > self.o["max_cond_number"] or p_f[2] < self.o["min_dist"] or p_f[2] > self.o["max_dist"] or np.isnan(np.linalg.norm(p_f)):
            return False
        feat.p_FinA = p_f
        feat.p_FinG = R_GtoA.T @ feat.p_FinA + p_AinG
        return True

    def single_triangulation_1d(self, feat, clones):            # :114-195
        self._anchor(feat)
        idx = len(feat.timestamps[feat.anchor_cam_id]) - 1
        A = b = 0.0
        R_GtoA, p_AinG = clones[feat.anchor_cam_id][feat.anchor_clone_timestamp]
        u = feat.uvs_norm[feat.anchor_cam_id][idx]
        bearing = np.array([u[0], u[1], 1.0], np.float64)
        bearing = bearing / np.linalg.norm(bearing)
        for cam, ts in feat.timestamps.items():
            for m, t in enumerate(ts):
                if cam == feat.anchor_cam_id and m == idx:
                    continue
                R_GtoCi, p_CiinG = clones[cam][t]
                R_AtoCi = R_GtoCi @ R_GtoA.T
                p_CiinA = R_GtoA @ (p_CiinG - p_AinG)
                b_i = np.array([feat.uvs_norm[cam][m][0], feat.uvs_norm[cam][m][1], 1.0], np.float64)
                b_i = R_AtoCi.T @ b_i
                b_i = b_i / np.linalg.norm(b_i)
                Bperp = skew_x(b_i)
                BB = Bperp @ bearing
                A += BB @ BB
                b += BB @ (Bperp @ p_CiinA)
        with np.errstate(all="ignore"):
            p_f = (np.float64(b) / np.float64(A)) * bearing
        if p_f[2] < self.o["min_dist"] or p_f[2] > self.o["max_dist"] or np.isnan(np.linalg.norm(p_f)):
            return False
        feat.p_FinA = p_f
        feat.p_FinG = R_GtoA.T @ feat.p_FinA + p_AinG
        return True

    def _terms(self, feat, clones, alpha, beta, rho, jac):
        err = 0.0
        Hess, grad = np.zeros((3, 3)), np.zeros(3)
        R_GtoA, p_AinG = clones[feat.anchor_cam_id][feat.anchor_clone_timestamp]
        for cam, ts in feat.timestamps.items():
            for m, t in enumerate(ts):
                R_GtoCi, p_CiinG = clones[cam][t]
                R = R_GtoCi @ R_GtoA.T
                p_CiinA = R_GtoA @ (p_CiinG - p_AinG)
                pA = -R @ p_CiinA
                hi1 = R[0, 0] * alpha + R[0, 1] * beta + R[0, 2] + rho * pA[0]
                hi2 = R[1, 0] * alpha + R[1, 1] * beta + R[1, 2] + rho * pA[1]
                hi3 = R[2, 0] * alpha + R[2, 1] * beta + R[2, 2] + rho * pA[2]
                with np.errstate(all="ignore"):
                    z = np.array([hi1 / hi3, hi2 / hi3]).astype(np.float32)          # Eigen::Matrix<float,2,1>
                    res = feat.uvs_norm[cam][m] - z                                   # float
                    rn = np.sqrt(res[0] * res[0] + res[1] * res[1], dtype=np.float32)  # res.norm() in float
                    err += float(rn) ** 2
                    if jac:
                        h32 = hi3 * hi3
                        H = np.array([[(R[0, 0] * hi3 - hi1 * R[2, 0]) / h32, (R[0, 1] * hi3 - hi1 * R[2, 1]) / h32, (pA[0] * hi3 - hi1 * pA[2]) / h32],
                                      [(R[1, 0] * hi3 - hi2 * R[2, 0]) / h32, (R[1, 1] * hi3 - hi2 * R[2, 1]) / h32, (pA[1] * hi3 - hi2 * pA[2]) / h32]])
                        grad += H.T @ res.astype(np.float64)
                        Hess += H.T @ H
        return err, Hess, grad

    def compute_error(self, clones, feat, alpha, beta, rho):     # :362-421
        return self._terms(feat, clones, alpha, beta, rho, False)[0]

    def single_gaussnewton(self, feat, clones):                  # :197-360
        o = self.o
        rho = 1 / feat.p_FinA[2]
        alpha, beta = feat.p_FinA[0] / feat.p_FinA[2], feat.p_FinA[1] / feat.p_FinA[2]
        lam, eps, runs, recompute = o["init_lamda"], 10000.0, 0, True
        Hess, grad = np.zeros((3, 3)), np.zeros(3)
        cost_old = self.compute_error(clones, feat, alpha, beta, rho)
        R_GtoA, p_AinG = clones[feat.anchor_cam_id][feat.anchor_clone_timestamp]
        guard = 0
        while runs < o["max_runs"] and lam < o["max_lamda"] and eps > o["min_dx"] and guard < 10000:
            guard += 1
            if recompute:
                _e, Hess, grad = self._terms(feat, clones, alpha, beta, rho, True)
            Hl = Hess.copy()
            for r in range(3):
                Hl[r, r] *= (1.0 + lam)
            try:
                dx = np.linalg.solve(Hl, grad)
            except np.linalg.LinAlgError:
                dx = np.full(3, np.nan)
            cost = self.compute_error(clones, feat, alpha + dx[0], beta + dx[1], rho + dx[2])
            with np.errstate(all="ignore"):
                converged = cost <= cost_old and (cost_old - cost) / cost_old < o["min_dcost"]
            if converged:
                alpha, beta, rho, eps = alpha + dx[0], beta + dx[1], rho + dx[2], 0
                break
            if cost <= cost_old:
                recompute, cost_old = True, cost
                alpha, beta, rho = alpha + dx[0], beta + dx[1], rho + dx[2]
                runs += 1
                lam = lam / o["lam_mult"]
                eps = np.linalg.norm(dx)
            else:
                recompute = False
                lam = lam * o["lam_mult"]
        feat.p_FinA = np.array([alpha / rho, beta / rho, 1 / rho])
        Q, _r = np.linalg.qr(feat.p_FinA.reshape(3, 1), mode="complete")
        base_max = 0.0
        for cam, ts in feat.timestamps.items():
            for t in ts:
                p_CiinA = R_GtoA @ (clones[cam][t][1] - p_AinG)
                base_max = max(base_max, float(np.linalg.norm(Q[:, 1:3].T @ p_CiinA)))
        n = np.linalg.norm(feat.p_FinA)
        with np.errstate(all="ignore"):
            bad = feat.p_FinA[2] < o["min_dist"] or feat.p_FinA[2] > o["max_dist"] or (n / base_max) > o["max_baseline"] or np.isnan(n)
        if bad:
            return False
        feat.p_FinG = R_GtoA.T @ feat.p_FinA + p_AinG
        return True
