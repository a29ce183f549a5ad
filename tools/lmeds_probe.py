"""Is cv::findFundamentalMat reproducible in its LMedS regime (10 <= N <= 14 points, what TrackKLT.cpp:848-873 reaches when a
camera holds that few tracks)?  OpenCV compares cv2 with ITSELF: the same inputs are given to cv2.findFundamentalMat(FM_RANSAC,
thr, 0.999) in child processes that differ only in the code path of the linear algebra underneath -- OPENBLAS_CORETYPE (this
cv2 build solves the 7-point null space with LAPACK from the OpenBLAS it bundles; OpenBLAS picks its kernels by CPU model) and
cv2.setUseOptimized -- and, separately, to inputs whose first coordinate is moved by one ulp.  Masks are compared per case.

    python tools/lmeds_probe.py [cases_per_N]      -> one JSON line (committed as profiles/r2_lmeds_probe.json)
"""
import json
import os
import subprocess
import sys

import numpy as np

NS = [10, 11, 12, 13, 14, 15, 20, 40]
VARIANTS = [("default", {}, True), ("Haswell", {"OPENBLAS_CORETYPE": "Haswell"}, True),
            ("Sandybridge", {"OPENBLAS_CORETYPE": "Sandybridge"}, True), ("Prescott", {"OPENBLAS_CORETYPE": "Prescott"}, True),
            ("SkylakeX", {"OPENBLAS_CORETYPE": "SkylakeX"}, True), ("no_opt", {}, False)]


def cases(per_n, seed=7):
    """Normalised correspondences of a moving camera looking at random 3-D points, 0.3 px of noise at f = 460, 0-2 gross outliers."""
    rng = np.random.default_rng(seed)
    out = []
    for n in NS:
        for _ in range(per_n):
            X = np.c_[rng.uniform(-2, 2, n), rng.uniform(-1.5, 1.5, n), rng.uniform(2, 8, n)]
            a = rng.normal(0, 0.02, 3)
            R = np.eye(3) + np.array([[0, -a[2], a[1]], [a[2], 0, -a[0]], [-a[1], a[0], 0]])
            t = rng.normal(0, 0.05, 3)
            Y = X @ R.T + t
            p0 = (X[:, :2] / X[:, 2:]) + rng.normal(0, 0.3 / 460, (n, 2))
            p1 = (Y[:, :2] / Y[:, 2:]) + rng.normal(0, 0.3 / 460, (n, 2))
            for k in range(int(rng.integers(0, 3))):
                p1[int(rng.integers(0, n))] += rng.normal(0, 0.05, 2)
            out.append((p0.astype(np.float32), p1.astype(np.float32)))
    return out


def child(path, use_opt):
    import cv2
    cv2.ipp.setUseIPP(False)
    cv2.setUseOptimized(bool(use_opt))
    cv2.setRNGSeed(0)
    z = np.load(path)
    res = []
    for i in range(int(z["count"])):
        for tag in ("a", "b"):       # a: as given, b: first coordinate + 1 ulp
            p0, p1 = z[f"p0_{i}"].copy(), z[f"p1_{i}"]
            if tag == "b":
                p0[0, 0] = np.nextafter(p0[0, 0], np.float32(10))
            try:
                _F, m = cv2.findFundamentalMat(p0.reshape(-1, 1, 2), p1.reshape(-1, 1, 2), cv2.FM_RANSAC, 2.0 / 460.0, 0.999)
                res.append("".join(str(int(v)) for v in m.reshape(-1)) if m is not None else "none")
            except cv2.error:
                res.append("raised")
    print(json.dumps(res))


def main():
    per_n = int(sys.argv[1]) if len(sys.argv) > 1 else 60
    cs = cases(per_n)
    path = "/tmp/lmeds_cases.npz"
    np.savez(path, count=len(cs), **{f"p0_{i}": c[0] for i, c in enumerate(cs)}, **{f"p1_{i}": c[1] for i, c in enumerate(cs)})
    results = {}
    for name, env, opt in VARIANTS:
        e = dict(os.environ)
        e.update(env)
        out = subprocess.run([sys.executable, os.path.abspath(__file__), "--child", path, str(int(opt))], env=e, capture_output=True, text=True)
        results[name] = json.loads(out.stdout.strip().splitlines()[-1])
    report = {"cases_per_N": per_n, "variants": [v[0] for v in VARIANTS], "per_N": {}}
    for k, n in enumerate(NS):
        idx = range(k * per_n, (k + 1) * per_n)
        base = results["default"]
        differ_variant = sum(any(results[v][2 * i] != base[2 * i] for v in results) for i in idx)
        differ_ulp = sum(base[2 * i] != base[2 * i + 1] for i in idx)
        raised = sum(base[2 * i] in ("raised", "none") for i in idx)
        report["per_N"][str(n)] = {"masks_differing_between_code_paths": differ_variant, "masks_differing_under_1ulp_input_change": differ_ulp,
                                   "raised_or_empty": raised, "cases": per_n}
    import cv2
    report["cv2"] = cv2.__version__
    print(json.dumps(report))


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "--child":
        child(sys.argv[2], int(sys.argv[3]))
    else:
        main()
