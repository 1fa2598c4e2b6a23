"""Generates tests/golden/p3p_cv2.json: P3P solution SETS from OpenCV's independent
solveP3P (cv2 4.x, SOLVEPNP_P3P) for seeded random configurations. Pins the oracle's
Lambda-Twist restatement (oracle/p3p_oracle.h), which the reference delegates to the
un-vendored PoseLib (calibrated_absolute_pose_estimator.cc:120-121)."""
import json
import os
import sys

import cv2
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from ransaclib_b200 import synth  # noqa: E402

rng = np.random.default_rng(2024)
cases = []
for trial in range(60):
    R = synth.random_rotation(rng)
    t = rng.uniform(-1, 1, 3)
    Xc = np.stack([rng.uniform(-1, 1, 3) * 3, rng.uniform(-1, 1, 3) * 3, rng.uniform(2, 10, 3)], 1)
    X = (Xc - t) @ R
    y = Xc / np.linalg.norm(Xc, axis=1, keepdims=True)
    img = (Xc[:, :2] / Xc[:, 2:3]).reshape(3, 1, 2)
    ok, rv, tv = cv2.solveP3P(X.reshape(3, 1, 3), img, np.eye(3), None, flags=cv2.SOLVEPNP_P3P)
    sols = []
    for r, tt in zip(rv, tv):
        Rc = cv2.Rodrigues(r)[0]
        tc = tt.reshape(3)
        if ((X @ Rc.T + tc)[:, 2] > 0).all():
            sols.append({"R": Rc.reshape(-1).tolist(), "t": tc.tolist()})
    cases.append({"y": y.reshape(-1).tolist(), "X": X.reshape(-1).tolist(), "R_gt": R.reshape(-1).tolist(),
                  "t_gt": t.tolist(), "cv2_solutions": sols})
path = os.path.join(HERE, "p3p_cv2.json")
json.dump({"generator": "tests/golden/make_p3p_golden.py, cv2 " + cv2.__version__, "cases": cases}, open(path, "w"))
print("wrote", path, os.path.getsize(path))
