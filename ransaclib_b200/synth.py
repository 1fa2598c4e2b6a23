"""Seeded synthetic instances with the distributions of the reference's example
generators (numpy restatements; the reference seeds from std::random_device, so
there is no stream to match -- only the distributions):

* line:    examples/line_estimation.cc:49-106
* hybrid:  examples/hybrid_line_estimation.cc:50-127
* pose:    examples/camera_pose_estimation.cc:52-127

Layouts are the reference's host layouts: points as (n,2)/(n,3)/(n,4) C-contiguous
float64 arrays == column-major Eigen::Matrix2Xd / Matrix4Xd == std::vector<Vector*d>.
"""
from __future__ import annotations

import math

import numpy as np


def _rot2(angle):
    c, s = math.cos(angle), math.sin(angle)
    return np.array([[c, -s], [s, c]])


def line_instance(seed: int, num_inliers: int, num_outliers: int, inlier_threshold: float = 0.005):
    """(n,2) points; inliers within +-inlier_threshold of a random line."""
    rng = np.random.default_rng(seed)
    n = num_inliers + num_outliers
    perm = rng.permutation(n)
    pts = np.empty((n, 2))
    x = np.arange(num_inliers) / max(num_inliers, 1)
    y = rng.uniform(-inlier_threshold, inlier_threshold, num_inliers)
    pts[perm[:num_inliers], 0] = x
    pts[perm[:num_inliers], 1] = y
    ox = rng.uniform(0.0, 1.0, num_outliers)
    oy = rng.uniform(-0.5, 0.5, num_outliers)
    bad = np.abs(oy) < 5.0 * inlier_threshold
    while bad.any():
        oy[bad] = rng.uniform(-0.5, 0.5, int(bad.sum()))
        bad = np.abs(oy) < 5.0 * inlier_threshold
    pts[perm[num_inliers:], 0] = ox
    pts[perm[num_inliers:], 1] = oy
    R = _rot2(rng.uniform(-0.5, 0.5) * math.pi)
    t = rng.uniform(-0.5, 0.5, 2) * 10.0
    return np.ascontiguousarray(pts @ R.T + t)


def hybrid_line_instance(seed: int, n0_in: int, n0_out: int, n1_in: int, n1_out: int,
                         inlier_threshold: float = 0.005):
    """Returns (points (n0,2), points_with_normals (n1,4)) on one random line."""
    rng = np.random.default_rng(seed)
    R = _rot2(rng.uniform(-0.5, 0.5) * math.pi)
    t = rng.uniform(-0.5, 0.5, 2) * 10.0

    def gen(n_in, n_out):
        n = n_in + n_out
        perm = rng.permutation(n)
        d = np.zeros((n, 4))
        d[perm[:n_in], 0] = np.arange(n_in) / max(n_in, 1)
        d[perm[:n_in], 1] = rng.uniform(-inlier_threshold, inlier_threshold, n_in)
        d[perm[:n_in], 3] = 1.0
        oy = rng.uniform(-0.5, 0.5, n_out)
        bad = np.abs(oy) < 5.0 * inlier_threshold
        while bad.any():
            oy[bad] = rng.uniform(-0.5, 0.5, int(bad.sum()))
            bad = np.abs(oy) < 5.0 * inlier_threshold
        d[perm[n_in:], 0] = rng.uniform(0.0, 1.0, n_out)
        d[perm[n_in:], 1] = oy
        ang = rng.uniform(0.0, 2.0 * math.pi, n_out)
        d[perm[n_in:], 2] = np.cos(ang)
        d[perm[n_in:], 3] = np.sin(ang)
        d[:, :2] = d[:, :2] @ R.T + t
        d[:, 2:] = d[:, 2:] @ R.T
        return d

    a = gen(n0_in, n0_out)
    b = gen(n1_in, n1_out)
    return np.ascontiguousarray(a[:, :2]), np.ascontiguousarray(b)


def random_rotation(rng):
    q = rng.normal(size=4)
    q /= np.linalg.norm(q)
    w, x, y, z = q
    return np.array([
        [1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
        [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
        [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)],
    ])


POSE_WIDTH = 640.0
POSE_HEIGHT = 320.0
POSE_FOCAL = (POSE_WIDTH * 0.5) / math.tan(60.0 * math.pi / 180.0)  # camera_pose_estimation.cc:151-153
POSE_THRESH_PX = 12.0


def pose_instance(seed: int, num_inliers: int, num_outliers: int, width: float = POSE_WIDTH,
                  height: float = POSE_HEIGHT, focal: float = POSE_FOCAL, pixel_noise: float = 2.0,
                  min_depth: float = 2.0, max_depth: float = 10.0):
    """Returns dict(points2D (n,2), points3D (n,3), R_gt, c_gt, focal).

    World points are X_w = Rg X_cam + tg, so the ground-truth [R|c] pose (R world->cam,
    c camera centre) is R = Rg^T, c = tg.
    """
    rng = np.random.default_rng(seed)
    n = num_inliers + num_outliers
    perm = rng.permutation(n)
    p2 = np.empty((n, 2))
    p3 = np.empty((n, 3))

    def backproject(px):
        d = np.concatenate([px / focal, np.ones((len(px), 1))], axis=1)
        d /= np.linalg.norm(d, axis=1, keepdims=True)
        return d * rng.uniform(min_depth, max_depth, (len(px), 1))

    pin = np.stack([rng.uniform(-width / 2, width / 2, num_inliers),
                    rng.uniform(-height / 2, height / 2, num_inliers)], axis=1)
    p3[perm[:num_inliers]] = backproject(pin)
    p2[perm[:num_inliers]] = pin + rng.uniform(-1.0, 1.0, (num_inliers, 2)) * pixel_noise

    pout = np.stack([rng.uniform(-width / 2, width / 2, num_outliers),
                     rng.uniform(-height / 2, height / 2, num_outliers)], axis=1)
    p3[perm[num_inliers:]] = backproject(pout)
    far = np.stack([rng.uniform(-width / 2, width / 2, num_outliers),
                    rng.uniform(-height / 2, height / 2, num_outliers)], axis=1)
    bad = np.linalg.norm(far - pout, axis=1) < 10.0 * (pixel_noise + 1.0)
    while bad.any():
        k = int(bad.sum())
        far[bad] = np.stack([rng.uniform(-width / 2, width / 2, k),
                             rng.uniform(-height / 2, height / 2, k)], axis=1)
        bad = np.linalg.norm(far - pout, axis=1) < 10.0 * (pixel_noise + 1.0)
    p2[perm[num_inliers:]] = far

    Rg = random_rotation(rng)
    tg = rng.uniform(-1.0, 1.0, 3) * rng.uniform(1.0, 2.0)
    p3 = p3 @ Rg.T + tg
    return {
        "points2D": np.ascontiguousarray(p2),
        "points3D": np.ascontiguousarray(p3),
        "R_gt": Rg.T.copy(),
        "c_gt": tg.copy(),
        "focal": float(focal),
    }


def hybrid_pose_instance(seed: int, n0_in: int, n0_out: int, n1_in: int, n1_out: int, depth_noise: float = 0.01):
    """2D-3D matches (type 0, as pose_instance) plus 3D-3D matches (type 1: a point measured in the camera frame
    against its world position) of the same camera. Returns dict(points2D, points3D, cam1 (n1,3), world1 (n1,3),
    R_gt, c_gt, focal)."""
    inst = pose_instance(seed, n0_in, n0_out)
    rng = np.random.default_rng(seed + 77)
    R, c = inst["R_gt"], inst["c_gt"]
    n1 = n1_in + n1_out
    perm = rng.permutation(n1)
    cam = np.empty((n1, 3))
    cam[:, 2] = rng.uniform(2.0, 10.0, n1)
    cam[:, 0] = rng.uniform(-1.5, 1.5, n1) * cam[:, 2]
    cam[:, 1] = rng.uniform(-0.8, 0.8, n1) * cam[:, 2]
    world = (cam @ R) + c  # x_cam = R (X - c)
    meas = cam.copy()
    meas[perm[:n1_in]] += rng.uniform(-1.0, 1.0, (n1_in, 3)) * depth_noise
    off = rng.normal(size=(n1_out, 3))
    off *= (rng.uniform(0.5, 3.0, (n1_out, 1)) / np.linalg.norm(off, axis=1, keepdims=True))
    meas[perm[n1_in:]] += off
    inst["cam1"] = np.ascontiguousarray(meas)
    inst["world1"] = np.ascontiguousarray(world)
    return inst


def pose_to_model(R, c):
    """[R|c] as the reference's column-major 3x4 (12 doubles)."""
    return np.concatenate([np.asarray(R).T.reshape(-1), np.asarray(c).reshape(-1)])


def perturbed_poses(seed: int, R, c, k: int, rot_sigma: float = 0.02, trans_sigma: float = 0.05):
    """k models around (R,c): a spread of good and bad hypotheses for scorer tests/benches."""
    rng = np.random.default_rng(seed)
    out = np.empty((k, 12))
    for i in range(k):
        w = rng.normal(size=3) * rot_sigma * (1 + 20 * (i % 7 == 0))
        th = np.linalg.norm(w)
        K = np.array([[0, -w[2], w[1]], [w[2], 0, -w[0]], [-w[1], w[0], 0]])
        dR = np.eye(3) + (math.sin(th) / th) * K + ((1 - math.cos(th)) / th ** 2) * (K @ K) if th > 0 else np.eye(3)
        out[i] = pose_to_model(dR @ R, c + rng.normal(size=3) * trans_sigma)
    return out
