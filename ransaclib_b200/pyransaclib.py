"""Drop-in for the reference's Python module ``pyransaclib`` (examples/pyransaclib.cc:38-46,
examples/ransaclib_localization.cc:58-165): same function name, keyword arguments and result
dictionary. Usage:  ``import pyransaclib`` (the one-file shim at the repository root) or
``from ransaclib_b200 import pyransaclib``.

Additions allowed by the survey: an optional trailing ``random_seed`` (the reference seeds from
std::random_device, ransaclib_localization.cc:109-110, so it is not reproducible) and the batched
sibling ``ransaclib_localization_batch``.
"""
from __future__ import annotations

from . import host


def ransaclib_localization(query_name, focal_x, focal_y, points2D_vec, points3D_vec, inlier_threshold,
                           number_lo_steps, min_num_iterations, max_num_iterations, random_seed=None,
                           device=0):
    """Absolute pose estimation with non-linear refinement.

    Returns ``{"success": False}`` for <= 3 matches, else ``{"success", "qvec" [w,x,y,z], "tvec",
    "num_inliers", "inliers"}`` (ransaclib_localization.cc:86-92,157-162)."""
    if random_seed is None:
        import random

        random_seed = random.SystemRandom().getrandbits(32)
    r = host.localize(focal_x, focal_y, points2D_vec, points3D_vec, inlier_threshold, number_lo_steps,
                      min_num_iterations, max_num_iterations, random_seed=random_seed, device=device)
    r.pop("_stats", None)
    return r


def ransaclib_localization_batch(query_names, focal_x, focal_y, points2D_list, points3D_list, inlier_threshold,
                                 number_lo_steps, min_num_iterations, max_num_iterations, random_seeds=None,
                                 device=0):
    """Independent queries (they share nothing: examples/localization.cc:190), one result dict per
    query, all of them estimated by ONE launch of the device-resident LO-MSAC kernel
    (``rlb_localize_batch``: one CTA per query, no host round trips)."""
    nq = len(query_names)
    if random_seeds is None:
        import random

        rng = random.SystemRandom()
        random_seeds = [rng.getrandbits(32) for _ in range(nq)]
    fx = focal_x if hasattr(focal_x, "__len__") else [focal_x] * nq
    fy = focal_y if hasattr(focal_y, "__len__") else [focal_y] * nq
    out = host.localize_batch(fx, fy, points2D_list, points3D_list, inlier_threshold, number_lo_steps,
                              min_num_iterations, max_num_iterations, random_seeds, device=device)
    for r in out:
        r.pop("_stats", None)
    return out
