"""The localization command-line tool (examples/localization.cc -> ransaclib_b200/lib/localization): the file formats
of the reference's examples/localization.cc, localization_with_gt.cc and localization_gc.cc (list file, match files,
poses / stats output). Parsing is checked on the CPU with --dry-run; on the GPU the poses written by the tool are
compared with the oracle's sequential LO-MSAC on the parsed matches, same seeds, same per-tool options."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

import oracle_lib as ol
from ransaclib_b200 import synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TOOL = os.path.join(ROOT, "ransaclib_b200", "lib", "localization")
POSTFIX = ".matches.txt"

# (model, intrinsics after "w h"; fx, fy as the tools see them are derived below)
MODELS = ["SIMPLE_RADIAL", "PINHOLE", "OPENCV", "VSFM", "BROWN_3_PARAMS"]


def quat_to_R(q):
    w, x, y, z = q
    return np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)],
                     [2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)],
                     [2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)]])


def R_to_quat(R):
    """Eigen::Quaterniond(R) (w, x, y, z), normalised."""
    t = np.trace(R)
    if t > 0:
        s = np.sqrt(t + 1.0)
        w = 0.5 * s
        s = 0.5 / s
        q = [w, (R[2, 1] - R[1, 2]) * s, (R[0, 2] - R[2, 0]) * s, (R[1, 0] - R[0, 1]) * s]
    else:
        i = 0
        if R[1, 1] > R[0, 0]:
            i = 1
        if R[2, 2] > R[i, i]:
            i = 2
        j, k = (i + 1) % 3, (i + 2) % 3
        s = np.sqrt(R[i, i] - R[j, j] - R[k, k] + 1.0)
        v = [0.0, 0.0, 0.0]
        v[i] = 0.5 * s
        s = 0.5 / s
        w = (R[k, j] - R[j, k]) * s
        v[j] = (R[j, i] + R[i, j]) * s
        v[k] = (R[k, i] + R[i, k]) * s
        q = [w] + v
    q = np.array(q)
    return q / np.linalg.norm(q)


def write_dataset(tmp, dialect, num=6, n_in=150, n_out=150, invert=True, centered=False, models=MODELS):
    """Writes a list file + match files; returns the per-query ground truth of what a tool must parse."""
    queries = []
    lines = []
    for i in range(num):
        model = models[i % len(models)]
        inst = synth.pose_instance(700 + i, n_in, n_out)
        f = inst["focal"]
        cx, cy = 320.0 + i, 160.0 - 0.5 * i
        fy_file = f * 1.25 if model in ("PINHOLE", "OPENCV", "BROWN_3_PARAMS") else f
        name = os.path.join(str(tmp), f"img_{i:03d}.jpg")
        if model in ("SIMPLE_RADIAL", "VSFM"):
            intr = [f, cx, cy, 0.01]
        elif model == "PINHOLE":
            intr = [f, fy_file, cx, cy]
        elif model == "OPENCV":
            intr = [f, fy_file, cx, cy, 0.01, -0.02, 0.001, 0.002]
        else:
            intr = [f, fy_file, cx, cy, 0.01, -0.02, 0.003]
        fy_seen = fy_file
        if dialect == "plain" and model == "PINHOLE":
            fy_seen = f  # localization.cc:91-92 overwrites fy with fx
        # the pixels a tool must end up with are generated for (f, fy_seen)
        p2 = inst["points2D"].copy()
        p2[:, 1] *= fy_seen / f
        p3 = inst["points3D"]
        R_gt, c_gt = inst["R_gt"], inst["c_gt"]
        gt = []
        if dialect == "gt":
            gt = list(R_to_quat(R_gt)) + list(c_gt)
        lines.append(" ".join([name, model, "640", "320"] + [repr(float(v)) for v in intr + gt]))
        u = p2[:, 0] + (0.0 if centered else cx)
        v = p2[:, 1] + (0.0 if centered else cy)
        sgn = -1.0 if invert else 1.0
        with open(name + POSTFIX, "w") as fh:
            for j in range(len(p2)):
                row = (u[j], v[j], p3[j, 0], sgn * p3[j, 1], sgn * p3[j, 2])
                fh.write(" ".join(repr(float(x)) for x in row) + "\n")
        # exactly what the tool computes from the text
        pu = u - (0.0 if centered else cx)
        pv = v - (0.0 if centered else cy)
        P3 = np.stack([p3[:, 0], sgn * (sgn * p3[:, 1]), sgn * (sgn * p3[:, 2])], 1)
        queries.append(dict(name=name, model=model, fx=f, fy=fy_seen, p2=np.stack([pu, pv], 1), p3=P3, R_gt=R_gt,
                            c_gt=c_gt))
    listfile = os.path.join(str(tmp), "list.txt")
    with open(listfile, "w") as fh:
        fh.write("\n".join(lines) + "\n")
    return listfile, queries


def run_tool(args, check=True):
    assert os.path.exists(TOOL), "run __graft_entry__.build() first"
    r = subprocess.run([TOOL] + [str(a) for a in args], capture_output=True, text=True, timeout=600)
    if check:
        assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    return r


def parse_dry(stdout):
    out = {}
    for ln in stdout.splitlines():
        if ln.startswith("DRY "):
            t = ln.split()
            out[t[1]] = dict(model=t[2], w=int(t[3]), h=int(t[4]), fx=float(t[5]), fy=float(t[6]), n=int(t[7]),
                             s2=float(t[8]), s3=float(t[9]))
    return out


@pytest.mark.parametrize("dialect,invert,centered", [("plain", True, False), ("plain", False, True), ("gt", True, False)])
def test_dry_run_parses_lists_and_matches(tmp_path, dialect, invert, centered):
    listfile, queries = write_dataset(tmp_path, dialect, invert=invert, centered=centered)
    # a query with 3 matches (skipped), one whose match file is missing, one blank line in a match file
    with open(queries[4]["name"] + POSTFIX) as fh:
        keep = fh.readlines()[:3]
    with open(queries[4]["name"] + POSTFIX, "w") as fh:
        fh.writelines(keep)
    os.remove(queries[5]["name"] + POSTFIX)
    with open(queries[0]["name"] + POSTFIX, "a") as fh:
        fh.write("\n   \n")
    flags = ["--gt"] if dialect == "gt" else []
    r = run_tool(flags + ["--dry-run", listfile, tmp_path / "out.txt", 12.0, 10, int(invert), int(centered), POSTFIX])
    got = parse_dry(r.stdout)
    assert sorted(got) == sorted(q["name"] for q in queries[:4])
    assert "skipping image" in r.stdout and "Could not load matches" in r.stderr
    for q in queries[:4]:
        g = got[q["name"]]
        assert g["model"] == q["model"] and (g["w"], g["h"]) == (640, 320)
        assert g["fx"] == q["fx"] and g["fy"] == q["fy"]
        assert g["n"] == len(q["p2"])
        s2 = 0.0
        for a, b in q["p2"]:
            s2 += a + b
        s3 = 0.0
        for a, b, c in q["p3"]:
            s3 += a + b + c
        assert g["s2"] == s2 and g["s3"] == s3  # same values, same summation order
    assert not os.path.exists(tmp_path / "out.txt")


def test_dry_run_gc_dialect_and_errors(tmp_path):
    listfile, queries = write_dataset(tmp_path, "gc", num=3, models=["SIMPLE_RADIAL", "PINHOLE"], invert=True,
                                      centered=True)
    r = run_tool(["--gc", "--dry-run", listfile, tmp_path / "o.txt", POSTFIX])
    got = parse_dry(r.stdout)
    assert len(got) == 3
    assert got[queries[1]["name"]]["fy"] == queries[1]["fy"] != queries[1]["fx"]  # gc keeps PINHOLE fy
    # OPENCV is not a model of the gc tool; an unknown model is an error, not uninitialised focals
    with open(listfile, "a") as fh:
        fh.write("x.jpg FISHEYE 640 320 100 320 160\n")
    r = run_tool(["--gc", "--dry-run", listfile, tmp_path / "o.txt", POSTFIX], check=False)
    assert r.returncode != 0 and "unknown camera model" in r.stderr
    r = run_tool(["--dry-run", tmp_path / "nope.txt", tmp_path / "o.txt", 12, 10, 1, 0], check=False)
    assert r.returncode != 0 and "cannot read the image list" in r.stderr
    assert run_tool([listfile], check=False).returncode != 0  # too few arguments: usage only


def oracle_pose(q, seed, **kw):
    O = ol.oracle()
    o = ol.Options.default(seed=seed, **kw)
    st = ol.Stats()
    model = np.zeros(12)
    inl = np.zeros(len(q["p2"]), np.int32)
    # the example solver takes one focal per axis
    O.lomsac_pose(C.byref(o), float(q["fx"]), float(q["fy"]), len(q["p2"]), ol.c64(q["p2"]), ol.c64(q["p3"]), model,
                  C.byref(st), inl)
    return model, st


def read_poses(path):
    out = {}
    with open(path) as fh:
        for ln in fh:
            t = ln.split()
            out[t[0]] = np.array([float(v) for v in t[1:]])
    return out


@pytest.mark.gpu
@pytest.mark.parametrize("dialect", ["plain", "gt", "gc"])
def test_tool_poses_equal_the_oracle(tmp_path, dialect):
    opts = {"plain": dict(min_iters=200, max_iters=100000, lo_start=60, final_lsq=1, sq_thr=144.0, lo_steps=10),
            "gt": dict(min_iters=100, max_iters=10000, lo_start=60, final_lsq=1, sq_thr=144.0, lo_steps=5,
                       thr_mult=16.0),
            "gc": dict(min_iters=20, max_iters=10000, lo_start=20, final_lsq=0, sq_thr=25.0, lo_steps=10)}[dialect]
    models = ["SIMPLE_RADIAL", "PINHOLE"] if dialect == "gc" else MODELS
    centered = dialect == "gc"
    listfile, queries = write_dataset(tmp_path, dialect, num=8, n_in=200, n_out=200, invert=True, centered=centered,
                                      models=models)
    with open(queries[6]["name"] + POSTFIX) as fh:   # too few matches for every tool
        keep = fh.readlines()[:3]
    with open(queries[6]["name"] + POSTFIX, "w") as fh:
        fh.writelines(keep)
    out = tmp_path / "poses.txt"
    if dialect == "gc":
        args = ["--gc", "--seed", 100, listfile, out, POSTFIX]
    else:
        args = (["--gt"] if dialect == "gt" else []) + ["--seed", 100, listfile, out, 12.0, opts["lo_steps"], 1, 0,
                                                         POSTFIX]
    r = run_tool(args)
    poses = read_poses(out)
    assert sorted(poses) == sorted(q["name"] for i, q in enumerate(queries) if i != 6)
    pos_err, rot_err = [], []
    for i, q in enumerate(queries):
        if i == 6:
            continue
        model, st = oracle_pose(q, 100 + i, **opts)
        M = model.reshape(4, 3).T  # column-major [R | c]
        R, c = M[:, :3], M[:, 3]
        want = np.concatenate([R_to_quat(R), -R @ c])
        np.testing.assert_allclose(poses[q["name"]], want, rtol=2e-5, atol=2e-5)  # 6 significant digits in the file
        assert f"LOMSAC found {st.best_num_inliers} inliers in {st.num_iterations} iterations" in r.stdout
        assert st.best_num_inliers >= 150
        pos_err.append(np.linalg.norm(c - q["c_gt"]))
        cosang = np.clip(0.5 * (np.trace(R.T @ q["R_gt"]) - 1.0), -1, 1)
        rot_err.append(np.degrees(np.arccos(cosang)))
    if dialect == "gt":
        stats = open(str(out) + ".stats.txt").read()
        big = np.finfo(np.float64).max
        med_pos = np.median(pos_err + [big])  # the skipped query keeps DBL_MAX (localization_with_gt.cc:224-227)
        med_rot = np.median(rot_err + [big])
        assert f"Median position error: {med_pos:.6g}m" in stats
        assert f"Median orientation error: {med_rot:.6g} deg" in stats
        within = sum(1 for p, a in zip(pos_err, rot_err) if p <= 0.05 and a <= 5.0)
        assert f"% images within 5cm and 5deg: {within / len(queries) * 100.0:.6g}" in stats
        assert len(open(str(out) + ".times.txt").read().split()) == 7
