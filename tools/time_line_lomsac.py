"""Device-resident line LO-MSAC: latency of one BASELINE configs[0] run (one CTA) and throughput of a
batch of such runs (one problem per CTA, 2 CTAs/SM), next to the reference's EstimateModel on one host
core. Inputs resident in HBM, CUDA events on the context stream. Prints one JSON object.

    python tools/time_line_lomsac.py [problems=592] [n=10000] [outlier_ratio=0.5]
"""
import ctypes as C
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import oracle_lib as ol  # noqa: E402  (test infrastructure: the timed CPU baseline)
import ransaclib_b200 as rlb  # noqa: E402
from ransaclib_b200 import _capi, synth  # noqa: E402


def main():
    P = int(sys.argv[1]) if len(sys.argv) > 1 else 592
    n = int(sys.argv[2]) if len(sys.argv) > 2 else 10000
    ratio = float(sys.argv[3]) if len(sys.argv) > 3 else 0.5
    n_out = int(n * ratio)
    distinct = 16
    xs = [synth.line_instance(1000 + q, n - n_out, n_out) for q in range(distinct)]
    xy = np.concatenate([xs[q % distinct] for q in range(P)])
    off = (np.arange(P + 1, dtype=np.int64) * n)
    seeds = (42 + np.arange(P)).astype(np.uint32)

    ctx = rlb.Context(0)
    stream = torch.cuda.Stream()
    ctx.set_stream(stream.cuda_stream)
    opt = ctx.lomsac_options(sq_thr=1e-4, min_iters=100, max_iters=100000)
    dev = torch.device("cuda:0")
    d_xy = torch.from_numpy(xy).to(dev)
    d_off = torch.from_numpy(off).to(dev)
    d_seeds = torch.from_numpy(seeds.view(np.int32)).to(dev)  # same 32 bits
    d_lines = torch.zeros(P, 3, dtype=torch.float64, device=dev)
    d_res = torch.zeros(P * _capi.LOCALIZE_RESULT.itemsize, dtype=torch.uint8, device=dev)
    d_inl = torch.zeros(P * n, dtype=torch.int32, device=dev)
    torch.cuda.synchronize()

    def run(count, opt=opt):
        d_lines.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(stream):
            e0.record(stream)
            ctx.line_lomsac_batch_dev(count, n, d_off.data_ptr(), d_xy.data_ptr(), d_seeds.data_ptr(), opt,
                                      d_lines.data_ptr(), d_res.data_ptr(), d_inl.data_ptr())
            e1.record(stream)
        e1.synchronize()
        return e0.elapsed_time(e1)

    modes = {}
    for mode in ("1", "0"):  # points staged in shared memory (1 CTA/SM) / read from global memory (2 CTAs/SM)
        os.environ["RLB_LL_SMEM"] = mode
        for _ in range(3):
            run(1)
        one_m = min(run(1) for _ in range(5))
        for _ in range(2):
            run(P)
        many_m = min(run(P) for _ in range(3))
        modes["smem" if mode == "1" else "global"] = {"one_problem_ms": one_m, "batch_ms": many_m,
                                                      "problems_per_s": P / (many_m * 1e-3)}
    os.environ.pop("RLB_LL_SMEM")
    run(1)
    one = min(run(1) for _ in range(5))
    many = min(run(P) for _ in range(3))
    res = np.frombuffer(d_res.cpu().numpy().tobytes(), _capi.LOCALIZE_RESULT)

    # per-phase cycle counters of problem 0 running alone (debug_flags bit 1)
    run(1, ctx.lomsac_options(sq_thr=1e-4, min_iters=100, max_iters=100000, debug_flags=2))
    tail = d_inl[:n].cpu().numpy()
    at = (n - 32) & ~1
    prof = np.frombuffer(tail[at:at + 32].tobytes(), np.int64)
    names = ["sampler", "solve+sweep", "exact_passes", "shuffles", "fits", "whole", "seeding", "-",
             "pass_eval", "pass_reduce_barrier", "pass_list", "pass_closing_barrier"]
    phases = {k: int(v) for k, v in zip(names, prof)}

    # CPU: the reference's own EstimateModel (oracle/_ref, reference flags) on one core, same problems
    ref = ol.reference(fast=True) or ol.oracle()
    kind = "reference" if ol.reference(fast=True) is not None else "port"
    m = min(P, 16)
    t0 = time.perf_counter()
    for q in range(m):
        o = ol.Options.default(sq_thr=1e-4, seed=int(seeds[q]), min_iters=100, max_iters=100000)
        st, mod, inl = ol.Stats(), np.zeros(3), np.zeros(n, np.int32)
        ref.lomsac_line(C.byref(o), n, ol.c64(xs[q % distinct]), mod, C.byref(st), inl)
        assert st.num_iterations == res["num_iterations"][q] and st.best_num_inliers == res["num_inliers"][q]
    cpu_s = (time.perf_counter() - t0) / m

    print(json.dumps({
        "what": "device-resident line LO-MSAC (BASELINE configs[0] shape), inputs resident, CUDA events",
        "n_points": n, "outlier_ratio": ratio, "problems": P,
        "one_problem_ms": one, "batch_ms": many, "problems_per_s": P / (many * 1e-3),
        "mean_iterations": float(res["num_iterations"].mean()), "mean_lo": float(res["num_lo"].mean()),
        "mean_passes": float(res["passes"].mean()), "mean_fits": float(res["lsq_calls"].mean()),
        "cpu_kind": kind, "cpu_one_core_ms": cpu_s * 1e3, "cpu_problems_per_s_one_core": 1.0 / cpu_s,
        "modes": modes, "phase_cycles_problem0": phases, "sms": ctx.sm_count}))
    ctx.close()


if __name__ == "__main__":
    main()
