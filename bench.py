#!/usr/bin/env python
"""bench.py — MSAC hypothesis x point evaluations / s on the scoring-stress workload.

One "step" = one pass of the hot path over one batch of synthetic input:
  P3P on K seeded 4-samples  ->  K x N MSAC sweep (score + inlier count)  ->  argmin
  ->  [N>1: one NCCL all_gather of (score, id, model) per rank]  ->  ordered inlier list of
  the winning hypothesis.
Workload (BASELINE.json configs[4], "scoring stress"): N = 1,000,000 2D-3D correspondences
(50 % outliers, camera_pose_estimation.cc:52-127 distributions), K = 65,536 P3P hypotheses
drawn by the reference's seeded UniformSampling (seed 42), hypotheses sharded K/G per GPU
(points replicated), so scaling is STRONG. configs[1] (N = 10k x K = 10k) is 1e8 evaluations
= ~0.1 ms of GPU work, i.e. a launch-latency measurement; it is a parity-test case and the
``--workload c2`` option here.

Usage:  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                        [--workload c5|c2] [--precision f32|f64]
Under torchrun (N > 1) one process per GPU; rank 0 prints ONE JSON line.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

FLOP_PER_EVAL = 29.0  # pose evaluation incl. min + accumulate
# dram__bytes_read.sum + dram__bytes_write.sum of one sweep_pose_tc_kernel launch on the default C5 run
# (ncu --set full): 80.2 MB read (72 MB of operand tiles) + the [splits][K] partials written: 18.4 MB scores only
# (profiles/history/r01s_ncu_sweep_tc.csv), 28.0 MB with counts (profiles/r01r_ncu_sweep_tc.csv)
TC_TRAFFIC = {False: 98.6e6, True: 108.3e6}
THR = 144.0           # 12 px squared, camera_pose_estimation.cc:154-156
WORKLOADS = {
    "c5": dict(n=1_000_000, k=65_536, name="scoring stress N=1M x K=65536 P3P hypotheses (configs[4])"),
    "c2": dict(n=10_000, k=10_000, name="camera_pose_estimation N=10k x K=10k (configs[1])"),
}


_REAL_STDOUT = None


def claim_stdout():
    """The contract is ONE JSON line on stdout. Libraries print there too (NCCL's version banner, for one), so
    fd 1 is pointed at stderr for the duration of the run and the line goes to the saved descriptor."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def emit(text):
    sys.stdout.flush()
    if _REAL_STDOUT is None:
        print(text, flush=True)
    else:
        os.write(_REAL_STDOUT, (text + "\n").encode())


def make_instance(n, data_seed=9000):
    from ransaclib_b200 import synth

    inst = synth.pose_instance(data_seed, n // 2, n - n // 2)
    return inst["points2D"], inst["points3D"], inst["focal"], inst


def make_samples(n, k, seed=42):
    from ransaclib_b200.sampling import UniformSampling

    return UniformSampling(seed, n, 4).sample_fast(k)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                 "-i", str(self.gpu)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._pump, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for ln in self.proc.stdout:
            self.lines.append((time.time(), ln.strip()))

    def stop(self, t0, t1):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        for ts, ln in self.lines:
            if ts < t0 or ts > t1 + 0.2:
                continue
            p = [x.strip() for x in ln.split(",")]
            if len(p) < 9:
                continue
            try:
                sm.append(float(p[1]))
                mx.append(float(p[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"),
                               p[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as fh:
            return json.load(fh), "measured"
    except Exception:
        return {"hbm_gbs": 6650.0}, "fallback"


# ------------------------------------------------------------------------ our arm
def run_ours(args):
    import torch

    import ransaclib_b200 as rlb
    from ransaclib_b200 import sharding

    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist = None
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=dev)
    wl = WORKLOADS[args.workload]
    n, k_total = wl["n"], (args.k_override or wl["k"])
    prec = rlb.RLB_F32 if args.precision == "f32" else rlb.RLB_F64
    id_lo, id_hi = sharding.hypothesis_range(k_total, rank, world)
    k = id_hi - id_lo

    p2, p3, f, inst = make_instance(n)
    samples_all = make_samples(n, k_total)
    my_samples = np.ascontiguousarray(samples_all[id_lo:id_hi])  # contiguous id ranges per rank

    ctx = rlb.Context(local)
    stream = torch.cuda.Stream(dev)  # one explicit stream for our kernels and the library's NCCL calls
    torch.cuda.set_stream(stream)
    ctx.set_stream(stream.cuda_stream)
    ctx.set_timing(True)
    # The data plane's collectives live in the library (rlb_comm_*: ncclBroadcast of the points, one
    # ncclAllGather per round); torch.distributed only carries the unique id, barriers and the timing max.
    uid = torch.zeros(128, dtype=torch.uint8, device=dev)
    if rank == 0 and world > 1:
        uid.copy_(torch.frombuffer(bytearray(rlb.Comm.unique_id()), dtype=torch.uint8))
    if world > 1:
        dist.broadcast(uid, 0)
    comm = rlb.Comm.from_rank(ctx, world, rank, bytes(uid.cpu().numpy().tobytes()) if world > 1 else None)
    # points resident in HBM before the timed region: one PCIe upload on rank 0, NVLink broadcast to the others
    cscorer = comm.pose_scorer(p2 if rank == 0 else n, p3 if rank == 0 else None, f, f, root=0)
    scorer = rlb.Scorer(ctx, cscorer.local_scorer_handle(0), borrowed=True)

    d_samples = torch.from_numpy(my_samples).to(dev)
    d_models = torch.empty((k, 12), dtype=torch.float64, device=dev)
    d_valid = torch.empty(k, dtype=torch.uint8, device=dev)
    d_scores = torch.empty(k, dtype=torch.float64, device=dev)
    d_counts = torch.empty(k, dtype=torch.int32, device=dev)
    d_best = torch.zeros(2, dtype=torch.float64, device=dev)  # {double score; int64 id}
    d_idx = torch.empty(n, dtype=torch.int32, device=dev)
    d_hdr = torch.zeros(4, dtype=torch.float64, device=dev)
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)  # > 126 MB L2

    d_winner = torch.zeros(14, dtype=torch.float64, device=dev)  # {score, global id, model[12]}

    def step(p=None):
        # rlb_comm_round_dev: solve + sweep + argmin of this rank's shard, ONE ncclAllGather of the ranks'
        # (score, global id, model) records, lexicographic winner on every rank; then the winner's inliers
        cscorer.round_dev(d_samples.data_ptr(), k, THR, id_lo, d_models.data_ptr(), d_valid.data_ptr(),
                          d_scores.data_ptr(), d_counts.data_ptr() if args.counts else 0, d_winner.data_ptr(),
                          precision=prec if p is None else p)
        scorer.get_inliers_dev(d_winner.data_ptr() + 16, THR, THR, None, d_idx.data_ptr(), d_hdr.data_ptr())

    for _ in range(args.warmup):
        step()
    torch.cuda.synchronize()

    # ---- timed region: device-resident inputs ------------------------------------
    launches0 = ctx.launch_count
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    clocks = ClockSampler(local)
    clocks.start()
    time.sleep(0.25)
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    sweep_ms, inl_ms = [], []
    wall0 = time.time()
    for s0, s1 in evs:
        flush.zero_()  # L2 flush between timed iterations (untimed)
        s0.record(stream)
        step()
        s1.record(stream)
        s1.synchronize()
        sweep_ms.append(ctx.kernel_time_ms(0))
        inl_ms.append(ctx.kernel_time_ms(1))
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    wall1 = time.time()
    launches = ctx.launch_count - launches0
    clk = clocks.stop(wall0, wall1)
    total_ms = sum(a.elapsed_time(b) for a, b in evs)
    t = torch.tensor([total_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)  # max over ranks
    total_ms = float(t.item())
    value = k_total * n * args.steps / (total_ms * 1e-3)

    hdr = d_hdr.cpu().numpy()
    n_inl = int(hdr.view(np.int32)[3])
    best_score = float(d_hdr[0].item())

    # ---- parity (once, untimed): f32 against f64 on ALL hypotheses, a sample against the reference's own
    #      loops, and (N > 1) the gathered winner against a single-GPU recomputation ---------------------
    parity = None
    if not args.profile:
        parity = measure_parity(args, rank, world, dev, ctx, scorer, step, d_samples, d_models, d_valid,
                                d_scores, d_counts, d_winner, samples_all, k_total, p2, p3, f, dist)

    # ---- e2e: the same step through the host-buffer C ABI, copies inside the timed region ----
    pin = lambda a: torch.from_numpy(a).pin_memory().numpy()  # noqa: E731
    h_p2, h_p3 = pin(p2), pin(p3)
    e2e_steps = max(1, min(args.steps, 5))

    h_all = pin(np.ascontiguousarray(samples_all))
    e2e_parts = []

    def e2e_step():
        # host buffers in, results out, everything inside: every rank holds the step's inputs in pinned host
        # memory and uploads 1/G of the points over its own PCIe link (RLB_COMM_ROOT_ALL), the slices are
        # exchanged over NVLink; every rank uploads its shard of the samples; one ncclAllGather; the winner's
        # ordered inlier list comes back from the first GPU of every process
        t0 = time.perf_counter()
        cs = comm.pose_scorer(h_p2, h_p3, f, f, root=-1)
        t1 = time.perf_counter()
        r = cs.hypothesize_and_score(h_all, THR, precision=prec, want_scores=True)
        t2 = time.perf_counter()
        inl, score = cs.get_inliers(r["best_model"], THR, THR)
        cs.close()
        t3 = time.perf_counter()
        e2e_parts.append((t1 - t0, t2 - t1, t3 - t2))
        return r, [len(inl)]

    if args.profile:
        if rank == 0:
            emit(json.dumps({"profile_run": True, "value": value, "ms_per_step": total_ms / args.steps,
                              "sweep_ms": float(np.mean(sweep_ms)), "inlier_ms": float(np.mean(inl_ms)),
                              "gpu_launches": int(launches)}))
        if world > 1:
            dist.destroy_process_group()
        return
    e2e_step()
    e2e_parts.clear()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        r_e2e, c_e2e = e2e_step()
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    te = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_value = k_total * n * e2e_steps / float(te.item())
    h2d = (p2.nbytes + p3.nbytes) // world + my_samples.nbytes + 12 * 8
    d2h = k * 8 + k * 4 + 16 + 96 + 16 + 4 * int(c_e2e[0])

    # ---- second half of BASELINE.json's metric: batched localization (configs[3]) on the same GPUs ----
    loc = None
    if args.workload == "c5" and not args.no_localization:
        loc = measure_c4(args, rank, world, local, dist, steps=max(1, min(args.steps, 5)),
                         warmup=max(1, min(args.warmup, 3)))
        torch.cuda.set_stream(stream)
        ctx.set_stream(stream.cuda_stream)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel (the K x N sweep) ------------------------
    peaks, peak_kind = measured_peaks()
    fma_dtype = 1 if prec == rlb.RLB_F32 else 0
    fma_peak = ctx.measure_fma_peak(fma_dtype)
    ffma2_peak = ctx.measure_fma_peak(2) if prec == rlb.RLB_F32 else None
    sweep_avg_ms = float(np.mean(sweep_ms))
    evals_per_launch = k * n
    achieved_tflops = evals_per_launch * FLOP_PER_EVAL / (sweep_avg_ms * 1e-3) / 1e12
    bpp = 20 if prec == rlb.RLB_F32 else 40
    if prec == rlb.RLB_F32 and ctx.last_sweep_plan()[0] == 9:
        bpp = 72  # tensor-core operand tiles: 16 + 2 floats per point
    algo_bytes = n * bpp + k * 96 + k * 12
    inl_avg_ms = float(np.mean(inl_ms))
    inl_bytes = n * 40 + 4 * n_inl
    variant, splits = ctx.last_sweep_plan()
    ffma_names = {0: "sweep_pose_f32x2_kernel<4>", 2: "sweep_pose_f32x2_kernel<2>", 3: "sweep_pose_f32x2_kernel<8>",
                  7: "sweep_kernel<PoseF,double,4>", -1: "sweep_kernel<PoseF,double,2>"}
    tc = (variant == 9)
    kernel_name = ("sweep_pose_tc_kernel<%s>" % ("true" if args.counts else "false")) if tc \
        else ffma_names.get(variant, "sweep variant %d" % variant)
    tensor = None
    if tc:
        # the projection of one evaluation is 3 rows x K = 16 TF32 MACs on the tensor core (hi/lo split)
        tf32 = evals_per_launch * 96.0 / (sweep_avg_ms * 1e-3) / 1e12
        tf32_peak = peaks.get("bf16_tflops", 0.0) / 2.0 if peaks.get("bf16_tflops") else None
        tensor = {"achieved_tf32_tflops": tf32, "peak_tf32_tflops": tf32_peak,
                  "frac": (tf32 / tf32_peak) if tf32_peak else None,
                  "note": "tcgen05.mma kind::tf32 M128 N32 K8; peak = MEASURED_PEAKS bf16 / 2. The tensor pipe is "
                          "not the bound: the CUDA-core epilogue (rsqrt, residual, MSAC test, sums; ~9 issue "
                          "slots per evaluation-warp) and the TMEM read-back are (DESIGN.md section 3c)"}
    roofline = {
        "bound": "fma", "kernel": kernel_name, "sweep_variant": variant, "splits": splits, "tensor": tensor,
        "achieved": achieved_tflops, "peak": fma_peak, "unit": "TFLOP/s", "frac": achieved_tflops / fma_peak,
        "peak_source": "rlb_measure_fma_peak: register-resident %s microbenchmark on this GPU, same run "
                       "(MEASURED_PEAKS.json has no FMA-pipe figure)" % ("FFMA" if fma_dtype else "DFMA"),
        "ffma2_peak_tflops": ffma2_peak,
        "frac_note": ("29 algorithmic flop per evaluation against the FFMA-pipe peak; with the tensor-core sweep "
                      "18 of them (the 3x4 projection) run on the tensor pipe, so the fraction is work-equivalent, "
                      "not FMA-pipe occupancy") if tc else None,
        "flop_per_eval": FLOP_PER_EVAL, "evals_per_launch": evals_per_launch, "launch_ms": sweep_avg_ms,
        "evals_per_s": evals_per_launch / (sweep_avg_ms * 1e-3),
        "share_of_step": sweep_avg_ms * args.steps / total_ms,
        # dram__bytes_read.sum + dram__bytes_write.sum of this kernel, one launch, from the ncu --set full
        # capture of the same command (profiles/history/r01b_ncu_sweep_h8.csv); only valid for the default f32 C5 run.
        "traffic": (TC_TRAFFIC[bool(args.counts)] if tc else 33.1e6) if (prec == rlb.RLB_F32 and args.workload == "c5" and world == 1)
        else None,
        "hbm": {"algorithmic_bytes": algo_bytes, "achieved_gbs": algo_bytes / (sweep_avg_ms * 1e-3) / 1e9,
                "peak_gbs": peaks.get("hbm_gbs"), "peak_kind": peak_kind,
                "note": "arithmetic intensity ~%.0f FLOP/B: HBM is not the bound of this kernel"
                        % (evals_per_launch * FLOP_PER_EVAL / algo_bytes)},
        "inlier_pass": {"bound": "hbm", "launch_ms": inl_avg_ms, "algorithmic_bytes": inl_bytes,
                        "achieved": inl_bytes / (inl_avg_ms * 1e-3) / 1e9, "peak": peaks.get("hbm_gbs"),
                        "unit": "GB/s", "frac": inl_bytes / (inl_avg_ms * 1e-3) / 1e9 / peaks.get("hbm_gbs", 6650.0),
                        "kernel": "inlier_range_kernel<PoseF>",
                        # dram__bytes_read + write of one launch at N = 1M (profiles/r02z_ncu_inlier_range.csv): the
                        # 2 MB inlier list stays in L2, so the traffic is the 40 MB of f64 point channels
                        "traffic": 40.07e6 if (args.workload == "c5" and n == 1_000_000) else None},
    }

    # ---- the inlier pass where its fixed costs are amortised: the same points eight times over ----
    if args.workload == "c5" and world == 1 and not args.no_extra:
        roofline["inlier_pass"]["large_n"] = measure_inlier_large_n(ctx, flush, p2, p3, f, d_winner, peaks)

    # ---- the reference-exact f64 sweep of the same workload (own leg, same run) ------
    f64_leg = None
    if prec == rlb.RLB_F32 and args.workload == "c5" and world == 1 and not args.no_extra:
        f64_leg = measure_f64_leg(args, ctx, step, stream, flush, k, n, peaks)
    other_sweeps = measure_line_sweeps(ctx, peaks) if (world == 1 and not args.no_extra) else None
    configs_table = run_configs_table() if (world == 1 and not args.no_extra) else None

    # ---- cpu baseline: the reference's loops on ONE host core, bounded sample -------
    cpu = cpu_baseline(p2, p3, f, d_models[:64].cpu().numpy(), d_valid[:64].cpu().numpy(), threads=1,
                       budget_s=12.0) if world == 1 else None

    line = {
        "metric": "MSAC hypothesis x point evals/sec", "value": value, "unit": "evals/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": total_ms / args.steps,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f32" if prec == rlb.RLB_F32 else "f64", "data": "synthetic",
        # the second half of BASELINE.json's metric up front (details under "localization")
        "localizations_per_s": loc["value"] if loc else None,
        "localizations_per_s_e2e": loc["e2e"]["value"] if loc else None,
        "e2e_evals_per_s": e2e_value,
        "config": {"workload": wl["name"], "n_points": n, "k_hypotheses": k_total, "k_per_gpu": k,
                   "per_hypothesis_counts": bool(args.counts),
                   "sq_thr": THR, "partitioning": "hypothesis-sharded, points replicated, 1 all_gather/step"
                   if world > 1 else "single GPU", "l2": "flushed (256 MiB write) between timed steps",
                   "step": "P3P solve + KxN MSAC sweep + argmin + ordered inlier list of the winner"},
        "e2e": {"value": e2e_value, "unit": "evals/s", "h2d_bytes_per_step": int(h2d),
                "d2h_bytes_per_step": int(d2h), "steps": e2e_steps,
                "breakdown_ms": {
                    "create_scorer (upload + NVLink exchange + repack + rays)": 1e3 * float(np.mean([q[0] for q in e2e_parts])),
                    "hypothesize_and_score (samples up, P3P, tile pack, sweep, all_gather, scores back)":
                        1e3 * float(np.mean([q[1] for q in e2e_parts])),
                    "get_inliers + destroy": 1e3 * float(np.mean([q[2] for q in e2e_parts]))},
                "path": "rlb_comm_scorer_create_pose (every GPU uploads 1/G of the points over its own PCIe link, "
                        "slices exchanged over NVLink) + "
                        "rlb_comm_hypothesize_and_score (shard upload, sweep, one ncclAllGather, scores + counts "
                        "back) + rlb_comm_get_inliers, pinned host buffers; bytes are rank 0's"},
        "gpu_launches": int(launches), "clocks": clk, "roofline": roofline, "cpu_baseline": cpu,
        "result": {"num_inliers": n_inl, "best_score": best_score}, "parity": parity,
        "f64": f64_leg, "other_sweeps": other_sweeps, "configs_table": configs_table,
        "comm": {"nranks": comm.nranks, "collectives_issued": comm.collective_count,
                 "note": "rlb_comm_*: NCCL inside the library (ncclBroadcast of the points, one ncclAllGather per "
                         "round); torch.distributed carries only the unique id, barriers and the timing max"},
    }
    if loc is not None:
        if world == 1:
            loc["cpu_baseline"] = cpu_baseline_c4(budget_queries=64, threads=1)
        line["localization"] = loc  # "P3P LO-MSAC localizations/sec" at this N (same run, same GPUs)
    emit(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def measure_inlier_large_n(ctx, flush, p2, p3, f, d_winner, peaks, reps=8, steps=5):
    """One K = 1 inlier pass over reps x the workload's points (8M for C5) with the winner of the timed step: the
    ~6 us of HBM time of the 1M-point pass is of the order of the launch, ticket and chain latencies; here those
    are amortised and the fraction of the HBM peak is the kernel's streaming rate. L2 flushed before every launch."""
    import torch

    dev = d_winner.device
    sc = ctx.pose_scorer(np.tile(p2, (reps, 1)), np.tile(p3, (reps, 1)), f, f)
    n = reps * len(p2)
    d_idx = torch.empty(n, dtype=torch.int32, device=dev)
    d_hdr = torch.zeros(4, dtype=torch.float64, device=dev)
    ms = []
    for it in range(steps + 2):
        flush.zero_()
        sc.get_inliers_dev(d_winner.data_ptr() + 16, THR, THR, None, d_idx.data_ptr(), d_hdr.data_ptr())
        torch.cuda.synchronize()
        if it >= 2:
            ms.append(ctx.kernel_time_ms(1))
    n_inl = int(d_hdr.cpu().numpy().view(np.int32)[3])
    del sc
    avg = float(np.mean(ms))
    nbytes = n * 40 + 4 * n_inl
    gbs = nbytes / (avg * 1e-3) / 1e9
    return {"n_points": n, "num_inliers": n_inl, "launch_ms": avg, "algorithmic_bytes": nbytes, "achieved": gbs,
            "unit": "GB/s", "frac": gbs / peaks.get("hbm_gbs", 6650.0), "kernel": "inlier_range_kernel<PoseF>"}


def measure_f64_leg(args, ctx, step, stream, flush, k, n, peaks):
    """The same step with f64 results: the filtered sweep (RLB_F64: f32 classification on the tensor cores, f64
    for the pairs that could be inliers) and the plain f64 sweep (RLB_F64_PLAIN) beside it; CUDA events, L2
    flushed, 3 timed steps each."""
    import torch

    import ransaclib_b200 as rlb

    def timed(prec):
        step(prec)
        torch.cuda.synchronize()
        steps = 3
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        sweep_ms = []
        for s0, s1 in evs:
            flush.zero_()
            s0.record(stream)
            step(prec)
            s1.record(stream)
            s1.synchronize()
            sweep_ms.append(ctx.kernel_time_ms(0))
        total_ms = sum(a.elapsed_time(b) for a, b in evs)
        variant, splits = ctx.last_sweep_plan()
        return steps, total_ms, float(np.mean(sweep_ms)), variant, splits

    steps, total_ms, sw, variant, splits = timed(rlb.RLB_F64)
    p_steps, p_total, p_sw, p_variant, p_splits = timed(rlb.RLB_F64_PLAIN)
    dfma = ctx.measure_fma_peak(0)
    ffma = ctx.measure_fma_peak(1)
    ach = k * n * FLOP_PER_EVAL / (sw * 1e-3) / 1e12
    p_ach = k * n * FLOP_PER_EVAL / (p_sw * 1e-3) / 1e12
    return {"metric": "MSAC hypothesis x point evals/sec", "dtype": "f64", "value": k * n * steps / (total_ms * 1e-3),
            "unit": "evals/s", "steps": steps, "ms_per_step": total_ms / steps,
            "roofline": {"bound": "fma",
                         "kernel": "sweep_pose_tcx_kernel<false> (+ prepass, partition, sweep_kernel<PoseF,double,4> "
                                   "over the dense list)" if variant == 11 else "sweep_kernel<PoseF,double,*>",
                         "sweep_variant": variant, "splits": splits, "achieved": ach, "peak": ffma,
                         "unit": "TFLOP/s", "frac": ach / ffma, "launch_ms": sw,
                         "peak_source": "rlb_measure_fma_peak: register-resident FFMA microbenchmark, same run",
                         "frac_note": "f64 RESULTS (counts identical to the plain f64 sweep, scores to 1e-12, both "
                                      "within 1e-9 of the reference: parity.oracle / parity.filtered_vs_plain), but the "
                                      "arithmetic is mostly f32: every pair is classified by the tensor-core front end, "
                                      "pairs that could be inliers are re-evaluated in f64, hypotheses with many such "
                                      "pairs go to the plain f64 sweep. Same 29-flop-equivalent / FFMA-peak convention "
                                      "as the f32 line; launch_ms covers all four launches of the sweep",
                         # dram__bytes of the sparse-list launch at K = 16,384 (profiles/r02z_ncu_tcx.csv): 72 B / point
                         # of f32 tiles + 40 B / point of f64 channels + the [splits][K] partials
                         "traffic_at_k16384": 121.7e6,
                         "hbm": {"algorithmic_bytes": n * 112 + k * 108,
                                 "achieved_gbs": (n * 112 + k * 108) / (sw * 1e-3) / 1e9,
                                 "peak_gbs": peaks.get("hbm_gbs")}},
            "plain": {"value": k * n * p_steps / (p_total * 1e-3), "unit": "evals/s", "ms_per_step": p_total / p_steps,
                      "roofline": {"bound": "fma", "kernel": "sweep_kernel<PoseF,double,%d>" % (4 if p_variant == 7 else 2),
                                   "sweep_variant": p_variant, "splits": p_splits, "achieved": p_ach, "peak": dfma,
                                   "unit": "TFLOP/s", "frac": p_ach / dfma, "launch_ms": p_sw,
                                   "peak_source": "rlb_measure_fma_peak: register-resident DFMA microbenchmark, same run",
                                   "fp64_pipe_instructions_per_eval": 20,
                                   "note": "every pair in f64 (RLB_F64_PLAIN): 29 algorithmic flop per evaluation = 14.5 "
                                           "FMA slots against 20 FP64-pipe instructions issued, so the pipe bounds it at "
                                           "0.725 of the DFMA peak in these units; ncu: FP64 pipe 72 % active, top stall "
                                           "math-pipe throttle (profiles/r02_f64_sweep_ab.txt)"}}}


def measure_line_sweeps(ctx, peaks):
    """K x N sweeps of the line (8 flop / evaluation) and hybrid-line (9) functors, SURVEY 8(d): N = 1M points
    (hybrid: 2 x 512k), K = 16,384 two-point hypotheses, f32 and f64, kernel time by CUDA events."""
    import ransaclib_b200 as rlb
    from ransaclib_b200 import synth
    from ransaclib_b200.sampling import UniformSampling

    out = {}
    n, k = 1_000_000, 16_384
    xy = synth.line_instance(77, n // 2, n - n // 2)
    samples = UniformSampling(5, n, 2).sample_fast(k)
    sc = ctx.line_scorer(xy)
    ffma, dfma = ctx.measure_fma_peak(1), ctx.measure_fma_peak(0)
    for name, prec, flop, peak in (("line_f32", rlb.RLB_F32, 8.0, ffma), ("line_f64", rlb.RLB_F64, 8.0, dfma)):
        for _ in range(3):
            sc.hypothesize_and_score(samples, 1e-4, precision=prec, want_models=False, want_scores=False)
        ms = ctx.kernel_time_ms(0)
        out[name] = {"evals_per_s": k * n / (ms * 1e-3), "launch_ms": ms, "flop_per_eval": flop,
                     "achieved_tflops": k * n * flop / (ms * 1e-3) / 1e12, "peak_tflops": peak,
                     "frac": k * n * flop / (ms * 1e-3) / 1e12 / peak, "sweep_plan": list(ctx.last_sweep_plan()),
                     "hbm_gbs": (n * (8 if prec == rlb.RLB_F32 else 16) + k * 36) / (ms * 1e-3) / 1e9}
    sc.close()
    h = n // 2
    xy0, p1 = synth.hybrid_line_instance(78, h // 2, h - h // 2, h // 2, h - h // 2)
    sc = ctx.hybrid_line_scorer(xy0, p1)
    s2 = UniformSampling(6, h, 2).sample_fast(k)
    for name, prec, peak in (("hybrid_line_f32", rlb.RLB_F32, ffma), ("hybrid_line_f64", rlb.RLB_F64, dfma)):
        for _ in range(3):
            sc.hypothesize_and_score(s2, [1e-4, 1e-4], weights=[2.0, 0.5], precision=prec, want_models=False,
                                     want_scores=False)
        ms = ctx.kernel_time_ms(0)
        out[name] = {"evals_per_s": k * n / (ms * 1e-3), "launch_ms": ms, "flop_per_eval": 9.0,
                     "achieved_tflops": k * n * 9.0 / (ms * 1e-3) / 1e12, "peak_tflops": peak,
                     "frac": k * n * 9.0 / (ms * 1e-3) / 1e12 / peak, "sweep_plan": list(ctx.last_sweep_plan())}
    sc.close()
    out["note"] = ("line: sweep_kernel<LineF,T,4>, 8 flop / evaluation of which 2 are multiply-adds of the residual: the "
                   "kernel is bound by issue slots (predicated add + count per evaluation), not by the FMA pipe")
    return out


def run_configs_table():
    """End-to-end EstimateModel wall time of BASELINE configs[0..2] through the drop-in layer next to the
    reference's EstimateModel on one host core (tests/e2e_configs.py in a subprocess, bounded)."""
    try:
        r = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "e2e_configs.py")], capture_output=True,
                           text=True, timeout=240, cwd=ROOT)
        if r.returncode != 0:
            return {"error": r.stderr[-400:]}
        tab = json.loads(r.stdout)
        rows = []
        for row in tab["rows"]:
            gpu = row.get("gpu_device_resident_s", min(row["gpu_f64_s"], row["gpu_f32_s"]))
            rows.append({"config": row["config"], "iterations": row["iterations"], "inliers": row["inliers"],
                         "reference_1core_ms": row["cpu_s"] * 1e3, "gpu_ms": gpu * 1e3,
                         "gpu_f64_ms": row["gpu_f64_s"] * 1e3, "gpu_f32_ms": row["gpu_f32_s"] * 1e3,
                         "speedup": row["cpu_s"] / gpu})
        return {"what": tab["what"], "cpu_kind": tab["cpu_kind"], "rows": rows}
    except Exception as e:  # noqa: BLE001
        return {"error": repr(e)[:400]}


def measure_parity(args, rank, world, dev, ctx, scorer, step, d_samples, d_models, d_valid, d_scores, d_counts,
                   d_winner, samples_all, k_total, p2, p3, f, dist):
    """Full-size parity of the timed workload, outside the timed region. Returns the dict on rank 0."""
    import torch

    import ransaclib_b200 as rlb

    k = d_samples.shape[0]
    if not args.counts:
        return {"skipped": "--no-counts run"}
    step(rlb.RLB_F32)
    torch.cuda.synchronize()
    w32 = d_winner.cpu().numpy().copy()
    s32, c32 = d_scores.clone(), d_counts.clone()
    step(rlb.RLB_F64)
    torch.cuda.synchronize()
    w64 = d_winner.cpu().numpy().copy()
    s64, c64 = d_scores.clone(), d_counts.clone()
    step(rlb.RLB_F64_PLAIN)
    torch.cuda.synchronize()
    wpl = d_winner.cpu().numpy().copy()
    spl, cpl = d_scores.clone(), d_counts.clone()
    ok = (d_valid == 1) & torch.isfinite(s64)
    rel = ((s32 - s64).abs() / s64.abs())[ok]
    relp = ((spl - s64).abs() / s64.abs())[ok]
    fvp = torch.tensor([float(relp.max().item()) if relp.numel() else 0.0], dtype=torch.float64, device=dev)
    cmis = torch.tensor([float((cpl != c64)[ok].sum().item())], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(fvp, op=dist.ReduceOp.MAX)
        dist.all_reduce(cmis, op=dist.ReduceOp.SUM)
    filtered_vs_plain = {"max_rel_score_diff": float(fvp.item()), "tolerance": 1e-12,
                         "count_mismatches": int(cmis.item()), "same_best_id": bool(wpl[1] == w64[1]),
                         "note": "RLB_F64 (filtered) against RLB_F64_PLAIN (every pair in f64) on all valid "
                                 "hypotheses of the timed workload: counts must be identical"}
    assert filtered_vs_plain["count_mismatches"] == 0 and filtered_vs_plain["max_rel_score_diff"] <= 1e-12 and \
        filtered_vs_plain["same_best_id"], filtered_vs_plain
    stats = torch.tensor([float(rel.max().item()) if rel.numel() else 0.0, float((c32 != c64)[ok].sum().item()),
                          float(ok.sum().item())], dtype=torch.float64, device=dev)
    if world > 1:
        mx = stats[:1].clone()
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        sm = stats[1:].clone()
        dist.all_reduce(sm, op=dist.ReduceOp.SUM)
        stats = torch.cat([mx, sm])
    out = {"f32_vs_f64": {"hypotheses": int(stats[2].item()), "same_best_id": bool(w32[1] == w64[1]),
                          "best_id": int(w64[1]), "max_rel_score_diff": float(stats[0].item()),
                          "tolerance": 1e-4, "count_mismatches": int(stats[1].item()),
                          "note": "all valid hypotheses of the timed workload; counts may differ only for points "
                                  "whose f32 error is within rounding of the threshold"}}
    out["filtered_vs_plain"] = filtered_vs_plain
    assert out["f32_vs_f64"]["same_best_id"], "f32 and f64 sweeps select different hypotheses"
    assert out["f32_vs_f64"]["max_rel_score_diff"] <= 1e-4, out
    if rank != 0:
        return None
    # the reference's own ScoreModel / GetInliers loops on a random sample of this rank's hypotheses
    ol, lib, kind = _load_ref(strict=True)
    rng = np.random.default_rng(7)
    valid_ids = torch.nonzero(ok).flatten().cpu().numpy()
    ids = np.sort(rng.choice(valid_ids, size=min(256, len(valid_ids)), replace=False))
    ids = np.unique(np.concatenate([ids, [int(w64[1]) - 0] if 0 <= int(w64[1]) < k else []]).astype(np.int64))
    models = d_models[torch.from_numpy(ids).to(dev)].cpu().numpy()
    t0 = time.perf_counter()
    ws, wc = _ref_scores(ol, lib, f, p2, p3, models, THR)
    dt = time.perf_counter() - t0
    g64 = s64[torch.from_numpy(ids).to(dev)].cpu().numpy()
    gc = c64[torch.from_numpy(ids).to(dev)].cpu().numpy()
    rel_o = float(np.max(np.abs(g64 - ws) / np.abs(ws)))
    out["oracle"] = {"kind": kind, "hypotheses": int(len(ids)), "max_rel_score_diff": rel_o, "tolerance": 1e-9,
                     "counts_exact": bool((gc == wc).all()), "seconds": dt,
                     "note": "f64 sweep vs the reference's ransac.h ScoreModel / GetInliers loops (oracle/_ref, "
                             "strict IEEE build) over all N points"}
    assert rel_o <= 1e-9 and out["oracle"]["counts_exact"], out["oracle"]
    if world > 1:
        # one GPU sweeps ALL k_total hypotheses: the N-GPU winner must be that winner
        d_all = torch.from_numpy(np.ascontiguousarray(samples_all)).to(dev)
        m_all = torch.empty((k_total, 12), dtype=torch.float64, device=dev)
        v_all = torch.empty(k_total, dtype=torch.uint8, device=dev)
        rec = torch.zeros(14, dtype=torch.float64, device=dev)
        scorer.sharded_round_dev(d_all.data_ptr(), k_total, THR, None, 0, m_all.data_ptr(), v_all.data_ptr(), 0, 0,
                                 rec.data_ptr(), precision=rlb.RLB_F64)
        torch.cuda.synchronize()
        r1 = rec.cpu().numpy()
        same = bool(r1[1] == w64[1] and abs(r1[0] - w64[0]) <= 1e-12 * abs(r1[0]) and (r1[2:] == w64[2:]).all())
        out["multi_gpu"] = {"winner_equals_single_gpu": same, "single_gpu_best_id": int(r1[1]),
                            "n_gpu_best_id": int(w64[1])}
        assert same, out["multi_gpu"]
    return out


def _ref_scores(ol, lib, f, p2, p3, models, thr, threads=0):
    """ScoreModel + inlier count of each model by the reference's loops, `threads` models at a time."""
    from concurrent.futures import ThreadPoolExecutor

    threads = threads or host_threads()
    chunks = [c for c in np.array_split(np.arange(len(models)), max(1, min(threads, len(models)))) if len(c)]

    def work(c):
        return ol.pose_score(lib, f, f, p2, p3, models[c], thr)

    with ThreadPoolExecutor(len(chunks)) as ex:
        res = list(ex.map(work, chunks))
    return np.concatenate([r[0] for r in res]), np.concatenate([r[1] for r in res])


# ------------------------------------------------------------------- CPU baselines
def _load_ref(strict=False):
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as ol

    lib = ol.reference(fast=not strict)
    kind = "reference"
    if lib is None:
        lib, kind = ol.oracle(), "port"
    return ol, lib, kind


def cpu_baseline(p2, p3, f, models, valid, threads, budget_s):
    """Times the reference's own ScoreModel loop (RansacLib/ransac.h:295-303, compiled from
    /root/reference into oracle/_ref with the reference's flags) on `threads` host threads over a
    bounded sample: all N points x as many hypotheses as fit the budget."""
    from concurrent.futures import ThreadPoolExecutor

    ol, lib, kind = _load_ref()
    models = np.ascontiguousarray(models[valid == 1] if (valid == 1).any() else models)
    n = len(p2)
    p2c, p3c = ol.c64(p2), ol.c64(p3)
    t_probe = lib.time_pose_scoring(f, f, n, p2c, p3c, 1, models[:1], THR, 1, None)
    k_each = int(max(1, min(4096, budget_s / max(t_probe, 1e-6))))
    reps = (k_each + len(models) - 1) // len(models)
    ms = np.ascontiguousarray(np.tile(models, (reps, 1))[:k_each])

    def work(_):
        return lib.time_pose_scoring(f, f, n, p2c, p3c, k_each, ms, THR, 1, None)

    t0 = time.perf_counter()
    if threads == 1:
        work(0)
    else:
        with ThreadPoolExecutor(threads) as ex:
            list(ex.map(work, range(threads)))
    dt = time.perf_counter() - t0
    return {"value": threads * k_each * n / dt, "unit": "evals/s", "cores": threads, "kind": kind,
            "sample": f"ScoreModel over all {n} points x {k_each} hypotheses per thread, "
                      f"{threads} thread(s), {dt:.1f} s; reference ransac.h loops + Eigen-free solver "
                      f"restatement (Eigen/PoseLib/Ceres absent), -O3 -march={ol.reference_fast_flags()} (the widest "
                      f"this host supports, i.e. what the reference's -march=native selects here), "
                      f"EvaluateModelOnPoint out-of-line as in the reference", "seconds": dt}


def run_reference(args):
    rank = int(os.environ.get("RANK", 0))
    if rank != 0:
        return
    wl = WORKLOADS[args.workload]
    n, k_total = wl["n"], wl["k"]
    p2, p3, f, inst = make_instance(n)
    from ransaclib_b200 import synth

    models = synth.perturbed_poses(1, inst["R_gt"], inst["c_gt"], 64)
    cores = os.cpu_count() or 1
    try:
        cores = len(os.sched_getaffinity(0))
    except Exception:
        pass
    # bounded sample per step: ~0.7 s of work per thread
    per_step_budget = 0.7
    res = []
    for i in range(args.warmup + args.steps):
        r = cpu_baseline(p2, p3, f, models, np.ones(len(models), np.uint8), threads=cores,
                         budget_s=per_step_budget)
        if i >= args.warmup:
            res.append(r)
    evals = sum(r["value"] * r["seconds"] for r in res)
    secs = sum(r["seconds"] for r in res)
    value = evals / secs
    world = int(os.environ.get("WORLD_SIZE", 1))
    line = {
        "impl": "reference", "metric": "MSAC hypothesis x point evals/sec", "value": value, "unit": "evals/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": secs / args.steps * 1e3,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": wl["name"], "n_points": n, "k_hypotheses": k_total,
                   "note": "CPU reference: each step is a bounded sample of the workload (all N points x a "
                           "subset of hypotheses) on every host core"},
        "cpu_baseline": {"value": value, "unit": "evals/s", "cores": cores, "kind": res[-1]["kind"],
                         "sample": res[-1]["sample"]},
        "e2e": {"value": value, "unit": "evals/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    if args.workload == "c5" and not args.no_localization:
        r = cpu_baseline_c4(max(cores * 4, 32), cores)
        line["localization"] = {"impl": "reference", "metric": "P3P LO-MSAC localizations/sec", "value": r["value"],
                                "unit": "localizations/s", "cpu_baseline": r,
                                "e2e": {"value": r["value"], "unit": "localizations/s", "h2d_bytes_per_step": 0,
                                        "d2h_bytes_per_step": 0}}
    emit(json.dumps(line))

# ------------------------------------------------- C4: batched localization (query-sharded)
C4 = dict(q=4096, n=2000, thr_px=12.0, lo_steps=10, min_iters=100, max_iters=10000,
          name="batched ransaclib_localization: 4096 queries x 2k matches, 50% outliers (configs[3])")


def c4_queries(indices):
    from ransaclib_b200 import synth

    out = []
    for q in indices:
        inst = synth.pose_instance(5000 + q, C4["n"] // 2, C4["n"] - C4["n"] // 2)
        out.append((inst["points2D"], inst["points3D"], np.float32(inst["focal"])))
    return out


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def c4_arrays(indices):
    """Concatenated (offsets, points2D, points3D, focal, seeds) of the given query indices."""
    data = c4_queries(indices)
    off = np.zeros(len(data) + 1, np.int64)
    for i, d in enumerate(data):
        off[i + 1] = off[i] + len(d[0])
    p2 = np.concatenate([d[0] for d in data])
    p3 = np.concatenate([d[1] for d in data])
    f = np.array([float(d[2]) for d in data])
    seeds = np.array([42 + q for q in indices], np.uint32)
    return off, p2, p3, f, seeds


def measure_c4(args, rank, world, local, dist, steps, warmup, q_total=None):
    """Batched localization on this rank's shard of the queries (query q on rank q mod world, NO
    collective: examples/localization.cc:190 has no cross-query state). Every query runs wholly on the
    device (rlb_localize_batch: one warp per query). Returns the JSON-ready dict on rank 0."""
    import torch

    import ransaclib_b200 as rlb
    from ransaclib_b200 import _capi, sharding

    dev = torch.device("cuda", local)
    q_total = q_total or args.queries or C4["q"]
    mine = sharding.query_indices(q_total, rank, world)
    off, p2, p3, f, seeds = c4_arrays(mine)
    nq = len(mine)
    pin = lambda a: torch.from_numpy(a).pin_memory().numpy()  # noqa: E731
    off, p2, p3, f, seeds_h = pin(off), pin(p2), pin(p3), pin(f), pin(seeds)

    ctx = rlb.Context(local)
    stream = torch.cuda.Stream(dev)
    torch.cuda.set_stream(stream)
    ctx.set_stream(stream.cuda_stream)
    opt = ctx.localize_options(C4["thr_px"], C4["lo_steps"], C4["min_iters"], C4["max_iters"])
    max_matches = int(np.max(np.diff(off)))

    # ---- value: matches already resident in HBM, CUDA events on the launching stream ----------
    d_off, d_p2, d_p3, d_f = (torch.from_numpy(a).to(dev) for a in (off, p2, p3, f))
    d_seeds = torch.from_numpy(seeds_h.view(np.int32)).to(dev)
    d_poses = torch.zeros((nq, 12), dtype=torch.float64, device=dev)
    d_res = torch.zeros(nq * _capi.LOCALIZE_RESULT.itemsize, dtype=torch.uint8, device=dev)
    d_inl = torch.zeros(int(off[-1]), dtype=torch.int32, device=dev)
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)

    def step_dev():
        ctx.localize_batch_dev(nq, max_matches, d_off.data_ptr(), d_p2.data_ptr(), d_p3.data_ptr(), d_f.data_ptr(),
                               d_f.data_ptr(), d_seeds.data_ptr(), opt, d_poses.data_ptr(), d_res.data_ptr(),
                               d_inl.data_ptr())

    for _ in range(warmup):
        step_dev()
    torch.cuda.synchronize()
    launches0 = ctx.launch_count
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    clocks = ClockSampler(local)
    clocks.start()
    time.sleep(0.25)
    w0 = time.time()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    for e0, e1 in evs:
        flush.zero_()  # L2 flush between timed steps (untimed)
        e0.record(stream)
        step_dev()
        e1.record(stream)
        e1.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    w1 = time.time()
    launches = ctx.launch_count - launches0
    clk = clocks.stop(w0, w1)
    total_ms = sum(a.elapsed_time(b) for a, b in evs)
    t = torch.tensor([total_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)  # timing only; the data path has no collective
    total_ms = float(t.item())
    value = q_total * steps / (total_ms * 1e-3)
    res_dev = d_res.cpu().numpy().view(_capi.LOCALIZE_RESULT)

    # ---- e2e: pinned host buffers through rlb_localize_batch (upload pipelined with the kernel,
    #      poses / results / inlier lists copied back), wall clock ------------------------------
    e2e_steps = max(1, min(steps, 5))
    ctx.localize_batch(off, p2, p3, f, f, seeds_h, opt)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        poses, res, inl = ctx.localize_batch(off, p2, p3, f, f, seeds_h, opt)
    e2e_s = time.perf_counter() - t0
    te = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_value = q_total * e2e_steps / float(te.item())
    assert (res["num_inliers"] == res_dev["num_inliers"]).all(), "host and device entry points disagree"
    if rank != 0:
        return None
    h2d = p2.nbytes + p3.nbytes + off.nbytes + 2 * f.nbytes + seeds_h.nbytes
    d2h = poses.nbytes + res.nbytes + inl.nbytes
    return {
        "metric": "P3P LO-MSAC localizations/sec", "value": value, "unit": "localizations/s", "n_gpus": world,
        "steps": steps, "warmup": warmup, "ms_per_step": total_ms / steps,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": C4["name"], "queries": q_total, "queries_per_gpu": nq,
                   "matches_per_query": C4["n"], "partitioning": "query q on rank q mod world, no collective",
                   "options": "ransaclib_localization.cc:100-107 (lo_start 60, final LSQ), seed 42+q, data seed 5000+q",
                   "l2": "flushed (256 MiB write) between timed steps",
                   "step": "one launch of the device-resident LO-MSAC kernel over all of this rank's queries"},
        "e2e": {"value": e2e_value, "unit": "localizations/s", "h2d_bytes_per_step": int(h2d),
                "d2h_bytes_per_step": int(d2h), "steps": e2e_steps,
                "path": "rlb_localize_batch with pinned host buffers (pyransaclib.ransaclib_localization_batch): "
                        "up to four chunks of queries, each chunk's grid on its own stream behind that chunk's copies "
                        "(cudaStreamWaitEvent; no kernel waits on the host), poses + stats + inlier lists back"},
        "gpu_launches": int(launches), "clocks": clk,
        "roofline": {"bound": "latency", "kernel": "localize_cta_kernel", "achieved": None, "peak": None, "unit": None,
                     "frac": None, "traffic": None,
                     "note": "control-flow kernel (one CTA = one whole LO-MSAC run: the master warp runs the control "
                             "flow, the four warps share every pass over the matches); per query on average "
                             "%.0f f64 passes over the matches, %.0f LM refinements, %.1f iterations, %.2f LO calls; "
                             "bound by the dependent-f64 chain of the LM refinements (62 %% of a loaded query), "
                             "per-phase counters and the timeline of a launch: profiles/r02_localize_ab.txt"
                             % (res["passes"].mean(), res["lsq_calls"].mean(), res["num_iterations"].mean(),
                                res["num_lo"].mean())},
        "result": {"mean_inliers": float(res["num_inliers"].mean()), "min_inliers": int(res["num_inliers"].min()),
                   "success": int(res["success"].sum())},
    }


def run_ours_c4(args):
    import torch

    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    line = measure_c4(args, rank, world, local, dist, args.steps, args.warmup)
    if rank == 0:
        line["cpu_baseline"] = cpu_baseline_c4(budget_queries=64, threads=1) if world == 1 else None
        emit(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def cpu_baseline_c4(budget_queries, threads):
    """The reference's LO-MSAC (its own ransac.h compiled in oracle/_ref) on `threads` host cores, one
    independent query per core at a time, with ransaclib_localization's options."""
    import ctypes as C
    from concurrent.futures import ThreadPoolExecutor

    ol, lib, kind = _load_ref()
    data = c4_queries(range(budget_queries))

    def one(i):
        p2, p3, f = data[i]
        o = ol.Options.default(sq_thr=float(np.float32(C4["thr_px"])) ** 2, seed=42 + i, final_lsq=1,
                               lo_steps=C4["lo_steps"], min_iters=C4["min_iters"], max_iters=C4["max_iters"],
                               lo_start=60)
        st = ol.Stats()
        model = np.zeros(12)
        inl = np.zeros(len(p2), np.int32)
        lib.lomsac_pose(C.byref(o), float(f), float(f), len(p2), ol.c64(p2), ol.c64(p3), model, C.byref(st), inl)
        return st.best_num_inliers

    t0 = time.perf_counter()
    if threads == 1:
        res = [one(i) for i in range(budget_queries)]
    else:
        with ThreadPoolExecutor(threads) as ex:
            res = list(ex.map(one, range(budget_queries)))
    dt = time.perf_counter() - t0
    return {"value": budget_queries / dt, "unit": "localizations/s", "cores": threads, "kind": kind,
            "sample": f"{budget_queries} of the {C4['q']} queries, {threads} thread(s), {dt:.1f} s; reference "
                      f"RansacLib/ransac.h EstimateModel + restated Eigen-free pose solver / LM refinement",
            "seconds": dt, "mean_inliers": float(np.mean(res))}


def run_reference_c4(args):
    if int(os.environ.get("RANK", 0)) != 0:
        return
    cores = host_threads()
    per_step = max(cores * 4, 32)
    res = []
    for i in range(args.warmup + args.steps):
        r = cpu_baseline_c4(per_step, cores)
        if i >= args.warmup:
            res.append(r)
    value = sum(per_step for _ in res) / sum(r["seconds"] for r in res)
    emit(json.dumps({
        "impl": "reference", "metric": "P3P LO-MSAC localizations/sec", "value": value, "unit": "localizations/s",
        "n_gpus": int(os.environ.get("WORLD_SIZE", 1)), "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": sum(r["seconds"] for r in res) / args.steps * 1e3, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": C4["name"], "note": f"each step = {per_step} queries on {cores} host cores"},
        "cpu_baseline": {"value": value, "unit": "localizations/s", "cores": cores, "kind": res[-1]["kind"],
                         "sample": res[-1]["sample"]},
        "e2e": {"value": value, "unit": "localizations/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c5", choices=list(WORKLOADS) + ["c4"])
    ap.add_argument("--queries", type=int, default=0, help="c4: total number of queries (default 4096)")
    ap.add_argument("--threads", type=int, default=0, help="c4: host threads per GPU (default: cores / gpus)")
    ap.add_argument("--precision", default="f32", choices=["f32", "f64"])
    ap.add_argument("--no-localization", action="store_true",
                    help="c5: skip the batched-localization leg that is reported under 'localization'")
    ap.add_argument("--no-counts", dest="counts", action="store_false",
                    help="scores only: pass no counts array to the sweep, which is what ScoreModel in ransac.h "
                         "computes (the winner's inliers come from the GetInliers pass either way). Default: "
                         "score AND inlier count per hypothesis, as BASELINE.json's north_star describes")
    ap.add_argument("--k-override", type=int, default=0, help="tuning only: total number of hypotheses")
    ap.add_argument("--no-extra", action="store_true",
                    help="skip the f64 leg, the line / hybrid sweep rooflines and the configs[0..2] table")
    ap.add_argument("--profile", action="store_true",
                    help="short run for ncu: skips the e2e leg, the CPU baseline and the peak microbenchmarks")
    args = ap.parse_args()
    claim_stdout()
    if args.workload == "c4":
        (run_reference_c4 if args.impl == "reference" else run_ours_c4)(args)
    elif args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
