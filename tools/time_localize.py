"""Quick timing of the batched localization for Q queries of the C4 shape.
usage: time_localize.py Q [dev|host] [reps]   (dev = inputs resident in HBM, CUDA-event timing)"""
import sys
import time

import numpy as np

sys.path.insert(0, ".")
import ransaclib_b200 as rlb
from ransaclib_b200 import _capi, synth

Q = int(sys.argv[1]) if len(sys.argv) > 1 else 512
mode = sys.argv[2] if len(sys.argv) > 2 else "host"
reps_t = int(sys.argv[3]) if len(sys.argv) > 3 else 3
base = 64
p2s, p3s, fs = [], [], []
for q in range(base):
    inst = synth.pose_instance(5000 + q, 1000, 1000)
    p2s.append(inst["points2D"]); p3s.append(inst["points3D"]); fs.append(float(np.float32(inst["focal"])))
reps = (Q + base - 1) // base
p2 = np.concatenate(p2s * reps)[: 2000 * Q]
p3 = np.concatenate(p3s * reps)[: 2000 * Q]
f = np.array((fs * reps)[:Q])
off = np.arange(Q + 1, dtype=np.int64) * 2000
seeds = np.arange(42, 42 + Q, dtype=np.uint32)
ctx = rlb.Context(0)
o = ctx.localize_options(12.0, 10, 100, 10000)


def report(dt, res):
    print(f"{mode} Q={Q} {dt*1e3:.2f} ms  {Q/dt:.0f} loc/s  mean inl {res['num_inliers'].mean():.1f} "
          f"iters {res['num_iterations'].mean():.1f} lo {res['num_lo'].mean():.2f} passes {res['passes'].mean():.0f} "
          f"lsq {res['lsq_calls'].mean():.0f}", flush=True)


if mode == "host":
    for it in range(reps_t):
        t0 = time.perf_counter()
        poses, res, inl = ctx.localize_batch(off, p2, p3, f, f, seeds, o)
        report(time.perf_counter() - t0, res)
else:
    import torch

    dev = torch.device("cuda", 0)
    st = torch.cuda.Stream(dev)
    torch.cuda.set_stream(st)
    ctx.set_stream(st.cuda_stream)
    d_off = torch.from_numpy(off).to(dev)
    d_p2 = torch.from_numpy(p2).to(dev)
    d_p3 = torch.from_numpy(p3).to(dev)
    d_f = torch.from_numpy(f).to(dev)
    d_seeds = torch.from_numpy(seeds.astype(np.int64)).to(dev).to(torch.int32)  # same 32 bits
    d_poses = torch.zeros((Q, 12), dtype=torch.float64, device=dev)
    d_res = torch.zeros(Q * _capi.LOCALIZE_RESULT.itemsize, dtype=torch.uint8, device=dev)
    d_inl = torch.zeros(2000 * Q, dtype=torch.int32, device=dev)
    for it in range(reps_t):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st)
        ctx.localize_batch_dev(Q, 2000, d_off.data_ptr(), d_p2.data_ptr(), d_p3.data_ptr(), d_f.data_ptr(),
                               d_f.data_ptr(), d_seeds.data_ptr(), o, d_poses.data_ptr(), d_res.data_ptr(),
                               d_inl.data_ptr())
        e1.record(st)
        e1.synchronize()
        res = d_res.cpu().numpy().view(_capi.LOCALIZE_RESULT)
        report(e0.elapsed_time(e1) * 1e-3, res)
