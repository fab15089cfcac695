"""Forward-only timing of the ensemble kernels over the BASELINE.json sweep (hidden width x batch):
  python tools/wide_bench.py [--widths 200,256,512,1024] [--rows 65536,262144,1000000] [--layers 4]
Prints, per point, the forward time (CUDA events inside the library around all launches of one forward),
algorithmic TFLOP/s (2*E*(in*H + (L-1)*H^2 + H*2D), unpadded) and the fraction of the measured
sustained bf16 peak."""
import argparse
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import bench
import constrainedmbpo_b200 as cm

ap = argparse.ArgumentParser()
ap.add_argument("--widths", default="200,256,512,1024")
ap.add_argument("--rows", default="65536,262144,1000000")
ap.add_argument("--layers", type=int, default=4)
ap.add_argument("--reps", type=int, default=5)
args = ap.parse_args()
peak = bench.measured_peaks()["bf16_sustained"]
for H in [int(x) for x in args.widths.split(",")]:
    w = dict(bench.WORKLOADS["antsafe_step_1M"], hidden=(H,) * args.layers)
    fe = bench.build_env(w, 0)
    lib, h = fe._model._lib, fe._model._handle
    pol = cm.StubPolicy(w["obs"], w["act"])
    F = bench.flops_per_row(w)
    for B in [int(x) for x in args.rows.split(",")]:
        if B * H > 1_100_000_000:
            continue
        obs = torch.from_numpy(bench.synth_start_states(w, B, 1)).cuda()
        act = pol(obs)
        out = {}
        for _ in range(2):
            fe.step_device(obs, act, want_ep_var=False, out=out)
        torch.cuda.synchronize()
        lib.cmbpo_profile(h, 1)
        for _ in range(args.reps):
            fe.step_device(obs, act, want_ep_var=False, out=out)
        torch.cuda.synchronize()
        ms, n = C.c_double(), C.c_int64()
        lib.cmbpo_profile_read(h, C.byref(ms), C.byref(n))
        lib.cmbpo_profile(h, 0)
        fwd = ms.value / n.value
        tf = F * B / fwd / 1e9
        print(f"hidden {args.layers}x{H} rows {B}: fwd {fwd * 1e3:9.1f} us  {tf:7.1f} TFLOP/s  ({100 * tf / peak:5.1f}% of {peak:.0f} sustained bf16)  "
              f"{B / fwd / 1e3:8.2f} M rows/s", flush=True)
        del obs, act, out
    del fe
    torch.cuda.empty_cache()
