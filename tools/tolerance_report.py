#!/usr/bin/env python
"""tf32 vs bf16 tolerance report over the width sweep of BASELINE.json configs[4]: BNN.predict (factored) of the CUDA
path against the fp64 oracle for antsafe-shaped ensembles of hidden width 200 / 256 / 512 / 1024 (4 layers, random init
per fc.py:154-185, fitted-looking scalers incl. two input dims at the 1e-2 floor), on 4096 seeded rows, for both operand
types.  Errors are in units of the output scaler's divisor s_out (the natural unit of a head column) and as the fraction
of elements inside the north_star tolerance |y - y_ref| <= rtol*|y_ref| + rtol*s_out with rtol = 1e-2 (bf16) / 1e-3 (tf32).
Also times the forward (CUDA events inside the library) and relates it to the measured torch.matmul peak of the same
operand type (same method as MEASURED_PEAKS.json: 8192^3, best of 10).
  python tools/tolerance_report.py > profiles/r02_tolerance_report.txt"""
import ctypes as C
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tests.helpers import build_cuda_env, make_inputs, make_params, orc, out_scale  # noqa: E402


def matmul_peak(dtype, tf32):
    torch.backends.cuda.matmul.allow_tf32 = tf32
    a = torch.randn(8192, 8192, device="cuda", dtype=dtype)
    b = torch.randn(8192, 8192, device="cuda", dtype=dtype)
    best = 1e9
    for _ in range(3):
        a @ b
    for _ in range(10):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); a @ b; e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    torch.backends.cuda.matmul.allow_tf32 = False
    return 2 * 8192 ** 3 / (best * 1e-3) / 1e12


B = 4096
peaks = {"bf16": matmul_peak(torch.bfloat16, False), "tf32": matmul_peak(torch.float32, True)}
print(f"# operands vs fp64 oracle, antsafe dims (in 37, out 2x30), 7 members, {B} rows, 4 hidden layers")
print(f"# measured torch.matmul 8192^3 burst peaks on this GPU: bf16 {peaks['bf16']:.0f} TFLOP/s, tf32 {peaks['tf32']:.0f} TFLOP/s")
print(f"{'dtype':>5} {'width':>6} {'kernel':>20} {'mean |err|/s':>13} {'p99.9 |err|/s':>14} {'max |err|/s':>12} "
      f"{'inside tol':>11} {'max |dlogvar|':>14} {'fwd us @262144':>15} {'TFLOP/s':>8} {'of peak':>8}")
for dt, rtol in (("bf16", 1e-2), ("tf32", 1e-3)):
    for H in (200, 256, 512, 1024):
        hidden = (H,) * 4
        p = make_params("antsafe", hidden)
        obs, act = make_inputs("antsafe", B)
        fe = build_cuda_env(p, "antsafe", hidden, operand_dtype=dt)
        x = np.concatenate((obs, act), -1)
        mean, var = fe._model.predict(x, factored=True, inc_var=True)
        rmean, rvar = orc.predict(p, x.astype(np.float64), factored=True, inc_var=True)
        s = out_scale(p)
        err = np.abs(mean - rmean) / s
        inside = (np.abs(mean - rmean) <= rtol * np.abs(rmean) + rtol * s).mean()
        dlv = np.abs(np.log(var.astype(np.float64)) - np.log(rvar))
        # forward timing at 262144 rows
        Bt = 262144
        o2, a2 = make_inputs("antsafe", Bt, seed=3)
        xt = torch.from_numpy(np.concatenate((o2, a2), -1)).cuda()
        m = fe._model
        for _ in range(2):
            m.predict_device([xt], 2, 0)
        torch.cuda.synchronize()
        lib, h = m._lib, m._handle
        lib.cmbpo_profile(h, 1)
        for _ in range(3):
            m.predict_device([xt], 2, 0)
        torch.cuda.synchronize()
        ms, n = C.c_double(), C.c_int64()
        lib.cmbpo_profile_read(h, C.byref(ms), C.byref(n))
        lib.cmbpo_profile(h, 0)
        fwd = ms.value / 3
        dims = [37] + list(hidden) + [60]
        F = 2 * 7 * sum(a * b for a, b in zip(dims[:-1], dims[1:]))
        tf = F * Bt / (fwd * 1e-3) / 1e12
        kern = ("ensemble_fwd", "wide_layer", "fused_wide")[lib.cmbpo_forward_path(h)] + (" tf32" if dt == "tf32" else "")
        print(f"{dt:>5} {H:>6} {kern:>20} {err.mean():13.2e} {np.quantile(err, 0.999):14.2e} {err.max():12.2e} {inside:11.6f} "
              f"{dlv.max():14.2e} {fwd * 1e3:15.1f} {tf:8.1f} {tf / peaks[dt]:8.3f}", flush=True)
        del fe, xt
        torch.cuda.empty_cache()
print("# tolerance: |y - y_ref| <= rtol*|y_ref| + rtol*s_out, rtol 1e-2 for bf16 operands, 1e-3 for tf32 (north_star).")
print("# tf32: fp32 operand images rounded to nearest tf32, layer 0 as x_hi*W_hi + x_lo*W_hi + x_hi*W_lo, exact sigmoid/tanh epilogue.")
