#!/usr/bin/env python
"""bf16 tolerance report over the width sweep of BASELINE.json configs[4]: BNN.predict (factored) of the CUDA path
against the fp64 oracle for antsafe-shaped ensembles of hidden width 200 / 256 / 512 / 1024 (4 layers, random init per
fc.py:154-185, fitted-looking scalers), on 4096 seeded rows.  Errors are reported in units of the output scaler's
divisor s_out (the natural unit of a head column) and as the fraction of elements inside the north_star tolerance
|y - y_ref| <= 1e-2*|y_ref| + 2e-2*s_out.  tf32 operands are not built: bf16 only.
  python tools/tolerance_report.py > profiles/r01_tolerance_report.txt"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tests.helpers import build_cuda_env, make_inputs, make_params, orc, out_scale  # noqa: E402

B = 4096
print(f"# bf16 operands / fp32 accumulate vs fp64 oracle, antsafe dims (in 37, out 2x30), 7 members, {B} rows, 4 hidden layers")
print(f"{'width':>6} {'kernel':>18} {'mean |err|/s_out':>17} {'p99.9 |err|/s_out':>18} {'max |err|/s_out':>16} "
      f"{'inside tol':>11} {'max |dlogvar|':>14} {'var rel err p99.9':>18}")
for H in (200, 256, 512, 1024):
    hidden = (H,) * 4
    p = make_params("antsafe", hidden)
    obs, act = make_inputs("antsafe", B)
    fe = build_cuda_env(p, "antsafe", hidden)
    x = np.concatenate((obs, act), -1)
    mean, var = fe._model.predict(x, factored=True, inc_var=True)
    rmean, rvar = orc.predict(p, x.astype(np.float64), factored=True, inc_var=True)
    s = out_scale(p)
    err = np.abs(mean - rmean) / s
    inside = (np.abs(mean - rmean) <= 1e-2 * np.abs(rmean) + 2e-2 * s).mean()
    dlv = np.abs(np.log(var.astype(np.float64)) - np.log(rvar))
    vrel = np.abs(var - rvar) / rvar
    kern = "ensemble_fwd (fused)" if H <= 254 else "wide_layer"
    print(f"{H:>6} {kern:>18} {err.mean():17.2e} {np.quantile(err, 0.999):18.2e} {err.max():16.2e} {inside:11.6f} "
          f"{dlv.max():14.2e} {np.quantile(vrel, 0.999):18.2e}")
