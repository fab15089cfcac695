"""Times one FakeEnv.step (ensemble forward + post-processing) at a given batch, reporting the forward
kernel alone (CUDA events inside the library) and the whole step."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import bench
import constrainedmbpo_b200 as cm

B = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 10
wname = sys.argv[3] if len(sys.argv) > 3 else "antsafe_uncertainty_H30_100k"
w = bench.WORKLOADS[wname]
p = bench.synth_params(w)
fe = cm.FakeEnv(cm.SimpleEnv(w["obs"], w["act"]), w["env"], num_networks=w["E"], num_elites=w["K"],
                hidden_dims=w["hidden"], cares_about_cost=w["cost"], seed=5)
fe._model.load_state(p.weights, p.biases, scaler_in=(p.mu_in, p.var_in), scaler_out=(p.mu_out, p.var_out), elites=p.elites)
obs = torch.from_numpy(bench.synth_start_states(w, B, 1)).cuda()
pol = cm.StubPolicy(w["obs"], w["act"])
act = pol(obs)
out = {}
for _ in range(3):
    fe.step_device(obs, act, want_ep_var=False, out=out)
torch.cuda.synchronize()
lib, h = fe._model._lib, fe._model._handle
lib.cmbpo_profile(h, 1)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(reps):
    fe.step_device(obs, act, want_ep_var=False, out=out)
e1.record()
torch.cuda.synchronize()
ms, n = C.c_double(), C.c_int64()
lib.cmbpo_profile_read(h, C.byref(ms), C.byref(n))
F = bench.flops_per_row(w)
fwd = ms.value / n.value
step = e0.elapsed_time(e1) / reps
print(f"B={B} fwd {fwd*1e3:.1f} us -> {F*B/fwd/1e9:.1f} TFLOP/s ({F*B/fwd/1e9/1372.7*100:.1f}% of 1372.7); "
      f"whole step {step*1e3:.1f} us -> {B/step/1e3:.2f} M rows/s")
