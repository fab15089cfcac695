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
wname = sys.argv[3] if len(sys.argv) > 3 else "antsafe_step_1M"
w = bench.WORKLOADS[wname]
fe = bench.build_env(w, 0)
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
pk = bench.measured_peaks()
tf = F * B / fwd / 1e9
print(f"{wname} B={B} fwd {fwd*1e3:.1f} us -> {tf:.1f} TFLOP/s ({tf/pk['bf16_burst']*100:.1f}% of burst {pk['bf16_burst']:.0f}, "
      f"{tf/pk['bf16_sustained']*100:.1f}% of sustained {pk['bf16_sustained']:.0f}); whole step {step*1e3:.1f} us -> {B/step/1e3:.2f} M rows/s")
