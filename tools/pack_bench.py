#!/usr/bin/env python
"""Times cmbpo_pack_records on the pool of one default-workload rollout (1 GPU)."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench, constrainedmbpo_b200 as cm
from constrainedmbpo_b200 import sharding, _lib

wname = sys.argv[1] if len(sys.argv) > 1 else bench.DEFAULT_WORKLOAD
w = bench.WORKLOADS[wname]
fe = bench.build_env(w, 0)
pol = cm.StubPolicy(w["obs"], w["act"], device="cuda:0")
lim = None
if w["mode"] == "uncertainty":
    lim = cm.compute_dynamics_dkl(fe, pol, torch.from_numpy(bench.synth_start_states(w, 5000, 99)).cuda(), depth=w["dkl_depth"])
ro = cm.ModelRollout(fe, w["T"], rollout_mode=w["mode"], dkl_lim=lim)
start = torch.from_numpy(bench.synth_start_states(w, w["rows"], 1)).cuda()
ro.reset(start); ro.run(pol); torch.cuda.synchronize()
E = w["E"] if w["mode"] == "iv_gae" else 1
W = sharding.record_width(w["obs"], w["act"], E)
n = int(ro.pool["populated"].sum())
out = torch.empty(n + 8, W, device="cuda"); count = torch.zeros(1, dtype=torch.int64, device="cuda")
scratch = torch.empty(int(_lib.load().cmbpo_pack_scratch_bytes(w["rows"])) + 64, dtype=torch.uint8, device="cuda")
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for i in range(6):
    if i == 1:
        e0.record()
    sharding.pack_records_cuda(ro.pool, 0, w["obs"], w["act"], E, out, count, scratch)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 5
print(f"{wname}: {n} records x {W} cols ({n*W*4/1e6:.0f} MB): pack {ms*1e3:.1f} us -> {2*n*W*4/ms/1e6:.0f} GB/s (read+write)")
