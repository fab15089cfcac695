import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import constrainedmbpo_b200 as cm
from tests.helpers import build_cuda_env, make_inputs, make_params
env, B, T = "halfcheetah", 4000, 5
p = make_params(env, (200,) * 4, parity_scalers=False, bias=False)
obs, _ = make_inputs(env, B)
fe = build_cuda_env(p, env, (200,) * 4, seed=3)
ro = cm.ModelRollout(fe, T, rollout_mode="schedule")
start = torch.from_numpy(obs).cuda()
ro.capture(start, cm.StubPolicy(17, 6))
ro.replay(); torch.cuda.synchronize()
base = ro.rng_base.clone()
snaps = []
for k in range(6):
    ro.rng_base.copy_(base)
    ro.replay(); torch.cuda.synchronize()
    snaps.append({n: v.clone() for n, v in ro.pool.items()})
for k in range(1, 6):
    for n in ("obs", "act", "next_obs", "rew", "dyn_error"):
        d = snaps[k][n] != snaps[0][n]
        if d.any():
            idx = d.nonzero()
            ts = torch.unique(idx[:, 1]).tolist()
            mag = (snaps[k][n] - snaps[0][n]).abs().max().item()
            print(f"replay {k} vs 0: {n}: {int(d.sum())} differ, t in {ts}, rows {idx[:3,0].tolist()} max abs {mag:.3e}")
print("done")
# eager (no graph) repeats
ro2 = cm.ModelRollout(fe, T, rollout_mode="schedule")
pol = cm.StubPolicy(17, 6)
outs = []
for k in range(4):
    ro2.rng_base.zero_(); fe._rng_offset = 50
    ro2.reset(start); ro2.run(pol); torch.cuda.synchronize()
    outs.append(ro2.pool["next_obs"].clone())
print("eager repeats equal:", [bool(torch.equal(outs[0], o)) for o in outs[1:]])
