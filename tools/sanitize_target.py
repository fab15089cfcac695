"""Target of the compute-sanitizer runs (profiles/r02_sanitizer_*.txt): one FakeEnv.step at the two benchmarked regimes of
the forward kernel -- 100 000 rows (G = 7 member groups, ~20 tile pairs per cluster) and 1 000 000 rows (G = 1, all 7 members
per cluster, ~27 pairs) -- plus one short 'uncertainty' rollout (K3 bookkeeping) and a 3-D step, results checked for finiteness.
  compute-sanitizer --tool memcheck  python tools/sanitize_target.py [rows ...]
  compute-sanitizer --tool racecheck python tools/sanitize_target.py [rows ...]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench, constrainedmbpo_b200 as cm

sizes = [int(a) for a in sys.argv[1:]] or [100_000, 1_000_000]
w = bench.WORKLOADS["antsafe_uncertainty_H30_100k"]
fe = bench.build_env(w, 0)
pol = cm.StubPolicy(w["obs"], w["act"])
for B in sizes:
    obs = torch.from_numpy(bench.synth_start_states(w, B, 1)).cuda()
    act = pol(obs).clone()
    o = fe.step_device(obs, act, want_ep_var=False)
    torch.cuda.synchronize()
    fe._model.check_status()
    assert torch.isfinite(o["next_obs"]).all() and torch.isfinite(o["dkl_path"]).all()
    print(f"step {B} rows ok: dkl_mean {float(o['scalars'][0]):.4f}", flush=True)
lim = cm.compute_dynamics_dkl(fe, pol, torch.from_numpy(bench.synth_start_states(w, 2000, 99)).cuda(), depth=2)
ro = cm.ModelRollout(fe, 8, rollout_mode="uncertainty", dkl_lim=lim)
ro.reset(torch.from_numpy(bench.synth_start_states(w, 20_000, 3)).cuda()); ro.run(pol); torch.cuda.synchronize()
print("rollout ok:", ro.counters(), flush=True)
w3 = bench.WORKLOADS["humanoidsafe_iv_gae_H12_125k"]
fe3 = bench.build_env(w3, 0)
o3 = torch.from_numpy(bench.synth_start_states(w3, 4000, 1)).cuda()[None].repeat(7, 1, 1).contiguous()
a3 = cm.StubPolicy(w3["obs"], w3["act"])(o3).clone()
r3 = fe3.step_device(o3, a3, perm=np.arange(7)[::-1].copy(), want_ep_var=False); torch.cuda.synchronize()
assert torch.isfinite(r3["next_obs"]).all()
print("3-D step ok", flush=True)
