"""Bit-reproducibility of the kernels: the same launch twice must give identical bits (no races, no atomics in the data path)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench, constrainedmbpo_b200 as cm

for wname, B in (("halfcheetah_schedule_L10_1M", 1_000_000), ("antsafe_uncertainty_H30_100k", 100_000), ("halfcheetah_schedule_L10_1M", 4000)):
    w = bench.WORKLOADS[wname]
    fe = bench.build_env(w, 0)
    pol = cm.StubPolicy(w["obs"], w["act"])
    obs = torch.from_numpy(bench.synth_start_states(w, B, 1)).cuda()
    act = pol(obs).clone()
    x = torch.cat([obs, act], -1).contiguous()
    bad = {"predict": 0, "step": 0, "policy": 0}
    m0, v0 = [t.clone() for t in fe._model.predict_device([x], 2, 0)]
    fe._rng_offset = 7
    s0 = {k: v.clone() for k, v in fe.step_device(obs, act, want_ep_var=False).items()}
    for rep in range(8):
        m1, v1 = fe._model.predict_device([x], 2, 0)
        bad["predict"] += int(not (torch.equal(m0, m1) and torch.equal(v0, v1)))
        fe._rng_offset = 7
        s1 = fe.step_device(obs, act, want_ep_var=False)
        bad["step"] += int(not all(torch.equal(s0[k], s1[k]) for k in ("next_obs", "rew", "term", "dkl_path", "ep_var_mean")))
        bad["policy"] += int(not torch.equal(act, pol(obs)))
    print(wname, B, "mismatching repeats of 8:", bad, flush=True)
    T = min(w["T"], 6)
    ro = cm.ModelRollout(fe, T, rollout_mode="schedule")
    ro.capture(obs[: min(B, 200_000)].contiguous(), pol)
    ro.replay(); torch.cuda.synchronize()
    base = ro.rng_base.clone()
    ro.replay(); torch.cuda.synchronize()
    ref = {k: v.clone() for k, v in ro.pool.items()}
    nbad = 0
    for rep in range(6):
        ro.rng_base.copy_(base)
        ro.replay(); torch.cuda.synchronize()
        diff = [k for k, v in ro.pool.items() if not torch.equal(v, ref[k])]
        nbad += int(bool(diff))
        if diff:
            k = diff[0]
            d = (ro.pool[k] != ref[k])
            idx = d.nonzero()[:3].tolist()
            print("   replay", rep, "differs in", diff, "first at", idx, "count", int(d.sum()))
    print(wname, "rollout replays differing:", nbad, "of 6", flush=True)
