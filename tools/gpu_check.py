"""Quick on-GPU numerical check of predict/step against the oracle (prints error statistics)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from tests.helpers import make_params, make_inputs, build_cuda_env, out_scale, orc

env, hidden = "antsafe", (200, 200, 200, 200)
B = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
p = make_params(env, hidden)
obs, act = make_inputs(env, B)
fe = build_cuda_env(p, env, hidden, cares_about_cost=True)
x = np.concatenate([obs, act], -1)
t0 = time.time()
mean, var = fe._model.predict(x, factored=True)
torch.cuda.synchronize()
print("predict ok", mean.shape, var.shape, "time", time.time() - t0)
rm, rv = orc.predict(p, x.astype(np.float64), factored=True)
s = out_scale(p)
err_m = np.abs(mean - rm) / (np.abs(rm) + s)
print("mean: max |err|/(|ref|+s_out) =", err_m.max(), " median", np.median(err_m))
err_v = np.abs(var - rv) / rv
print("var: max rel err =", err_v.max(), " median", np.median(err_v))
idx = np.random.default_rng(7).choice(p.elites, size=B)
no, rw, tm, info = fe.step(obs, act, elite_idx=idx)
rno, rrw, rtm, rinfo = orc.fake_env_step(p, env, obs.astype(np.float64), act.astype(np.float64), elite_idx=idx, cares_about_cost=True)
print("next_obs max abs err", np.abs(no - rno).max(), "rew", np.abs(rw - rrw).max())
print("term mismatches", int((tm != rtm).sum()), "of", B, " cost mismatches", int((info['cost'] != rinfo['cost']).sum()))
dk, rdk = info['ensemble_dkl_path'], rinfo['ensemble_dkl_path']
print("dkl_path rel err max", (np.abs(dk - rdk) / (np.abs(rdk) + 1e-6)).max(), "median ref", np.median(rdk))
ev, rev = info['ensemble_ep_var'], rinfo['ensemble_ep_var']
print("ep_var rel err max", (np.abs(ev - rev) / (np.abs(rev) + 1e-9)).max())
for k in ('ensemble_dkl_mean', 'ensemble_mean_var', 'rew_mean', 'cost_mean'):
    print(k, float(info[k]), float(rinfo[k]))
print("status", fe._model._lib.cmbpo_status(fe._model._handle, None))
