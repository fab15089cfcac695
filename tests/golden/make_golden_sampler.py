#!/usr/bin/env python
"""Golden fixtures from the reference's OWN sampler and model pool, run in this container.

Runs the unmodified ``ModelSampler.reset / sample / finish_all_paths`` (softlearning/samplers/
model_sampler.py:220-474, :497-570) against the unmodified ``ModelBuffer`` (softlearning/replay_pools/
modelbuffer.py:45-346) in all three rollout modes ('schedule' with max_samples, 'uncertainty', 'iv_gae'),
with a fake env / policy that feed seeded step outputs.  The package ``__init__`` files of the reference
pull in gym / tensorflow / ray, so the needed modules are loaded by FILE under their own dotted names,
with stubs for what is absent here (mpi4py at world size 1, matplotlib, the logger, the real-env sampler
base class -- none of which is on the path).

Everything the hot path owns is recorded: per sampler step the arrays handed to / returned by
``env.step`` and the policy outputs, and at the end the pool buffers (store layout, populated mask,
GAE / inverse-variance GAE outputs) and the sampler's accumulators.

  python tests/golden/make_golden_sampler.py      # writes tests/golden/sampler_<mode>.npz
"""
import importlib.util
import os
import sys
import types

import numpy as np

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

if not hasattr(np, "int"):   # removed in NumPy 1.24; the reference pins 1.17 (safe_utils/utils.py:159)
    np.int = int
if not hasattr(np, "bool"):
    np.bool = bool


def _stub(name, **attrs):
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    m.__path__ = []
    sys.modules[name] = m
    return m


def _load(name, rel):
    spec = importlib.util.spec_from_file_location(name, os.path.join(REF, rel))
    m = importlib.util.module_from_spec(spec)
    sys.modules[name] = m
    spec.loader.exec_module(m)
    return m


def load_reference():
    class _Comm:
        def Get_rank(self): return 0
        def Get_size(self): return 1
        def Allreduce(self, x, buf, op=None): buf[...] = x
        def Bcast(self, x, root=0): pass
    _stub("mpi4py", MPI=types.SimpleNamespace(COMM_WORLD=_Comm(), SUM="sum", MIN="min", MAX="max"))
    sys.modules["mpi4py.MPI"] = sys.modules["mpi4py"].MPI
    _stub("matplotlib"); _stub("matplotlib.pyplot")
    for pkg in ("softlearning", "softlearning.policies", "softlearning.policies.safe_utils",
                "softlearning.replay_pools", "softlearning.samplers"):
        _stub(pkg)
    _stub("softlearning.samplers.cpo_sampler", CpoSampler=type("CpoSampler", (object,), {}))
    _stub("softlearning.policies.safe_utils.logx", EpochLogger=type("EpochLogger", (object,), {}))
    _load("softlearning.policies.safe_utils.utils", "softlearning/policies/safe_utils/utils.py")
    _load("softlearning.policies.safe_utils.mpi_tools", "softlearning/policies/safe_utils/mpi_tools.py")
    _load("softlearning.replay_pools.cpobuffer", "softlearning/replay_pools/cpobuffer.py")
    mb = _load("softlearning.replay_pools.modelbuffer", "softlearning/replay_pools/modelbuffer.py")
    _load("softlearning.samplers.base_sampler", "softlearning/samplers/base_sampler.py")
    ms = _load("softlearning.samplers.model_sampler", "softlearning/samplers/model_sampler.py")
    return ms.ModelSampler, mb.ModelBuffer


from tests.helpers import make_inputs, make_params, orc  # noqa: E402

ENV, OBS, ACT, E = "antsafe", 29, 8, 7
N_V = 3   # value ensemble size in the 2-D modes (cpo_policy: 3 nets)


class _Space:
    def __init__(self, n): self.shape = (n,)


class FakeEnv:
    """Seeded stand-in for FakeEnv: the step arithmetic is the oracle's (its parity is pinned elsewhere); what
    matters here is the exact arrays the reference sampler receives and what it does with them."""
    def __init__(self, seed):
        self.num_networks = E
        self.dyn_target_var = 1
        self.observation_space, self.action_space = _Space(OBS), _Space(ACT)
        self.p = make_params(ENV, (24, 24), E=E, parity_scalers=False, bias=False, dtype=np.float32)
        self.rng = np.random.default_rng(seed)
        self.calls = []

    def step(self, obs, act):
        if obs.ndim == 3:
            perm = self.rng.permutation(E)
            out = orc.fake_env_step(self.p, ENV, obs, act, perm=perm, cares_about_cost=True)
        else:
            idx = self.rng.choice(self.p.elites, size=obs.shape[0])
            out = orc.fake_env_step(self.p, ENV, obs, act, elite_idx=idx, cares_about_cost=True)
        next_obs, rew, term, info = out
        info = {k: v for k, v in info.items() if not k.startswith("_")}
        self.calls.append(dict(cur=np.array(obs), act=np.array(act), next_obs=np.array(next_obs), rew=np.array(rew),
                               term=np.array(term), cost=np.array(info["cost"]),
                               dkl=np.array(info["ensemble_dkl_path"]), ep_var=np.array(info["ensemble_ep_var"])))
        return next_obs, rew, term, info


class FakePolicy:
    """Deterministic functions of the observation, shaped like CPOPolicy.get_action_outs / get_v / get_vc
    (cpo_policy.py:487-506, :931-944): 2-D obs -> v [N_V, B]; 3-D obs [E, B, .] -> v [E, B]."""
    gamma = cost_gamma = 0.99

    def __init__(self, seed):
        r = np.random.default_rng(seed)
        self.P = (r.standard_normal((OBS, ACT)) / np.sqrt(OBS)).astype(np.float32)
        self.wv = (r.standard_normal((E, OBS)) / np.sqrt(OBS)).astype(np.float32)
        self.wc = (r.standard_normal((E, OBS)) / np.sqrt(OBS)).astype(np.float32)
        self.agent = types.SimpleNamespace(reward_penalized=False)
        self.rng = np.random.default_rng(seed + 1)

    def reset(self): pass

    def _vals(self, obs, w):
        if obs.ndim == 3:
            return np.sin(np.einsum("ebd,ed->eb", obs, w)).astype(np.float32)
        return np.sin(obs @ w[:N_V].T).T.astype(np.float32)

    def get_v(self, obs, factored=True): return self._vals(obs, self.wv)
    def get_vc(self, obs, factored=True): return 0.5 * self._vals(obs, self.wc) + 0.5

    def get_action_outs(self, obs, factored=True, inc_var=True):
        v, vc = self.get_v(obs), self.get_vc(obs)
        return {"pi": np.tanh(obs @ self.P).astype(np.float32), "logp_pi": np.zeros(obs.shape[:-1], np.float32),
                "pi_info": {}, "v": v, "vc": vc, "v_var": 0.01 * np.abs(v) + 1e-3, "vc_var": 0.01 * np.abs(vc) + 1e-3}

    def random_v_inds(self, n): return self.rng.integers(0, N_V, n)
    def random_vc_inds(self, n): return self.rng.integers(0, N_V, n)


def run(mode, B, T, seed, max_samples=None, dkl_lim=None, max_unc=(5.5, 5.5), max_calls=None):
    ModelSampler, ModelBuffer = load_reference()
    np.random.seed(seed)   # ModelBuffer.model_ind / sampler.model_inds use the global RNG
    env, pol = FakeEnv(seed), FakePolicy(seed + 10)
    pool = ModelBuffer(B, env, T, E, rollout_mode=mode, cares_about_cost=True,
                       max_uncertainty_r=max_unc[0], max_uncertainty_c=max_unc[1])
    pool.initialize({}, gamma=0.99, lam=0.95, cost_gamma=0.97, cost_lam=0.9)   # distinct cost discounts on purpose
    sampler = ModelSampler(T, batch_size=B, cares_about_cost=True, rollout_mode=mode, logger=object())
    sampler.initialize(env, pol, pool)
    if dkl_lim is not None:
        sampler.set_rollout_dkl(dkl_lim)
    obs, _ = make_inputs(ENV, B, seed=seed)
    sampler.reset(obs)
    # log what the sampler hands to finish_path_multiple (path ids, bootstrap values, pool column at that moment)
    fin_ptr = np.full(B, -1, np.int64)
    lead = (E,) if mode == "iv_gae" else ()
    last_val, last_cval = np.zeros(lead + (B,), np.float32), np.zeros(lead + (B,), np.float32)
    orig_finish = pool.finish_path_multiple

    def logged_finish(term_mask, last_val_=0, last_cval_=0):
        if term_mask.any():
            fin = np.where(pool.alive_paths)[0][term_mask]
            fin_ptr[fin] = pool.ptr
            last_val[..., fin] = last_val_
            last_cval[..., fin] = last_cval_
        return orig_finish(term_mask, last_val_, last_cval_)
    pool.finish_path_multiple = logged_finish
    alive_hist = []
    n_calls = 0
    while pool.alive_paths.any():
        if max_calls is not None and n_calls >= max_calls:
            break      # CMBPO._train leaves the loop early on alive_ratio / the sample budget (cmbpo.py:339-348):
                       # finish_all_paths then bootstraps every still-alive path (model_sampler.py:550-570)
        alive_hist.append(int(pool.alive_paths.sum()))
        sampler.sample(max_samples=max_samples)
        n_calls += 1
        assert n_calls <= T + 2
    alive_at_break = np.array(pool.alive_paths)
    sampler.finish_all_paths()
    out = dict(alive_at_break=alive_at_break, start_obs=obs, mode=mode, T=T, E=E, max_samples=max_samples or 0, dkl_lim=dkl_lim or 0.0,
               model_ind=pool.model_ind, alive_hist=np.array(alive_hist), n_steps=len(env.calls),
               total_samples=sampler._total_samples, path_length=sampler._path_length,
               dyn_dkl_path=sampler._dyn_dkl_path, ptr=pool.ptr, max_uncertainty=np.array(max_unc),
               gamma=0.99, lam=0.95, cost_gamma=0.97, cost_lam=0.9, finish_ptr=fin_ptr, last_val=last_val, last_cval=last_cval)
    for k in ("obs_buf", "act_buf", "nextobs_buf", "rew_buf", "cost_buf", "term_buf", "dyn_error_buf",
              "populated_mask", "terminated_paths_mask", "val_buf", "cval_buf", "val_var_buf", "cval_var_buf",
              "adv_buf", "ret_buf", "cadv_buf", "cret_buf", "ret_var_buf", "cret_var_buf", "roll_lengths_buf",
              "croll_lengths_buf"):
        out[k] = np.array(getattr(pool, k))
    # ModelBuffer.get (modelbuffer.py:348-405): advantage normalisation + populated-mask selection (resets the pool)
    res, diag = pool.get()
    names = ["obs", "act", "adv", "ret_var", "cadv", "cret_var", "ret", "cret", "logp", "val", "val_var", "cval",
             "cval_var", "cost"]
    for k, v in zip(names, res):
        out[f"get_{k}"] = np.array(v)
    for k, v in diag.items():
        out[f"diag_{k}"] = np.array(v)
    for i, c in enumerate(env.calls):
        for k, v in c.items():
            out[f"step{i}_{k}"] = v
    return out


def main():
    cfgs = {
        "schedule": dict(mode="schedule", B=48, T=7, seed=21, max_samples=200),
        "uncertainty": dict(mode="uncertainty", B=64, T=9, seed=22, dkl_lim=None),
        "iv_gae": dict(mode="iv_gae", B=40, T=7, seed=23, max_unc=(1.6, 1.6)),
        # the rollout loop left after 3 sampler steps with paths still alive: finish_all_paths does real work
        "earlybreak": dict(mode="uncertainty", B=56, T=10, seed=24, dkl_lim=None, max_calls=3),
    }
    # calibrate a dkl limit that cuts paths at different lengths: mean first-step path KL x 3
    probe = run("schedule", 64, 3, 22)
    cfgs["uncertainty"]["dkl_lim"] = float(np.mean(probe["step0_dkl"]) * 3.0)
    cfgs["earlybreak"]["dkl_lim"] = float(np.mean(probe["step0_dkl"]) * 6.0)
    for name, kw in cfgs.items():
        out = run(**kw)
        path = os.path.join(HERE, f"sampler_{name}.npz")
        np.savez_compressed(path, **out)
        pop = out["populated_mask"]
        print(f"{name}: steps={out['n_steps']} alive={list(out['alive_hist'])} stored={int(out['total_samples'])} "
              f"populated={int(pop.sum())} -> {path} ({os.path.getsize(path) / 1e3:.0f} kB)")


if __name__ == "__main__":
    main()
