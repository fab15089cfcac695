"""The package driven by the reference's CALLERS through the drop-in boundary (NumPy in, NumPy out):

* the sampler loop of ``ModelSampler.sample`` (softlearning/samplers/model_sampler.py:265-474) -- its restatement in
  ``oracle.rollout`` is pinned bit for bit to the reference's REAL ``ModelSampler`` + ``ModelBuffer``
  (tests/test_oracle_golden.py) -- with ``env = constrainedmbpo_b200.FakeEnv``, free-running, against the pools the REAL
  sampler produced around the fp64 oracle env with the same injected draws (tests/golden/sampler_*.npz);
* ``MBPO._rollout_model`` (mbpo/algorithms/mbpo.py:509-533), replayed line for line, with ``fake_env`` = this package's
  ``FakeEnv`` vs the oracle env, the elite draws coming from the global NumPy RNG exactly as upstream
  (``BNN.random_inds``, mbpo/models/bnn.py:437-439).

/root/reference is not available on the GPU box, so the real classes cannot be imported here; the fixtures were written by
them in the build container (tests/golden/make_golden_sampler.py)."""
import os

import numpy as np
import pytest

from tests.helpers import (ENV_DIMS, build_cuda_env, make_inputs, make_params, near_threshold_rows, orc, out_scale)

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.mark.parametrize("name,seed", [("schedule", 21), ("uncertainty", 22)])
def test_sampler_loop_drives_b200_fakeenv_like_the_reference_sampler(name, seed):
    g = np.load(os.path.join(GOLDEN, f"sampler_{name}.npz"), allow_pickle=False)
    B, T, E = g["start_obs"].shape[0], int(g["T"]), int(g["E"])
    hidden = (24, 24)
    p = make_params("antsafe", hidden, E=E, parity_scalers=False, bias=False, dtype=np.float32)
    fe = build_cuda_env(p, "antsafe", hidden, cares_about_cost=True)
    r = np.random.default_rng(seed + 10)                       # FakePolicy of the fixture: a = tanh(obs @ P)
    P = (r.standard_normal((29, 8)) / np.sqrt(29)).astype(np.float32)
    policy = lambda obs: np.tanh(np.asarray(obs, np.float32) @ P).astype(np.float32)
    draws = np.random.default_rng(seed)                        # the fixture env's elite draws, same call order
    n_calls = {"n": 0}

    def step_fn(cur, a, step):
        idx = draws.choice(p.elites, size=cur.shape[0])
        n_calls["n"] += 1
        no, rw, tm, info = fe.step(np.asarray(cur, np.float32), np.asarray(a, np.float32), elite_idx=idx)
        assert isinstance(no, np.ndarray) and no.dtype == np.float32 and tm.dtype == np.bool_      # the boundary: NumPy out
        return no, rw, tm, info

    pool = orc.rollout(p, "antsafe", g["start_obs"], policy, T, name, dkl_lim=float(g["dkl_lim"]) or None,
                       max_samples=int(g["max_samples"]) or None, step_fn=step_fn)
    ref_pop, pop = g["populated_mask"], pool["populated"]
    same_len = pop.sum(1) == ref_pop.sum(1)
    # bf16 operands move next_obs by ~1e-2 * scale per step: a path whose termination / KL-cutoff decision sat within that
    # of a threshold may end one step earlier or later; everything else must follow the reference's bookkeeping exactly
    assert same_len.mean() >= 0.9, f"{int((~same_len).sum())} of {B} paths differ in length"
    if same_len.all():
        assert n_calls["n"] == int(g["n_steps"]) and pool["total_samples"] == int(g["total_samples"])
        assert np.array_equal(pop, ref_pop)
    s = out_scale(p)
    rows = np.where(same_len)[0]
    m = pop[rows]
    for ours, ref, scale in (("obs", "obs_buf", s[:-1]), ("next_obs", "nextobs_buf", s[:-1]), ("rew", "rew_buf", s[-1]),
                             ("act", "act_buf", 1.0)):
        a, b = np.asarray(pool[ours])[rows][m], g[ref][rows][m]
        t_idx = np.broadcast_to(np.arange(T)[None, :], m.shape)[m]
        tol = (2e-2 * (t_idx + 1))                             # errors compound over the steps of a free-running path
        tol = tol[:, None] if a.ndim == 2 else tol
        assert (np.abs(a - b) <= tol * (np.abs(b) + scale)).all(), (ours, float(np.abs(a - b).max()))
    # first sampler step: same inputs on both sides => plain single-step tolerance and identical decisions off thresholds
    first_no, ref_no = np.asarray(pool["next_obs"])[:, 0], g["nextobs_buf"][:, 0]
    assert (np.abs(first_no - ref_no) <= 1e-2 * np.abs(ref_no) + 2e-2 * s[:-1]).all()
    near = near_threshold_rows("antsafe", ref_no, 1e-2 * np.abs(ref_no) + 2e-2 * s[:-1])
    mism = np.asarray(pool["term"])[:, 0] != g["term_buf"][:, 0]
    assert not (mism & ~near).any()


@pytest.mark.parametrize("env", ["halfcheetah", "antsafe"])
def test_mbpo_rollout_model_replay(env):
    """mbpo/algorithms/mbpo.py:509-533, the smallest real caller of ``FakeEnv.step``: NumPy observations in, NumPy
    transitions out, terminated rows dropped with a boolean mask, elite members drawn from the global NumPy RNG."""
    hidden = (200, 200, 200, 200)
    obs_dim, act_dim = ENV_DIMS[env]
    p = make_params(env, hidden)
    fe = build_cuda_env(p, env, hidden)
    P = (np.random.default_rng(4).standard_normal((obs_dim, act_dim)) / np.sqrt(obs_dim)).astype(np.float32)
    actions_np = lambda o: np.tanh(o @ P).astype(np.float32)

    class OracleEnv:                                           # what the reference FakeEnv computes, in fp64
        def step(self, obs, act):
            idx = np.random.choice(p.elites, size=obs.shape[0])            # BNN.random_inds, bnn.py:437-439
            return orc.fake_env_step(p, env, obs.astype(np.float64), act.astype(np.float64), elite_idx=idx)

    def rollout_model(fake_env, obs, rollout_length):          # mbpo.py:514-533
        pool, steps_added = [], []
        for i in range(rollout_length):
            act = actions_np(obs)
            next_obs, rew, term, info = fake_env.step(obs, act)
            steps_added.append(len(obs))
            pool.append({'observations': obs, 'actions': act, 'next_observations': next_obs, 'rewards': rew, 'terminals': term})
            nonterm_mask = ~term.squeeze(-1)
            if nonterm_mask.sum() == 0:
                break
            obs = next_obs[nonterm_mask]
        return pool, steps_added

    start, _ = make_inputs(env, 4000, seed=31)
    np.random.seed(123)
    got, n_got = rollout_model(fe, start, 4)
    np.random.seed(123)
    ref, n_ref = rollout_model(OracleEnv(), start, 4)
    s = out_scale(p)
    assert got[0]['next_observations'].dtype == np.float32 and got[0]['terminals'].dtype == np.bool_
    assert got[0]['rewards'].shape == (4000, 1) and got[0]['terminals'].shape == (4000, 1)
    # step 0: identical inputs and draws
    a, b = got[0], ref[0]
    assert (np.abs(a['next_observations'] - b['next_observations']) <= 1e-2 * np.abs(b['next_observations']) + 1e-2 * s[:-1]).all()
    near = near_threshold_rows(env, b['next_observations'], 1e-2 * np.abs(b['next_observations']) + 1e-2 * s[:-1])
    mism = (a['terminals'] != b['terminals'])[:, 0]
    assert not (mism & ~near).any()
    if env == "halfcheetah":                                   # never terminates: row sets (and RNG draws) stay aligned
        assert n_got == n_ref == [4000] * 4
        for t in range(1, 4):
            x, y = got[t]['next_observations'], ref[t]['next_observations']
            assert (np.abs(x - y) <= (t + 1) * (1e-2 * np.abs(y) + 1e-2 * s[:-1])).all(), t
    else:                                                      # antsafe: most rows terminate within a few steps
        assert n_got[0] == n_ref[0] and abs(n_got[1] - n_ref[1]) <= max(2, int(near.sum()))
        assert n_got[1] < 4000
