"""CPU oracle for the CMBPO model-rollout hot path.  TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this module.  The product (``constrainedmbpo_b200``)
never does: it fails loudly when its CUDA extension is missing.

PARITY UNPINNED: the reference (anyboby/ConstrainedMBPO) ships no tests, golden vectors or
fixtures for this path, and its model code needs TensorFlow 1.14, which cannot be imported
here.  What *is* pinned: the static termination / cost functions (``mbpo/static/*.py`` are
pure NumPy; ``tests/golden/static_fns.npz`` was produced by importing those files, see
``tests/golden/make_golden.py``) and ``average_dkl`` (same script executes the reference's own
source lines ``mbpo/models/utils.py:7-57``).  Everything else below is a restatement reviewed
line by line against the cited reference code; all citations are relative to /root/reference.

Arithmetic runs in the dtype of the parameters handed in: float64 as the arbiter, float32 to
mirror the reference's own precision (TF variables and placeholders are tf.float32,
``mbpo/models/bnn.py:210-213,345-347``).
"""
from __future__ import annotations

import numpy as np

EPS_KL = 1e-10  # mbpo/models/utils.py:7


# --------------------------------------------------------------------------------------
# activations: mbpo/models/fc.py:14-21
# --------------------------------------------------------------------------------------
def _sigmoid(x):
    return 1.0 / (1.0 + np.exp(-x))


def _softmax(x):
    z = x - np.max(x, axis=-1, keepdims=True)
    e = np.exp(z)
    return e / np.sum(e, axis=-1, keepdims=True)


ACTIVATIONS = {
    None: lambda x: x,
    "ReLU": lambda x: np.maximum(x, 0),
    "tanh": np.tanh,
    "sigmoid": _sigmoid,
    "softmax": _softmax,
    "swish": lambda x: x * _sigmoid(x),  # fc.py:20
}


# --------------------------------------------------------------------------------------
# ensemble parameters
# --------------------------------------------------------------------------------------
class EnsembleParams:
    """State the forward pass reads: per-layer weights ``[E,in,out]`` and biases
    ``[E,1,out]`` (fc.py:154-185), layer activations, the two scalers' ``mu``/``var``
    ``[1,dim]`` (utils.py:98-115; ``var`` is the TF variable *named* ``_std``), ``sc_factor``,
    whether the head is doubled into mean|logvar (bnn.py:111-112, :190-191), the output
    activation applied to the mean (bnn.py:194-196, :880-881) and the elite list
    (bnn.py:98-102, :428-431)."""

    def __init__(self, weights, biases, activations, mu_in=None, var_in=None, mu_out=None,
                 var_out=None, sc_factor=1.0, probabilistic=True, end_act=None, elites=None):
        self.weights = list(weights)
        self.biases = list(biases)
        self.activations = list(activations)
        self.mu_in, self.var_in = mu_in, var_in
        self.mu_out, self.var_out = mu_out, var_out
        self.sc_factor = sc_factor
        self.probabilistic = probabilistic
        self.end_act = end_act
        self.num_nets = self.weights[0].shape[0]
        self.elites = list(range(self.num_nets)) if elites is None else list(elites)

    @property
    def dtype(self):
        return self.weights[0].dtype

    @property
    def in_dim(self):
        return self.weights[0].shape[1]

    @property
    def out_dim(self):
        w = self.weights[-1].shape[2]
        return w // 2 if self.probabilistic else w

    def astype(self, dt):
        c = lambda a: None if a is None else np.asarray(a, dtype=dt)
        return EnsembleParams([c(w) for w in self.weights], [c(b) for b in self.biases],
                              self.activations, c(self.mu_in), c(self.var_in), c(self.mu_out),
                              c(self.var_out), self.sc_factor, self.probabilistic, self.end_act,
                              self.elites)


def truncated_normal(rng, shape, std):
    """tf.truncated_normal_initializer: N(0, std) re-drawn until |x| <= 2 std (fc.py:157)."""
    x = rng.standard_normal(shape)
    bad = np.abs(x) > 2.0
    while bad.any():
        x[bad] = rng.standard_normal(int(bad.sum()))
        bad = np.abs(x) > 2.0
    return x * std


def init_ensemble(rng, in_dim, out_dim, hidden_dims, num_networks, probabilistic=True,
                  activation="swish", output_activation=None, use_scaler_in=False,
                  use_scaler_out=False, sc_factor=1.0, logit_bias_std=1.0, elites=None,
                  dtype=np.float32):
    """Random ensemble of the shape ``construct_model`` builds (constructor.py:41-67):
    weights ~ TruncNormal(0, 1/(2 sqrt(in))) (last layer scaled by ``logit_bias_std``,
    bnn.py:216-219), zero biases (fc.py:180-184), identity scalers (utils.py:108-115).
    Distribution only -- TF's RNG stream is not reproduced."""
    dims = [in_dim] + list(hidden_dims) + [2 * out_dim if probabilistic else out_dim]
    acts = [activation] * len(hidden_dims) + [None]  # last activation stripped, bnn.py:194-196
    W, B = [], []
    for li in range(len(dims) - 1):
        std = 1.0 / (2.0 * np.sqrt(dims[li]))
        if li == len(dims) - 2:
            std *= logit_bias_std
        W.append(truncated_normal(rng, (num_networks, dims[li], dims[li + 1]), std).astype(dtype))
        B.append(np.zeros((num_networks, 1, dims[li + 1]), dtype=dtype))
    mk = lambda d, v: np.full((1, d), v, dtype=dtype)
    return EnsembleParams(
        W, B, acts,
        mu_in=mk(in_dim, 0.0) if use_scaler_in else None,
        var_in=mk(in_dim, 1.0) if use_scaler_in else None,
        mu_out=mk(out_dim, 0.0) if use_scaler_out else None,
        var_out=mk(out_dim, 1.0) if use_scaler_out else None,
        sc_factor=sc_factor, probabilistic=probabilistic, end_act=output_activation, elites=elites)


# --------------------------------------------------------------------------------------
# scalers: mbpo/models/utils.py:141-192
# --------------------------------------------------------------------------------------
def scaler_divisor(var, sc_factor):
    """max(sqrt(var) * sc_factor, 1e-2) -- utils.py:161, :172, :192 (quirk Q6)."""
    return np.maximum(np.sqrt(var) * sc_factor, np.asarray(1e-2, dtype=var.dtype))


def scaler_transform(x, mu, var, sc_factor):  # utils.py:141-162
    return (x - mu) / scaler_divisor(var, sc_factor)


def scaler_inverse(x, mu, var, sc_factor):  # utils.py:164-172
    return scaler_divisor(var, sc_factor) * x + mu


def scaler_inverse_logvar(x, var, sc_factor):  # utils.py:184-192
    return 2.0 * np.log(scaler_divisor(var, sc_factor)) + x


# --------------------------------------------------------------------------------------
# forward pass: fc.py:83-114 driven by bnn.py:852-903
# --------------------------------------------------------------------------------------
def fc_forward(x, W, b, activation):
    """2-D input: einsum('ij,ajk->aik') (fc.py:105-106); 3-D input with leading E:
    batched matmul (fc.py:107-108); anything else raises (fc.py:110-111)."""
    if x.ndim == 2:
        raw = np.matmul(x, W) + b  # == einsum("ij,ajk->aik"), BLAS-backed
    elif x.ndim == 3 and x.shape[0] == W.shape[0]:
        raw = np.matmul(x, W) + b
    else:
        raise ValueError("Invalid input dimension.")
    return ACTIVATIONS[activation](raw)


def compile_outputs(p: EnsembleParams, inputs, ret_log_var=False, logvar_bound=None):
    """bnn.py:852-903 with ``scale_output=True`` (the form create_prediction_tensors uses,
    bnn.py:781-782).  ``logvar_bound`` is None for the live reference (softplus bounding is
    commented out, bnn.py:888-889) or ``(max_logvar, min_logvar)`` for the optional PETS form
    (bnn_chua.py:400-401)."""
    dt = p.dtype
    cur = np.asarray(inputs, dtype=dt)
    if p.mu_in is not None:
        cur = scaler_transform(cur, p.mu_in, p.var_in, p.sc_factor)  # bnn.py:867-868
    for W, b, act in zip(p.weights, p.biases, p.activations):
        cur = fc_forward(cur, W, b, act)  # bnn.py:873-876
    dim_output = p.weights[-1].shape[2]
    mean = cur[..., : dim_output // 2] if p.probabilistic else cur  # bnn.py:878
    if p.end_act is not None:
        mean = ACTIVATIONS[p.end_act](mean)  # bnn.py:880-881
    if p.mu_out is not None:
        mean = scaler_inverse(mean, p.mu_out, p.var_out, p.sc_factor)  # bnn.py:883-884
    if not p.probabilistic:
        return mean
    logvar = cur[..., dim_output // 2:]  # bnn.py:891
    if logvar_bound is not None:  # bnn_chua.py:400-401 (off in the live reference)
        max_lv, min_lv = (np.asarray(v, dtype=dt) for v in logvar_bound)
        logvar = max_lv - np.logaddexp(0.0, max_lv - logvar)
        logvar = min_lv + np.logaddexp(0.0, logvar - min_lv)
    if p.mu_out is not None:
        logvar = scaler_inverse_logvar(logvar, p.var_out, p.sc_factor)  # bnn.py:892-893
    if ret_log_var:
        return mean, logvar
    with np.errstate(over="ignore"):
        return mean, np.exp(logvar)  # bnn.py:898


def predict(p: EnsembleParams, inputs, factored=False, inc_var=True, logvar_bound=None):
    """BNN.predict dispatch, bnn.py:671-776 (shape table: SURVEY appendix B)."""
    inputs = np.asarray(inputs)
    if inputs.ndim == 2:
        if p.probabilistic:
            mean_fac, var_fac = compile_outputs(p, inputs, logvar_bound=logvar_bound)
        else:
            mean_fac = compile_outputs(p, inputs)
        if factored:
            if inc_var and p.probabilistic:
                return mean_fac, var_fac  # :713-717
            if inc_var:
                rep = np.repeat(np.var(mean_fac, axis=0)[None], mean_fac.shape[0], axis=0)
                return mean_fac, rep  # :718-724
            return mean_fac  # :725-729
        if inc_var and p.probabilistic:
            m = np.mean(mean_fac, axis=0)  # bnn.py:350
            v = np.mean(var_fac, axis=0) + np.mean(np.square(mean_fac - m), axis=0)  # :351-352
            return m, v
        if inc_var:
            return np.mean(mean_fac, axis=0), np.var(mean_fac, axis=0)  # :736-743
        return np.mean(mean_fac, axis=0)  # :744-748
    # 3-D inputs, bnn.py:749-776
    if inc_var and p.probabilistic:
        return compile_outputs(p, inputs, logvar_bound=logvar_bound)  # :750-754
    if inc_var:
        mean3d = compile_outputs(p, inputs)
        flat = inputs.reshape(-1, inputs.shape[-1])  # :762-763
        mean2d = compile_outputs(p, flat)
        var3d = np.var(mean2d, axis=0).reshape(inputs.shape[0], inputs.shape[1], -1)  # :768-770
        return mean3d, var3d
    out = compile_outputs(p, inputs)
    return out[0] if p.probabilistic else out  # :772-776


# --------------------------------------------------------------------------------------
# ensemble KL: mbpo/models/utils.py:15-57
# --------------------------------------------------------------------------------------
def gaussian_kl(mu0, log_std0, mu1, log_std1):  # utils.py:15-25
    var0, var1 = np.exp(2 * log_std0), np.exp(2 * log_std1)
    pre_sum = 0.5 * (((mu1 - mu0) ** 2 + var0) / (var1 + EPS_KL) - 1) + log_std1 - log_std0
    return np.clip(pre_sum, 0, 1 / EPS_KL)


def average_dkl(mu, std):
    """utils.py:30-57: sum over ALL ordered pairs (i==j included, quirk Q5), divided by
    E(E-1)+EPS; ``log std`` clipped to [-100, 1e8]."""
    with np.errstate(divide="ignore", over="ignore", invalid="ignore"):
        log_std = np.clip(np.log(std), -100, 1e8)
        n = len(mu)
        d_kl = None
        for i in range(n):
            for j in range(n):
                t = gaussian_kl(mu[i], log_std[i], mu[j], log_std[j])
                d_kl = t if d_kl is None else d_kl + t
    return d_kl / (n * (n - 1) + EPS_KL)


# --------------------------------------------------------------------------------------
# static termination / cost functions: mbpo/static/*.py (quirks Q1, Q2 reproduced)
# --------------------------------------------------------------------------------------
def _antsafe_notdone(next_obs):  # antsafe.py:9-19 -- note the precedence: (a*b*c*z_rot) >= -0.7
    z = next_obs[..., 0]
    q = next_obs[..., 1:5]
    z_rot = 1 - 2 * (q[..., 1] ** 2 + q[..., 2] ** 2)
    with np.errstate(invalid="ignore", over="ignore"):
        return np.isfinite(next_obs).all(axis=-1) * (z >= 0.2) * (z <= 1.0) * z_rot >= -0.7


def term_antsafe(obs, act, next_obs):  # antsafe.py:6-22
    return (~_antsafe_notdone(next_obs))[..., None]


def cost_antsafe(obs, act, next_obs):  # antsafe.py:25-46
    y_dist = next_obs[..., -1:]
    obj_cost = np.any(abs(y_dist) > 3.2, axis=-1)[..., None] * 1.0
    done_cost = term_antsafe(obs, act, next_obs) * 1.0
    return np.clip(done_cost + obj_cost, 0, 1)


def term_ant(obs, act, next_obs):  # ant.py:6-16
    x = next_obs[..., 0]
    not_done = np.isfinite(next_obs).all(axis=-1) * (x >= 0.2) * (x <= 1.0)
    return (~not_done)[..., None]


def term_never(obs, act, next_obs):  # halfcheetah.py:6-11, halfcheetahsafe.py:6-11
    return np.zeros(obs.shape[:-1], dtype=bool)[..., None]


def cost_halfcheetahsafe(obs, act, next_obs):  # halfcheetahsafe.py:14-20
    xdist = next_obs[..., -1] * 10
    return np.array(np.abs(xdist) < 2.0, dtype=np.float32)[..., None]


def term_humanoid(obs, act, next_obs):  # humanoid.py:8-15, humanoidsafe.py:8-15
    z = next_obs[..., 0]
    return ((z < 1.0) + (z > 2.0))[..., None]


def cost_humanoidsafe(obs, act, next_obs):  # humanoidsafe.py:18-33
    done_cost = term_humanoid(obs, act, next_obs) * 1.0
    y_dist = next_obs[..., -1:]
    obj_cost = np.any(abs(y_dist) > 3, axis=-1)[..., None] * 1.0
    return np.clip(done_cost + obj_cost, 0, 1)


def term_hopper(obs, act, next_obs):  # hopper.py:6-18, hoppersafe.py:6-18 (Q2: abs of a bool)
    height = next_obs[..., 0]
    angle = next_obs[..., 1]
    not_done = np.isfinite(next_obs).all(axis=-1) \
        * np.abs(next_obs[..., 1:] < 100).all(axis=-1) \
        * (height > .7) * (np.abs(angle) < .2)
    return (~not_done)[..., None]


def cost_hoppersafe(obs, act, next_obs):  # hoppersafe.py:21-26
    xdist = next_obs[..., -1]
    return np.array(np.abs(xdist) < 1, dtype=np.float32)[..., None]


def term_walker2d(obs, act, next_obs):  # walker2d.py:6-17
    height = next_obs[..., 0]
    angle = next_obs[..., 1]
    not_done = (height > 0.8) * (height < 2.0) * (angle > -1.0) * (angle < 1.0)
    return (~not_done)[..., None]


STATIC_FNS = {  # name -> (termination_fn, cost_f or None); mbpo/static/__init__.py:7-24
    "ant": (term_ant, None),
    "antsafe": (term_antsafe, cost_antsafe),
    "halfcheetah": (term_never, None),
    "halfcheetahsafe": (term_never, cost_halfcheetahsafe),
    "humanoid": (term_humanoid, None),
    "humanoidsafe": (term_humanoid, cost_humanoidsafe),
    "hopper": (term_hopper, None),
    "hoppersafe": (term_hopper, cost_hoppersafe),
    "walker2d": (term_walker2d, None),
}


# --------------------------------------------------------------------------------------
# FakeEnv.step: mbpo/models/fake_env.py:131-222
# --------------------------------------------------------------------------------------
def fake_env_step(p: EnsembleParams, env: str, obs, act, elite_idx=None, perm=None,
                  cares_about_cost=False, cost_params: EnsembleParams | None = None,
                  logvar_bound=None, noise=None):
    """One rollout step.  RNG draws are injected (SURVEY §8b "Threading"): ``elite_idx`` is
    what ``BNN.random_inds(B)`` returned (values from the elite list, bnn.py:437-439), ``perm``
    the permutation ``forward_shuffle`` drew (fake_env.py:309-316; 3-D inputs only).

    ``noise`` (``[B,D]`` 2-D / ``[E,B,D]`` 3-D standard normals, default None) is the optional
    north_star mode ``sample = mean + noise*std``; the live reference adds none
    (fake_env.py:164-170) so None is the parity setting.
    """
    dt = p.dtype
    obs = np.asarray(obs, dtype=dt)
    act = np.asarray(act, dtype=dt)
    assert obs.ndim == act.ndim  # fake_env.py:132
    depth = obs.ndim
    assert depth in (2, 3), "1-D inputs hit a NameError in the reference (quirk Q3)"
    inputs = np.concatenate((obs, act), axis=-1)  # :145
    if depth == 3:
        inputs = inputs[perm]  # forward_shuffle :147-148
    means, variances = predict(p, inputs, factored=True, inc_var=True, logvar_bound=logvar_bound)
    means = np.array(means)
    if depth == 3:  # "inverse" shuffle re-applies perm (quirk Q4) :152-153, :318-324
        means, variances = means[perm], variances[perm]
    means[:, :, :-1] += obs  # :156
    with np.errstate(invalid="ignore"):
        stds = np.sqrt(variances)  # :157
    dkl_path = np.mean(average_dkl(means, stds), axis=-1)  # :158
    dkl_mean = np.mean(dkl_path)  # :159
    ep_var = np.var(means, axis=0)  # :160
    if depth == 3:
        samples = means  # :164-165
        if noise is not None:
            samples = means + np.asarray(noise, dtype=dt) * stds
    else:
        B = means.shape[1]
        samples = means[np.asarray(elite_idx), np.arange(B)]  # :167-170
        if noise is not None:
            samples = samples + np.asarray(noise, dtype=dt) * stds[np.asarray(elite_idx), np.arange(B)]
    next_obs = samples[..., :-1]  # :173
    rewards = samples[..., -1:]  # :180
    term_fn, cost_fn = STATIC_FNS[env]
    terminals = term_fn(obs, act, next_obs)  # :185
    if cares_about_cost:
        if cost_fn is not None:
            costs = cost_fn(obs, act, next_obs)  # :189
        elif cost_params is not None:
            inputs_cost = np.concatenate((obs, act, next_obs), axis=-1)  # :194
            costs = predict(cost_params, inputs_cost, factored=False, inc_var=False)  # :196
        else:
            raise ValueError("cares_about_cost without a static cost_f needs a cost model")
    else:
        costs = np.zeros_like(rewards)  # :202
    with np.errstate(over="ignore", invalid="ignore"):
        info = {
            "ensemble_dkl_mean": dkl_mean,
            "ensemble_dkl_path": dkl_path,
            "ensemble_mean_var": variances.mean(),
            "ensemble_ep_var": ep_var,
            "rew": rewards,
            "rew_mean": rewards.mean(),
            "cost": costs,
            "cost_mean": np.asarray(costs).mean(),
            # not part of the reference's dict; kept for tolerance analysis in the tests
            "_means": means,
            "_vars": variances,
        }
    return next_obs, rewards, terminals, info


# --------------------------------------------------------------------------------------
# rollout cutoff + model-pool layout:
#   softlearning/samplers/model_sampler.py:220-263, :265-474
#   softlearning/replay_pools/modelbuffer.py:45-173
# --------------------------------------------------------------------------------------
def rollout(p: EnsembleParams, env: str, start_obs, policy, max_path_length, rollout_mode,
            dkl_lim=None, max_samples=None, elite_idx_fn=None, cares_about_cost=False,
            max_steps=None, step_fn=None, trace=None):
    """Restatement of the loop ``CMBPO._train`` runs (cmbpo.py:335-348): ``sampler.reset`` then
    ``sampler.sample`` until no path is alive, restricted to what the hot path owns: the
    step, the DKL-accumulate cutoff (:311-313), the ``max_samples`` early termination
    (:327-332), the length mask (:437-438, quirk Q8), the terminal mask (:453-456) and the
    pool writes of ``store_multiple`` (modelbuffer.py:130-173).  Critic values, log-probs and
    GAE are out of scope and not modelled.

    ``policy(obs) -> act`` and ``elite_idx_fn(step, n_alive) -> int[n_alive]`` are injected.
    ``step_fn(cur, act, step) -> (next_obs, rew, term, info)`` replaces ``fake_env_step`` when
    given (teacher-forced bookkeeping tests replay recorded step outputs through it; the CPU
    baseline plugs its BLAS-backed step in).  ``trace``: optional list that receives the number of
    alive rows at the start of every sampler step.
    2-D modes only ('schedule' / 'uncertainty'); returns the pool arrays.
    """
    assert rollout_mode in ("schedule", "uncertainty")
    dt = p.dtype
    start_obs = np.asarray(start_obs, dtype=dt)
    B, obs_dim = start_obs.shape
    T = max_path_length
    act_dim = None
    pool = {
        "obs": np.zeros((B, T, obs_dim), np.float32),
        "next_obs": np.zeros((B, T, obs_dim), np.float32),
        "rew": np.zeros((B, T), np.float32),
        "cost": np.zeros((B, T), np.float32),
        "term": np.zeros((B, T), bool),
        "dyn_error": np.zeros((B, T), np.float32),  # mean_d(ep_var), model_sampler.py:428
        "populated": np.zeros((B, T), bool),
    }
    terminated = np.zeros(B, bool)  # modelbuffer.py:114 terminated_paths_mask
    ptr = 0
    cur = start_obs
    path_length = np.zeros(B)
    dkl_accum = np.zeros(B)  # _dyn_dkl_path, model_sampler.py:240
    total_samples = 0
    step = 0
    while (~terminated).any() and (max_steps is None or step < max_steps):
        alive = ~terminated
        n_alive = int(alive.sum())
        a = np.asarray(policy(cur), dtype=dt)
        if act_dim is None:
            act_dim = a.shape[-1]
            pool["act"] = np.zeros((B, T, act_dim), np.float32)
        if trace is not None:
            trace.append(n_alive)
        if step_fn is not None:
            next_obs, rew, term, info = step_fn(cur, a, step)
        else:
            eidx = elite_idx_fn(step, n_alive)
            next_obs, rew, term, info = fake_env_step(p, env, cur, a, elite_idx=eidx,
                                                      cares_about_cost=cares_about_cost)
        rew = np.squeeze(rew, -1)
        c = np.squeeze(info["cost"], -1)
        term = np.squeeze(term, -1)
        dkl_path = info["ensemble_dkl_path"]
        ep_var = info["ensemble_ep_var"]
        if rollout_mode == "uncertainty":  # :311-313
            too_uncertain = dkl_accum[alive] + dkl_path >= dkl_lim
        else:
            too_uncertain = np.zeros(n_alive, bool)  # :315
        if max_samples:  # :327-332
            n = total_samples + n_alive - int(too_uncertain.sum())
            n = max(n - max_samples, 0)
            early = np.zeros(int((~too_uncertain).sum()), bool)
            early[:n] = True
            too_uncertain[~too_uncertain] = early
        # _finish_paths(too_uncertain): pool.finish_path_multiple marks them terminated
        idx_alive = np.where(alive)[0]
        terminated[idx_alive[too_uncertain]] = True
        remaining = ~too_uncertain
        alive = ~terminated
        if not alive.any():  # :339-342
            break
        cur_r, a_r, nobs_r = cur[remaining], a[remaining], next_obs[remaining]
        rew_r, c_r, term_r = rew[remaining], c[remaining], term[remaining]
        dkl_r, epv_r = dkl_path[remaining], ep_var[remaining]
        total_samples += int(alive.sum())  # :386
        path_length[alive] += 1  # :400
        dkl_accum[alive] += dkl_r  # :415
        # store_multiple, modelbuffer.py:130-173
        pool["obs"][alive, ptr] = cur_r
        pool["act"][alive, ptr] = a_r
        pool["next_obs"][alive, ptr] = nobs_r
        pool["rew"][alive, ptr] = rew_r
        pool["cost"][alive, ptr] = c_r
        pool["term"][alive, ptr] = term_r
        pool["dyn_error"][alive, ptr] = np.mean(epv_r, axis=-1)
        pool["populated"][alive, ptr] = True
        ptr += 1
        cur = nobs_r  # :435
        path_end = (path_length >= max_path_length - 1)[alive]  # :437 (Q8)
        idx_alive = np.where(alive)[0]
        terminated[idx_alive[path_end]] = True
        rem2 = ~path_end
        if not rem2.any():  # :439-441
            break
        cur = cur[rem2]
        prem_term = term_r  # :452 (all alive paths share one length, so rem2 is all-True here)
        alive = ~terminated
        idx_alive = np.where(alive)[0]
        terminated[idx_alive[prem_term]] = True  # :456
        rem3 = ~prem_term
        if not rem3.any():
            break
        cur = cur[rem3]  # :465
        step += 1
    pool["ptr"] = ptr
    pool["total_samples"] = total_samples
    pool["terminated"] = terminated
    pool["dkl_accum"] = dkl_accum
    pool["path_length"] = path_length
    return pool


def rollout_iv_gae(p: EnsembleParams, env: str, start_obs, policy, max_path_length, perm_fn=None,
                   cares_about_cost=False, max_steps=None, step_fn=None, trace=None):
    """The same sampler loop in ``rollout_mode == 'iv_gae'`` (model_sampler.py:225-226, :283-288, :348-362,
    :388-392, :444-445, :463-464; modelbuffer.py:48-51, :134-150): the start states are tiled over the
    ensemble axis and every member rolls its own trajectory through the 3-D branch of ``FakeEnv.step``
    (fake_env.py:147-153, :164-165).  No uncertainty cutoff (:315) and no ``max_samples`` cap (:327) in
    this mode; a path ends at the length mask (:437, Q8) or when ANY member's state terminates (:445).
    Per-transition pools carry a leading ``E``; ``dyn_error`` / ``populated`` stay ``[B, T]``.

    ``policy(obs[E,n,obs]) -> act[E,n,act]``; ``perm_fn(step) -> permutation of range(E)`` is the injected
    ``forward_shuffle`` draw.  The inverse-variance GAE that consumes these pools is SURVEY §8f N2.
    """
    dt = p.dtype
    start_obs = np.asarray(start_obs, dtype=dt)
    B, obs_dim = start_obs.shape
    E = len(p.weights[0])
    T = max_path_length
    pool = {
        "obs": np.zeros((E, B, T, obs_dim), np.float32),
        "next_obs": np.zeros((E, B, T, obs_dim), np.float32),
        "rew": np.zeros((E, B, T), np.float32),
        "cost": np.zeros((E, B, T), np.float32),
        "term": np.zeros((E, B, T), bool),
        "dyn_error": np.zeros((B, T), np.float32),
        "populated": np.zeros((B, T), bool),
    }
    terminated = np.zeros(B, bool)
    ptr = 0
    cur = np.tile(start_obs[None], (E, 1, 1))  # :226
    path_length = np.zeros(B)
    dkl_accum = np.zeros(B)
    total_samples = 0
    step = 0
    while (~terminated).any() and (max_steps is None or step < max_steps):
        alive = ~terminated
        n_alive = int(alive.sum())
        a = np.asarray(policy(cur), dtype=dt)
        if "act" not in pool:
            pool["act"] = np.zeros((E, B, T, a.shape[-1]), np.float32)
        if trace is not None:
            trace.append(n_alive)
        if step_fn is not None:
            next_obs, rew, term, info = step_fn(cur, a, step)
        else:
            next_obs, rew, term, info = fake_env_step(p, env, cur, a, perm=np.asarray(perm_fn(step)),
                                                      cares_about_cost=cares_about_cost)
        rew = np.squeeze(rew, -1)            # [E, n]
        c = np.squeeze(info["cost"], -1)     # [E, n]
        term = np.squeeze(term, -1)          # [E, n]
        dkl_path = info["ensemble_dkl_path"]
        ep_var = info["ensemble_ep_var"]
        # too_uncertain is all-False (:315), max_samples is skipped (:327) => every alive path is stored
        total_samples += n_alive  # :386
        path_length[alive] += 1  # :400
        dkl_accum[alive] += dkl_path  # :415
        pool["obs"][:, alive, ptr] = cur  # modelbuffer.py:134-150
        pool["act"][:, alive, ptr] = a
        pool["next_obs"][:, alive, ptr] = next_obs
        pool["rew"][:, alive, ptr] = rew
        pool["cost"][:, alive, ptr] = c
        pool["term"][:, alive, ptr] = term
        pool["dyn_error"][alive, ptr] = np.mean(ep_var, axis=-1)  # :168
        pool["populated"][alive, ptr] = True
        ptr += 1
        cur = next_obs  # :435
        path_end = (path_length >= max_path_length - 1)[alive]  # :437
        idx_alive = np.where(alive)[0]
        terminated[idx_alive[path_end]] = True
        rem2 = ~path_end
        if not rem2.any():
            break
        cur = cur[:, rem2]  # :444
        prem_term = np.any(term[:, rem2], axis=0)  # :445
        idx_alive = np.where(~terminated)[0]
        terminated[idx_alive[prem_term]] = True
        if prem_term.all():
            break
        cur = cur[:, ~prem_term]  # :464
        step += 1
    pool["ptr"] = ptr
    pool["total_samples"] = total_samples
    pool["terminated"] = terminated
    pool["dkl_accum"] = dkl_accum
    pool["path_length"] = path_length
    return pool


# --------------------------------------------------------------------------------------
# ModelBuffer.finish_path_multiple: GAE and inverse-variance GAE
#   softlearning/replay_pools/modelbuffer.py:176-346
#   softlearning/policies/safe_utils/utils.py:18-188 (discount_cumsum, discount_cumsum_weighted,
#   disc_cumsum_matrix, triu_indices_t_h)
# --------------------------------------------------------------------------------------
GAE_EPS = 1e-8  # EPS of safe_utils/utils.py


def _iv_gae_path(x, v_ext, m, gamma_mat, gamma, lam):
    """One path, one signal (reward or cost).  ``x`` [E, L] per-member signal, ``v_ext`` [E, L+1] per-member
    values with the bootstrap value appended, ``m`` the member whose trajectory supplies the deltas
    (modelbuffer.py:210), ``gamma_mat`` the discount of the partial-sum matrix (the reference passes
    ``self.gamma`` for BOTH signals, modelbuffer.py:225,:286), ``gamma``/``lam`` the signal's own discounts.
    Returns (adv[L], ret_var[L], roll_lengths[L]) in float64."""
    E, L = x.shape
    x = x.astype(np.float64)
    v_ext = v_ext.astype(np.float64)
    # deltas of the chosen member, in the buffers' float32 (modelbuffer.py:210: float32 arrays and Python floats)
    xm, vm = x[m].astype(np.float32), v_ext[m].astype(np.float32)
    deltas = (xm + np.float32(gamma) * vm[1:] - vm[:-1]).astype(np.float64)
    var = np.zeros((L, L))
    for t in range(L):
        for h in range(L - t):
            # disc_cumsum_matrix entry (utils.py:128-180) + bootstrapped value (modelbuffer.py:226)
            s = np.zeros(E)
            for j in range(h + 1):
                s += gamma_mat ** j * x[:, t + j]
            s += v_ext[:, t + h + 1] * gamma ** (h + 1)
            var[t, h] = np.var(s)                                   # :227 epistemic variance over members
    w = np.zeros((L, L))
    for t in range(L):
        for h in range(L - t):
            w[t, h] = 1.0 / (var[t, h] + GAE_EPS) * (lam * gamma ** 2) ** h     # :230-233
        w[t, L - 1 - t] *= 1.0 / (1.0 - lam * gamma ** 2 + GAE_EPS)              # :234 last h of every row
    adv, ret_var, roll = np.zeros(L), np.zeros(L), np.zeros(L)
    for t in range(L):
        dw = np.cumsum(w[t, ::-1])[::-1]                                        # :237 sum from h to the end
        norm = 1.0 / dw[0]                                                       # :238
        for h in range(L - t):
            adv[t] += gamma ** h * deltas[t + h] * dw[h] * norm                  # :249 discount_cumsum_weighted
            ret_var[t] += w[t, h] * norm * var[t, h]                             # :245,:251-252
            roll[t] += (t + h) * w[t, h]                                         # :254-255
        roll[t] = roll[t] * norm - t
    return adv, ret_var, roll


def finish_paths(rew, val, cost, cval, finish_ptr, last_val, last_cval, mode, gamma, lam, cost_gamma, cost_lam,
                 model_ind=0, cares_about_cost=True, max_uncertainty_r=5.5, max_uncertainty_c=5.5, populated=None):
    """``ModelBuffer.finish_path_multiple`` for every path at once.  Path ``b`` was finished when the pool column
    was ``finish_ptr[b]`` (its slice is ``[0:finish_ptr[b]]``, modelbuffer.py:199) with bootstrap values
    ``last_val[..., b]`` / ``last_cval[..., b]``.  Arrays are ``[B, T]`` ('schedule' / 'uncertainty') or
    ``[E, B, T]`` ('iv_gae'); outputs are float32 ``[B, T]`` like the reference's buffers.  In 'iv_gae' mode the
    populated mask is trimmed at the first column whose normalised return variance exceeds the limits
    (:329-343) -- including the reference's quirk that a path with NO such column gets horizon
    ``argmax(all False) == 0`` and is dropped entirely."""
    iv = mode == "iv_gae"
    B, T = rew.shape[-2:]
    out = {k: np.zeros((B, T), np.float32) for k in ("adv", "ret", "cadv", "cret", "ret_var", "cret_var",
                                                      "roll_lengths", "croll_lengths")}
    pop = None if populated is None else np.array(populated, bool)
    for b in range(B):
        L = int(finish_ptr[b])
        if L <= 0:   # never stored anything (modelbuffer.py:198 `if self.ptr>0`)
            continue
        if not iv:
            for sig, v, lv, g, l, ka, kr in ((rew, val, last_val, gamma, lam, "adv", "ret"),
                                            (cost, cval, last_cval, cost_gamma, cost_lam, "cadv", "cret")):
                x_ext = np.append(sig[b, :L], lv[b]).astype(np.float32)     # :201 (rews also gets last_val appended)
                v_ext = np.append(v[b, :L], lv[b]).astype(np.float32)
                deltas = x_ext[:-1] + np.float32(g) * v_ext[1:] - v_ext[:-1]   # float32, :264
                a = np.zeros(L)
                run = 0.0
                for t in range(L - 1, -1, -1):                              # discount_cumsum (utils.py:44-47)
                    run = float(deltas[t]) + g * l * run
                    a[t] = run
                out[ka][b, :L] = a
                out[kr][b, :L] = out[ka][b, :L] + v[b, :L]                   # :268 float32 + float32
            continue
        v_ext = np.concatenate((val[:, b, :L], last_val[:, b, None]), axis=-1)
        a, rv, rl = _iv_gae_path(rew[:, b, :L], v_ext, model_ind, gamma, gamma, lam)
        out["adv"][b, :L], out["ret_var"][b, :L], out["roll_lengths"][b, :L] = a, rv, rl
        out["ret"][b, :L] = out["adv"][b, :L] + val[model_ind, b, :L]
        cv_ext = np.concatenate((cval[:, b, :L], last_cval[:, b, None]), axis=-1)
        a, rv, rl = _iv_gae_path(cost[:, b, :L], cv_ext, model_ind, gamma, cost_gamma, cost_lam)
        out["cadv"][b, :L], out["cret_var"][b, :L], out["croll_lengths"][b, :L] = a, rv, rl
        out["cret"][b, :L] = out["cadv"][b, :L] + cval[model_ind, b, :L]
        if pop is not None:
            # :329-343, evaluated on the float32 buffers
            ncv = out["cret_var"][b, :L] / (out["cret_var"][b, 0:1] + GAE_EPS)
            nrv = out["ret_var"][b, :L] / (out["ret_var"][b, 0:1] + GAE_EPS)
            mask = nrv > max_uncertainty_r
            if cares_about_cost:
                mask = np.logical_or(ncv > max_uncertainty_c, mask)
            horizon = int(np.argmax(mask))
            pop[b, :] &= np.arange(T) < horizon
    out["populated"] = pop
    return out


def normalize_advantages(adv, cadv, populated):
    """``ModelBuffer.get``'s advantage trick (modelbuffer.py:356-372) with ``mpi_statistics_scalar`` at world size 1
    (safe_utils/mpi_tools.py:71-93): float32 mean / population std over the populated entries; ``adv`` standardised,
    ``cadv`` centred only.  Returns (adv, cadv, adv_mean, adv_std, cadv_mean) without touching the inputs."""
    adv, cadv = np.array(adv, np.float32), np.array(cadv, np.float32)
    pop = np.asarray(populated, bool)
    if pop.sum() == 0:
        return adv, cadv, 0.0, 0.0, 0.0

    def stats(x):
        x = np.array(x, dtype=np.float32)
        s, n = np.asarray([np.sum(x), len(x)], dtype=np.float32)
        mean = s / n
        ss = np.asarray(np.sum((x - mean) ** 2), dtype=np.float32)
        return mean, np.sqrt(ss / n)
    am, asd = stats(adv[pop].flatten())
    adv[pop] = (adv[pop] - am) / (asd + GAE_EPS)
    cm, _ = stats(cadv[pop].flatten())
    cadv[pop] -= cm
    return adv, cadv, am, asd, cm


# --------------------------------------------------------------------------------------
# Start-state sampling: softlearning/replay_pools/cpobuffer.py:531-542, :559-610
# --------------------------------------------------------------------------------------
def boltz_dist(epoch_archive, kls, alpha=1):
    """``CPOBuffer.boltz_dist``: Boltzmann weights over epochs, spread evenly over each epoch's rows (float32)."""
    epoch_archive = np.asarray(epoch_archive)
    ep_probs = np.exp(alpha * np.negative(kls))
    ep_probs /= np.sum(ep_probs)
    sample_p = np.bincount(epoch_archive[epoch_archive >= 0]).astype(np.float32)
    sample_p[sample_p > 0] = ep_probs / sample_p[sample_p > 0]
    return np.where(epoch_archive >= 0, sample_p[epoch_archive], 0)


def choice_indices(dist, uniforms):
    """What ``np.random.choice(len(dist), size, p=dist)`` does with its uniforms (legacy RandomState.choice):
    cdf = cumsum(float64 p), normalised by its last entry; index = searchsorted(cdf, u, side='right')."""
    cdf = np.asarray(dist, dtype=np.float64).cumsum()
    cdf /= cdf[-1]
    return cdf.searchsorted(np.asarray(uniforms, np.float64), side="right")


# --------------------------------------------------------------------------------------
# Gaussian actor head: softlearning/models/ac_network.py:27-35 (mlp = tf.layers.dense stack),
# :50-52 (gaussian_likelihood, EPS = 1e-8 at :8), :105-117 (mlp_gaussian_policy), :129-146 (3-D: one draw tiled over E)
# --------------------------------------------------------------------------------------
def actor_outs(weights, biases, log_std, obs, noise, activation="tanh"):
    """``pi``, ``logp_pi``, ``mu`` of ``mlp_gaussian_policy`` for ``obs`` [B, obs] or [E, B, obs]; ``weights`` [in, out] and
    ``biases`` [out] per dense layer, ``noise`` [B, act] standard normals (the same row of it for every slot of a 3-D obs)."""
    x = np.asarray(obs)
    act = ACTIVATIONS[activation]
    for i, (W, b) in enumerate(zip(weights, biases)):
        x = x @ np.asarray(W, x.dtype) + np.asarray(b, x.dtype)
        if i < len(weights) - 1:
            x = act(x)
    mu = x
    ls = np.asarray(log_std, x.dtype)
    std = np.exp(ls)
    pi = mu + np.asarray(noise, x.dtype) * std                    # broadcasts [B, act] over a leading E
    pre = -0.5 * (((pi - mu) / (std + 1e-8)) ** 2 + 2 * ls + np.log(2 * np.pi))
    return pi, pre.sum(-1), mu
