"""Timed CPU baseline of the rollout path.  TEST / MEASUREMENT INFRASTRUCTURE ONLY.

kind = "port": the reference's own implementation needs TensorFlow 1.14 (absent here and on the
GPU box), so this module restates it op for op in fp32 with the same library roles: the ensemble
GEMMs run as ``torch.einsum('ij,ajk->aik')`` / ``torch.baddbmm`` on the host's BLAS threads (the
role TF's Eigen contraction plays at mbpo/models/fc.py:105-108), everything after ``sess.run``
is the NumPy code of ``oracle.cmbpo_oracle`` (49-pass ``average_dkl`` loop, fancy-index elite
gather, static termination / cost functions, boolean-mask compaction of the sampler).

Only ``bench.py`` (``cpu_baseline`` leg, ``--impl reference``) and ``tests/`` import this file.
"""
from __future__ import annotations

import os
import time

import numpy as np
import torch

from . import cmbpo_oracle as orc

ROW_CHUNK = 65536  # rows per GEMM chunk so the [E,B,H] fp32 activations stay ~0.4 GB


class TorchEnsemble:
    """fp32 forward of ``BNN._compile_outputs`` (bnn.py:852-903) on CPU threads."""

    def __init__(self, p: orc.EnsembleParams, threads=None):
        self.p = p.astype(np.float32)
        self.threads = threads or os.cpu_count() or 1
        torch.set_num_threads(self.threads)
        self.W = [torch.from_numpy(np.ascontiguousarray(w)) for w in self.p.weights]
        self.b = [torch.from_numpy(np.ascontiguousarray(b)) for b in self.p.biases]
        q = self.p
        if q.mu_in is not None:
            self.mu_in = torch.from_numpy(q.mu_in)
            self.s_in = torch.from_numpy(orc.scaler_divisor(q.var_in, np.float32(q.sc_factor)).astype(np.float32))
        else:
            self.mu_in = None
        if q.mu_out is not None:
            self.mu_out = torch.from_numpy(q.mu_out)
            self.s_out = torch.from_numpy(orc.scaler_divisor(q.var_out, np.float32(q.sc_factor)).astype(np.float32))
        else:
            self.mu_out = None

    @staticmethod
    def _act(name, x):
        if name == "swish":
            return x * torch.sigmoid(x)  # fc.py:20
        if name == "ReLU":
            return torch.relu(x)
        if name == "tanh":
            return torch.tanh(x)
        if name == "sigmoid":
            return torch.sigmoid(x)
        return x

    def _forward_chunk(self, x):
        cur = x
        if self.mu_in is not None:
            cur = (cur - self.mu_in) / self.s_in  # utils.py:161
        for i, (W, b, act) in enumerate(zip(self.W, self.b, self.p.activations)):
            if cur.dim() == 2:
                cur = torch.einsum("ij,ajk->aik", cur, W) + b  # fc.py:105-106
            else:
                cur = torch.baddbmm(b, cur, W)  # fc.py:107-108
            cur = self._act(act, cur)
        D = self.p.out_dim
        if not self.p.probabilistic:
            return (cur * self.s_out + self.mu_out if self.mu_out is not None else cur), None
        mean, logvar = cur[..., :D], cur[..., D:]
        if self.mu_out is not None:
            mean = mean * self.s_out + self.mu_out  # utils.py:172
            logvar = logvar + 2.0 * torch.log(self.s_out)  # utils.py:192
        return mean, torch.exp(logvar)  # bnn.py:898

    def predict_factored(self, inputs: np.ndarray):
        """[B,in] or [E,B,in] fp32 -> (mean[E,B,D], var[E,B,D]) NumPy fp32 (bnn.py:711-717, :749-754)."""
        x = torch.from_numpy(np.ascontiguousarray(inputs, dtype=np.float32))
        B = x.shape[-2]
        means, vars_ = [], []
        with torch.no_grad():
            for r0 in range(0, B, ROW_CHUNK):
                m, v = self._forward_chunk(x[..., r0:r0 + ROW_CHUNK, :])
                means.append(m)
                vars_.append(v)
        mean = torch.cat(means, dim=1).numpy()
        var = torch.cat(vars_, dim=1).numpy() if vars_[0] is not None else None
        return mean, var


def fake_env_step(ens: TorchEnsemble, env: str, obs, act, elite_idx, cares_about_cost=False, perm=None):
    """FakeEnv.step (fake_env.py:131-222), fp32: GEMMs in torch, the rest in NumPy exactly as
    ``oracle.cmbpo_oracle.fake_env_step`` does it.  3-D inputs ([E,B,.], 'iv_gae') take ``perm`` (the
    forward_shuffle draw) and keep every member's own sample (fake_env.py:147-153, :164-165)."""
    obs = np.asarray(obs, dtype=np.float32)
    act = np.asarray(act, dtype=np.float32)
    inputs = np.concatenate((obs, act), axis=-1)
    if obs.ndim == 3:
        inputs = inputs[perm]
    means, variances = ens.predict_factored(inputs)
    means = np.array(means)
    if obs.ndim == 3:
        means, variances = means[perm], variances[perm]
    means[:, :, :-1] += obs
    stds = np.sqrt(variances)
    dkl_path = np.mean(orc.average_dkl(means, stds), axis=-1)
    ep_var = np.var(means, axis=0)
    B = means.shape[1]
    samples = means if obs.ndim == 3 else means[np.asarray(elite_idx), np.arange(B)]
    next_obs, rewards = samples[..., :-1], samples[..., -1:]
    term_fn, cost_fn = orc.STATIC_FNS[env]
    terminals = term_fn(obs, act, next_obs)
    if cares_about_cost and cost_fn is not None:
        costs = cost_fn(obs, act, next_obs)
    else:
        costs = np.zeros_like(rewards)
    info = {"ensemble_dkl_mean": np.mean(dkl_path), "ensemble_dkl_path": dkl_path,
            "ensemble_mean_var": variances.mean(), "ensemble_ep_var": ep_var, "rew": rewards,
            "rew_mean": rewards.mean(), "cost": costs, "cost_mean": costs.mean()}
    return next_obs, rewards, terminals, info


def rollout(ens: TorchEnsemble, env, start_obs, policy_np, max_path_length, rollout_mode, dkl_lim=None,
            elite_rng=None, cares_about_cost=False):
    """The sampler loop of ``oracle.cmbpo_oracle.rollout`` with the torch-backed step; returns
    (total transitions stored, seconds)."""
    rng = elite_rng or np.random.default_rng(5)
    if rollout_mode == "iv_gae":
        E = ens.p.num_nets
        step_fn = lambda cur, a, step: fake_env_step(ens, env, cur, a, None, cares_about_cost, perm=rng.permutation(E))
        t0 = time.perf_counter()
        pool = orc.rollout_iv_gae(ens.p, env, start_obs, policy_np, max_path_length, step_fn=step_fn,
                                  cares_about_cost=cares_about_cost)
        return int(pool["total_samples"]), time.perf_counter() - t0, pool
    step_fn = lambda cur, a, step: _step_tuple(ens, env, cur, a, rng, cares_about_cost)
    t0 = time.perf_counter()
    pool = orc.rollout(ens.p, env, start_obs, policy_np, max_path_length, rollout_mode, dkl_lim=dkl_lim,
                       step_fn=step_fn, cares_about_cost=cares_about_cost)
    return int(pool["total_samples"]), time.perf_counter() - t0, pool


def _step_tuple(ens, env, cur, a, rng, cares_about_cost):
    idx = rng.choice(ens.p.elites, size=cur.shape[0])  # bnn.py:437-439
    no, rw, tm, info = fake_env_step(ens, env, cur, a, idx, cares_about_cost)
    return no, rw, tm, info


def compute_dynamics_dkl(ens: TorchEnsemble, env, obs_batch, policy_np, depth, rng=None, elite_fn=None, totals=None):
    """model_sampler.py:160-176 on the CPU port.  ``totals`` = [total_dkl, total_samples] carried across calls like the
    sampler's ``_total_dkl`` / ``_total_samples`` (:51,:61); ``elite_fn(step, n)`` injects the ``random_inds`` draws."""
    rng = rng or np.random.default_rng(6)
    obs = np.asarray(obs_batch, dtype=np.float32)
    tot = totals if totals is not None else [0.0, 0]
    for k in range(depth):
        if obs.shape[0] == 0:
            break
        a = policy_np(obs)
        idx = elite_fn(k, obs.shape[0]) if elite_fn is not None else rng.choice(ens.p.elites, size=obs.shape[0])
        no, _, tm, info = fake_env_step(ens, env, obs, a, idx)
        n_paths = no.shape[0]                                       # :167
        tot[0] += float(info["ensemble_dkl_mean"]) * n_paths        # :168
        tot[1] += n_paths                                           # :169
        obs = no[~np.squeeze(tm, -1)]                               # :171
    return tot[0] / (tot[1] + 1e-8) * depth                         # :173-174, dyn_dkl :192-194 (EPS = 1e-8, :17)
