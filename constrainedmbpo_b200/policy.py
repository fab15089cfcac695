"""The CPO actor inside the device rollout loop.

``CPOPolicy.get_action_outs`` (softlearning/policies/cpo_policy.py:883-925) is called by ``ModelSampler.sample`` every step
right before ``env.step`` (softlearning/samplers/model_sampler.py:275-294).  Its actor is ``mlp_gaussian_policy``
(softlearning/models/ac_network.py:105-149): a tanh MLP (``a_hidden_layer_sizes``, 256 x 256 in examples/development/
base.py:30) producing ``mu``, a state-independent ``log_std`` variable (initialised to -0.5), ``pi = mu + N(0,1) * std`` and
``logp_pi = gaussian_likelihood(pi, mu, log_std)``.  Here the MLP is a 1-member non-probabilistic ensemble on the forward
kernels of libcmbpo_b200.so (``tf.layers.dense`` == one FC layer), the head is ``cmbpo_actor_head``.  Training stays with
the reference: hand the trained variables over with ``load_state``.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib
from .constructor import construct_model


class GaussianActor:
    """``policy(obs, n_rows_dev)`` for ``ModelRollout`` and a ``get_action_outs`` with the reference's keys
    (``pi``, ``logp_pi``, ``pi_info['mu' | 'log_std']``).  ``obs`` [B, obs] or, for 'iv_gae' states, [E, B, obs] (the
    flattened batch goes through the net, one noise draw per row is shared by the E slots, ac_network.py:129-146)."""

    def __init__(self, obs_dim, act_dim, hidden_sizes=(256, 256), device="cuda:0", seed=None, log_std_init=-0.5):
        self.obs_dim, self.act_dim = int(obs_dim), int(act_dim)
        self.device = torch.device(device)
        self.net = construct_model(in_dim=obs_dim, out_dim=act_dim, name="pi", hidden_dims=tuple(hidden_sizes),
                                   num_networks=1, num_elites=1, loss="MSE", activation="tanh", output_activation=None,
                                   use_scaler_in=False, use_scaler_out=False, device=device, seed=seed)
        self.log_std = torch.full((act_dim,), float(log_std_init), dtype=torch.float32, device=self.device)   # :110
        self.deterministic = False
        self._seed = int(seed) if seed is not None else int(np.random.randint(2 ** 31 - 1))
        self._rng_offset = 0
        self._lib = _lib.load()
        self._buf = {}
        self.last = None

    def load_state(self, weights, biases, log_std):
        """``tf.layers.dense`` kernels [in, out] and biases [out] of the 'pi' scope, in layer order, + ``log_std`` [act]."""
        ws = [np.asarray(w, np.float32)[None] for w in weights]
        bs = [np.asarray(b, np.float32).reshape(1, 1, -1) for b in biases]
        self.net.load_state(ws, bs)
        self.log_std = torch.as_tensor(np.asarray(log_std, np.float32)).to(self.device).contiguous()

    def _out(self, key, shape):
        t = self._buf.get(key)
        if t is None or tuple(t.shape) != tuple(shape):
            t = self._buf[key] = torch.empty(shape, dtype=torch.float32, device=self.device)
        return t

    def get_action_outs_device(self, obs, n_rows_dev=None, noise=None, row_ids=None, rng_offset_dev=None):
        obs = obs.contiguous()
        lead = tuple(obs.shape[:-1])                       # (B,) or (E, B)
        B = lead[-1]
        S = lead[0] if len(lead) == 2 else 1
        flat = obs.reshape(-1, self.obs_dim)
        # 3-D states: every slot holds `capacity` rows of which the first n are live -> the flattened batch is not a prefix;
        # the net then runs on all S * B rows (the head below honours the live count)
        mu = self.net.predict_device([flat], 2, 0, want_var=False, n_rows_dev=n_rows_dev if S == 1 else None)[0][0]
        pi, logp = self._out("pi", lead + (self.act_dim,)), self._out("logp", lead)
        io = _lib.cmbpo_actor_io()
        io.mu_dev, io.log_std_dev = mu.data_ptr(), self.log_std.data_ptr()
        if noise is not None:
            noise = noise.to(self.device, torch.float32).contiguous()
            io.noise_dev = noise.data_ptr()
        io.n_rows, io.act_dim, io.n_slots = B, self.act_dim, S
        io.n_rows_dev = n_rows_dev.data_ptr() if n_rows_dev is not None else None
        io.deterministic = 1 if self.deterministic else 0
        io.row_id_dev = row_ids.data_ptr() if row_ids is not None else None
        io.seed, io.rng_offset = self._seed, self._rng_offset
        self._rng_offset += 1
        io.rng_offset_dev = rng_offset_dev.data_ptr() if rng_offset_dev is not None else None
        io.pi_dev, io.logp_pi_dev = pi.data_ptr(), logp.data_ptr()
        stream = C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        _lib.check(self._lib.cmbpo_actor_head(C.byref(io), stream))
        self._keep = (mu, noise)
        self.last = {"pi": pi, "logp_pi": logp, "pi_info": {"mu": mu.reshape(lead + (self.act_dim,)),
                                                             "log_std": self.log_std.expand(lead + (self.act_dim,))}}
        return self.last

    takes_rollout_context = True     # ModelRollout passes the global row ids and its device Philox-stream word

    def __call__(self, obs, n_rows_dev=None, row_ids=None, rng_offset_dev=None):
        return self.get_action_outs_device(obs, n_rows_dev=n_rows_dev, row_ids=row_ids, rng_offset_dev=rng_offset_dev)["pi"]

    def get_action_outs(self, obs, noise=None):
        """NumPy in -> NumPy out, the dict ``ModelSampler.sample`` reads (model_sampler.py:275-288)."""
        o = torch.as_tensor(np.asarray(obs, np.float32)).to(self.device)
        nz = None if noise is None else torch.as_tensor(np.asarray(noise, np.float32))
        r = self.get_action_outs_device(o, noise=nz)
        torch.cuda.synchronize(self.device)
        return {"pi": r["pi"].cpu().numpy(), "logp_pi": r["logp_pi"].cpu().numpy(),
                "pi_info": {k: v.cpu().numpy() for k, v in r["pi_info"].items()}}

    def actions_np(self, obs):     # cpo_policy.py:875-877
        return self.get_action_outs(obs)["pi"]
