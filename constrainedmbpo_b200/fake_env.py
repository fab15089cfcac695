"""B200-native ``FakeEnv`` with the surface of mbpo/models/fake_env.py:18-365 (forward path).

``step(obs, act)`` is one fused device pass (ensemble forward + post-processing kernels of
libcmbpo_b200.so) instead of ``sess.run`` + ~10 NumPy passes + the 49-pass ``average_dkl``
loop.  NumPy in -> NumPy out with the reference's shapes and dtypes; CUDA tensors in -> CUDA
tensors out (no host round trip), which is what the device-resident rollout uses.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib
from .constructor import construct_model
from .static import resolve as _resolve_static


class FakeEnv:
    supports_rng_offset_dev = True   # step_device takes the device Philox-stream word of ModelRollout

    def __init__(self, true_environment,
                 static_fns, num_networks=7,
                 num_elites=5, hidden_dims=(220, 220, 220),
                 dyn_discount=1,
                 cost_m_discount=1,
                 cares_about_cost=False,
                 session=None,
                 device="cuda:0",
                 seed=None,
                 sample_noise=False,
                 logvar_bound=None,
                 operand_dtype="bf16"):
        # fake_env.py:28-44
        self.domain = true_environment.domain if getattr(true_environment, "domain", None) else ""
        self.env = true_environment
        self.obs_dim = int(np.prod(self.observation_space.shape))
        self.act_dim = int(np.prod(self.action_space.shape))
        self.active_obs_dim = int(self.obs_dim / getattr(self.env, "stacks", 1))
        self._session = session
        self.cares_about_cost = cares_about_cost
        self.rew_dim = 1
        self.num_networks = num_networks
        self.num_elites = num_elites
        self.static_fns = static_fns
        self._static = _resolve_static(static_fns)
        self.static_c = self._static.has_cost_f
        self.prior_dim = 0  # prior_f / post_f exist only for safety-gym (out of scope)
        self.device = torch.device(device)

        input_dim_dyn = self.obs_dim + self.prior_dim + self.act_dim
        input_dim_c = 2 * self.obs_dim + self.act_dim + self.prior_dim
        output_dim_dyn = self.active_obs_dim + self.rew_dim
        self.dyn_loss = 'MSPE'
        # fake_env.py:61-75
        self._model = construct_model(in_dim=input_dim_dyn, out_dim=output_dim_dyn, name='BNN',
                                      loss=self.dyn_loss, hidden_dims=hidden_dims, lr=1e-3,
                                      num_networks=num_networks, num_elites=num_elites,
                                      weighted=dyn_discount < 1, use_scaler_in=True, use_scaler_out=True,
                                      decay=1e-6, max_logvar=.5, min_logvar=-10, session=session,
                                      device=device, seed=seed, operand_dtype=operand_dtype)
        if self.cares_about_cost and not self.static_c:
            # fake_env.py:76-94: separate CostNN ensemble, MSE (non-probabilistic), 64x64
            self.cost_m_loss = 'MSE'
            self._cost_model = construct_model(in_dim=input_dim_c, out_dim=1, loss=self.cost_m_loss,
                                               name='CostNN', hidden_dims=(64, 64), lr=1e-4,
                                               output_activation=None, num_networks=num_networks,
                                               num_elites=num_elites, weighted=cost_m_discount < 1,
                                               use_scaler_in=False, use_scaler_out=False, decay=1e-6,
                                               session=session, device=device,
                                               seed=None if seed is None else seed + 1, operand_dtype=operand_dtype)
        else:
            self._cost_model = None
        # options beyond the live reference (both off by default): Gaussian sampling of the next state in the output
        # kernel (Philox4x32-10, keyed by global row id) and PETS soft-bounding of the log-variance head to
        # (max_logvar, min_logvar) -- commented out upstream (bnn.py:888-889)
        self.sample_noise = bool(sample_noise)
        self.logvar_bound = tuple(float(x) for x in logvar_bound) if logvar_bound is not None else None
        self.dyn_target_var_rm = 1
        self._rng_offset = 0
        self._seed = int(seed) if seed is not None else int(np.random.randint(2 ** 31 - 1))

    @property
    def observation_space(self):
        return self.env.observation_space

    @property
    def action_space(self):
        return self.env.action_space

    @property
    def dyn_target_var(self):  # fake_env.py:270-272
        return self.dyn_target_var_rm

    def random_inds(self, size):  # fake_env.py:277-278
        return self._model.random_inds(batch_size=size)

    # ---- device-level step ------------------------------------------------------------------
    def step_device(self, obs, act, elite_idx=None, perm=None, noise=None, n_rows_dev=None,
                    row_ids=None, want_ep_var=True, out=None, rng_offset_dev=None):
        """CUDA tensors in, dict of CUDA tensors out.  ``elite_idx`` (int32 [B]) injects the
        ``random_inds`` draw; None -> in-kernel Philox over the elite list.  ``perm`` injects the
        ``forward_shuffle`` permutation of the 3-D path; ``perm="device"`` draws it on the GPU (cmbpo_draw_permutation,
        Philox keyed like the elite draws: fresh on every replay of a captured rollout); an int32 CUDA tensor [E] is used
        as is.  ``n_rows_dev``: optional int32 device
        scalar with the live row count (rollout loop; shapes stay at capacity).  ``rng_offset_dev``: optional uint64 /
        int64 device word the kernel adds to the Philox stream id (a captured rollout advances it per replay)."""
        m = self._model
        assert obs.dim() == act.dim()  # fake_env.py:132
        ndim = obs.dim()
        if ndim not in (2, 3):
            raise ValueError("obs must be [B,obs] or [E,B,obs]; 1-D inputs are broken upstream (fake_env.py:204-210)")
        obs = obs.contiguous()
        act = act.contiguous()
        B = obs.shape[-2]
        E, D, OD = m.num_nets, m.out_dim, self.obs_dim
        lead = (E,) if ndim == 3 else ()
        dev = self.device
        o = out if out is not None else {}

        def buf(key, shape, dtype=torch.float32):
            t = o.get(key)
            if t is None or tuple(t.shape) != tuple(shape) or t.dtype != dtype:
                t = torch.empty(shape, dtype=dtype, device=dev)
                o[key] = t
            return t
        next_obs = buf("next_obs", lead + (B, OD))
        rew = buf("rew", lead + (B, 1))
        term = buf("term", lead + (B, 1), torch.uint8)
        cost = buf("cost", lead + (B, 1))
        dkl_path = buf("dkl_path", (B,))
        ep_var_mean = buf("ep_var_mean", (B,))
        ep_var = buf("ep_var", (B, D)) if want_ep_var else None
        scalars = buf("scalars", (4,))
        io = _lib.cmbpo_step_io()
        io.obs_dev, io.act_dev = obs.data_ptr(), act.data_ptr()
        io.n_rows = B
        io.n_rows_dev = n_rows_dev.data_ptr() if n_rows_dev is not None else None
        io.ndim = ndim
        perm_arr = None
        draw_perm = False
        if ndim == 3:
            if isinstance(perm, str):
                if perm != "device":
                    raise ValueError("perm must be an array, a CUDA int32 tensor, None or 'device'")
                if getattr(self, "_perm_dev", None) is None:
                    self._perm_dev = torch.empty(E, dtype=torch.int32, device=dev)
                io.perm_dev, draw_perm = self._perm_dev.data_ptr(), True
            elif torch.is_tensor(perm):
                perm_arr = perm.to(device=dev, dtype=torch.int32).contiguous()
                io.perm_dev = perm_arr.data_ptr()
            else:
                if perm is None:
                    perm = np.random.permutation(E)  # fake_env.py:314
                perm_arr = np.ascontiguousarray(np.asarray(perm, dtype=np.int32))
                io.perm = perm_arr.ctypes.data_as(C.POINTER(C.c_int32))
        keep = [perm_arr]
        if elite_idx is not None:
            elite_idx = elite_idx.to(device=dev, dtype=torch.int32).contiguous()
            io.elite_idx_dev = elite_idx.data_ptr()
        if row_ids is not None:
            io.row_id_dev = row_ids.data_ptr()
        if noise is not None:
            noise = noise.to(device=dev, dtype=torch.float32).contiguous()
            io.noise_dev = noise.data_ptr()
        io.seed, io.rng_offset, io.row_offset = self._seed, self._rng_offset, 0
        self._rng_offset += 1
        if rng_offset_dev is not None:
            io.rng_offset_dev = rng_offset_dev.data_ptr()
        if draw_perm:
            _lib.check(m._lib.cmbpo_draw_permutation(E, io.seed, io.rng_offset, io.rng_offset_dev, 0, io.perm_dev, m._stream()))
        io.sample_mode = 1 if self.sample_noise else 0
        if self.logvar_bound is not None:
            io.bound_logvar, io.max_logvar, io.min_logvar = 1, self.logvar_bound[0], self.logvar_bound[1]
        io.env_id = self._static.env_id
        io.cost_mode = 1 if (self.cares_about_cost and self.static_c) else 0
        io.next_obs_dev, io.rew_dev, io.term_dev = next_obs.data_ptr(), rew.data_ptr(), term.data_ptr()
        io.cost_dev, io.dkl_path_dev = cost.data_ptr(), dkl_path.data_ptr()
        io.ep_var_mean_dev = ep_var_mean.data_ptr()
        io.ep_var_dev = ep_var.data_ptr() if ep_var is not None else None
        io.scalars_dev = scalars.data_ptr()
        ws = m._workspace(B)
        io.workspace_dev, io.workspace_bytes = ws.data_ptr(), ws.numel()
        _lib.check(m._lib.cmbpo_step(m._handle, C.byref(io), m._stream()))
        del keep
        if self.cares_about_cost and not self.static_c:
            # fake_env.py:191-196: CostNN on concat(obs, act, next_obs), ensemble mean
            cmean, _ = self._cost_model.predict_device([obs, act, next_obs], ndim, 1, want_var=False)
            cost.copy_(cmean.reshape(cost.shape))
            scalars[3] = cost.mean()
        return o

    # ---- reference surface ------------------------------------------------------------------
    def step(self, obs, act, deterministic=False, elite_idx=None, perm=None, noise=None):
        """fake_env.py:131-222.  ``deterministic`` is accepted and ignored, as upstream."""
        is_np = not torch.is_tensor(obs)
        if is_np:
            obs = np.asarray(obs)
            act = np.asarray(act)
        assert len(obs.shape) == len(act.shape)
        if len(obs.shape) == 1:
            raise NameError("name 'return_means' is not defined (1-D step is broken upstream, fake_env.py:204-210)")
        to_dev = lambda a: (torch.as_tensor(np.asarray(a, dtype=np.float32)) if is_np else a.to(torch.float32)).to(self.device)
        obs_d, act_d = to_dev(obs), to_dev(act)
        if obs_d.dim() == 2 and elite_idx is None and is_np:
            elite_idx = self._model.random_inds(obs_d.shape[0])  # fake_env.py:168 (global NumPy RNG)
        if elite_idx is not None and not torch.is_tensor(elite_idx):
            elite_idx = torch.as_tensor(np.asarray(elite_idx, dtype=np.int32))
        if noise is not None and not torch.is_tensor(noise):
            noise = torch.as_tensor(np.asarray(noise, dtype=np.float32))
        o = self.step_device(obs_d, act_d, elite_idx=elite_idx, perm=perm, noise=noise)
        term_b = o["term"].bool()
        if is_np:      # one batch of device -> pinned-host copies, one synchronisation (see _lib.to_host)
            sc, next_obs, rewards, terminals, costs, dkl_path, ep_var = _lib.to_host(
                [o["scalars"], o["next_obs"], o["rew"], term_b, o["cost"], o["dkl_path"], o["ep_var"]])
        else:
            sc, next_obs, rewards, terminals, costs = o["scalars"], o["next_obs"], o["rew"], term_b, o["cost"]
            dkl_path, ep_var = o["dkl_path"], o["ep_var"]
        info = {
            'ensemble_dkl_mean': sc[0],
            'ensemble_dkl_path': dkl_path,
            'ensemble_mean_var': sc[1],
            'ensemble_ep_var': ep_var,
            'rew': rewards,
            'rew_mean': sc[2],
            'cost': costs,
            'cost_mean': sc[3],
        }
        return next_obs, rewards, terminals, info

    def train(self, samples, **kwargs):
        return self.train_dyn_model(samples, **kwargs)

    def train_dyn_model(self, samples, **kwargs):
        raise NotImplementedError("model training stays with the reference (fake_env.py:231-247); "
                                  "hand the trained state over with env._model.load_state(...)")

    def train_cost_model(self, samples, **kwargs):
        if self.static_c:
            return {}  # fake_env.py:254-255
        raise NotImplementedError("cost-model training stays with the reference (fake_env.py:249-268)")

    def reset_model(self):
        raise NotImplementedError("re-initialisation belongs to the training side (fake_env.py:280-283)")

    def close(self):
        pass
