"""Device-resident batched model rollout.

What ``CMBPO._train`` drives through ``ModelSampler.reset`` / ``sample`` / ``ModelBuffer``
(mbpo/algorithms/cmbpo.py:335-351; softlearning/samplers/model_sampler.py:220-474;
softlearning/replay_pools/modelbuffer.py:45-173), restricted to what the hot path owns:
``env.step``, the rollout-length cutoff ('schedule' length mask, 'uncertainty' accumulated-KL
limit, ``max_samples`` early termination, terminal mask) and the pool writes.  Critic values,
log-probs and GAE stay with the reference.

Nothing in the loop synchronises with the host: the live row count is a device scalar that the
kernels read; shapes stay at capacity.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib

MODES = {"schedule": 0, "uncertainty": 1, "iv_gae": 2}


class StubPolicy:
    """Synthetic benchmark policy ``a = tanh(obs @ P)``, ``P ~ N(0, 1/obs_dim)``; stands
    in for ``policy.get_action_outs`` (model_sampler.py:275) in throughput runs only."""

    def __init__(self, obs_dim, act_dim, device="cuda:0", seed=4):
        g = torch.Generator().manual_seed(seed)
        self.P = (torch.randn(obs_dim, act_dim, generator=g) / np.sqrt(obs_dim)).to(device).contiguous()
        self.obs_dim, self.act_dim = obs_dim, act_dim
        self._act = None

    def __call__(self, obs, n_rows_dev=None):
        lib = _lib.load()
        B = obs.shape[-2]
        shape = tuple(obs.shape[:-1]) + (self.act_dim,)   # [B, act] or, for 'iv_gae' states, [E, B, act]
        if self._act is None or tuple(self._act.shape) != shape:
            self._act = torch.empty(shape, dtype=torch.float32, device=obs.device)
        stream = C.c_void_p(torch.cuda.current_stream(obs.device).cuda_stream)
        _lib.check(lib.cmbpo_stub_policy(
            C.c_void_p(obs.data_ptr()), C.c_void_p(self.P.data_ptr()), C.c_void_p(self._act.data_ptr()),
            C.c_void_p(n_rows_dev.data_ptr()) if n_rows_dev is not None else None,
            B, self.obs_dim, self.act_dim, obs.shape[0] if obs.dim() == 3 else 1, stream))
        return self._act


class CriticValues:
    """``value_fn`` for ``ModelRollout`` built from the policy's two value ensembles as device models
    (cpo_policy.py:883-945: V and VC are ``construct_model`` ensembles -- 3 nets, 128x128 swish, MSE, i.e. the
    non-probabilistic ``BNN.predict`` branch, bnn.py:718-724 -- so the same forward kernel serves them).
    2-D modes: every path keeps the critic member drawn for it at reset (``policy.random_v_inds`` /
    ``random_vc_inds``, model_sampler.py:248-249), the value of a row is that member's prediction
    (``get_action_outs['v'][v_inds, batch_inds]``, :290).  The per-path gather is an index op on the
    ``[E_v, B]`` predictions; all arithmetic is the ensemble forward."""

    def __init__(self, v_model, vc_model, n_paths, v_inds=None, vc_inds=None, seed=None):
        self.v, self.vc = v_model, vc_model
        dev = v_model.device
        g = torch.Generator().manual_seed(int(seed)) if seed is not None else None
        draw = lambda m: torch.randint(0, m.num_nets, (n_paths,), generator=g)
        as_idx = lambda x, m: (draw(m) if x is None else torch.as_tensor(np.asarray(x))).to(dev, torch.int64)
        self.v_inds, self.vc_inds = as_idx(v_inds, v_model), as_idx(vc_inds, vc_model)

    def __call__(self, obs, n_rows_dev=None, path_ids=None):
        if obs.dim() == 3:
            # 'iv_gae' states [E, B, obs]: get_v feeds them to BNN.predict as 3-D input (cpo_policy.py:927-945), so value
            # member e scores dynamics member e's state -- the value ensembles must have E members (bnn.py:749-754)
            if self.v.num_nets != obs.shape[0] or self.vc.num_nets != obs.shape[0]:
                raise ValueError(f"'iv_gae' states need value ensembles with one member per dynamics member "
                                 f"({obs.shape[0]}), got {self.v.num_nets} / {self.vc.num_nets}")
            out = [m.predict_device([obs], 3, 0, want_var=False)[0][:, :, 0] for m in (self.v, self.vc)]
            return out[0], out[1]
        B = obs.shape[0]
        pid = (path_ids if path_ids is not None else torch.arange(B, device=obs.device)).long().clamp_(0, self.v_inds.shape[0] - 1)
        out = []
        for model, inds in ((self.v, self.v_inds), (self.vc, self.vc_inds)):
            mean, _ = model.predict_device([obs], 2, 0, want_var=False, n_rows_dev=n_rows_dev)          # [E_v, B, 1]
            out.append(mean[:, :, 0].gather(0, inds[pid][None, :])[0])
        return out[0], out[1]


EPS = 1e-8   # model_sampler.py:17


class DklAccumulators:
    """The two sampler accumulators ``compute_dynamics_dkl`` adds to (``_total_dkl`` / ``_total_samples``,
    model_sampler.py:51,61; zeroed by ``ModelSampler.reset``, :251,:260).  They persist across calls upstream, so
    the value returned by a second calibration call is the running mean over BOTH calls (and over whatever
    ``sample`` added in between, :386); pass one object to successive calls to reproduce that."""

    def __init__(self):
        self.reset()

    def reset(self):
        self.total_dkl, self.total_samples = 0.0, 0

    @property
    def dyn_dkl(self):   # model_sampler.py:192-194
        return self.total_dkl / (self.total_samples + EPS)


def compute_dynamics_dkl(env, policy, obs_batch, depth=1, accum=None, elite_fn=None):
    """``ModelSampler.compute_dynamics_dkl`` (model_sampler.py:160-176): roll ``depth`` steps from
    ``obs_batch``, dropping terminated rows, and return (running mean of ``ensemble_dkl_mean``
    weighted by rows) * depth -- the value CMBPO hands to ``set_rollout_dkl`` (cmbpo.py:283-285).
    ``accum``: a ``DklAccumulators`` carried across calls (default: fresh, i.e. right after ``reset``).
    ``elite_fn(step, n_rows) -> int array``: injects the ``BNN.random_inds`` draw of each ``env.step``
    (parity tests); None = the in-kernel Philox draw.  Calibration only (cold path): synchronises every step."""
    dev = env.device
    if not torch.is_tensor(obs_batch):
        obs_batch = torch.as_tensor(np.asarray(obs_batch, dtype=np.float32))
    obs = obs_batch.to(dev, dtype=torch.float32).contiguous()
    acc = accum if accum is not None else DklAccumulators()
    for k in range(depth):
        n_paths = obs.shape[0]
        if n_paths == 0:     # upstream would step an empty batch; its contribution is 0 * 0
            break
        act = policy(obs).clone()
        idx = None
        if elite_fn is not None:
            idx = torch.as_tensor(np.asarray(elite_fn(k, n_paths), dtype=np.int32)).to(dev)
        o = env.step_device(obs, act, elite_idx=idx, want_ep_var=False)
        acc.total_dkl += float(o["scalars"][0].item()) * n_paths
        acc.total_samples += n_paths
        obs = o["next_obs"][~o["term"].reshape(-1).bool()].contiguous()
    return acc.dyn_dkl * depth


class ModelRollout:
    """``reset(start_obs)`` then ``run(policy)``; the pool arrays have the reference's logical
    layout ``[B, max_path_length, dim]`` with row = start-state index (modelbuffer.py:45-114)."""

    def __init__(self, env, max_path_length, rollout_mode="schedule", dkl_lim=None, max_samples=None,
                 row_offset=0, perm_fn=None, value_fn=None, pool_alloc=None, sharded=False):
        if rollout_mode not in MODES:
            raise ValueError("rollout_mode must be 'schedule', 'uncertainty' or 'iv_gae' (cmbpo.py:89)")
        if rollout_mode == "uncertainty" and dkl_lim is None:
            raise ValueError("'uncertainty' mode needs dkl_lim (model_sampler.py:178-179 set_rollout_dkl)")
        self.env = env
        self.T = int(max_path_length)
        self.mode = rollout_mode
        self.dkl_lim = float(dkl_lim) if dkl_lim is not None else 0.0
        self.max_samples = int(max_samples) if max_samples else 0
        self.row_offset = int(row_offset)
        # `sharded`: this rollout holds one row block of a batch that is split over several ranks.  The max_samples cap cuts
        # "the first n still-certain rows" in GLOBAL row order against the GLOBAL sample count (model_sampler.py:327-332);
        # a rank-local cap would admit up to world x the budget and cut different rows than the single-device run, so the
        # combination is refused (the cap needs an exclusive prefix of the per-step certain counts over the ranks).
        if (sharded or self.row_offset) and self.max_samples:
            raise NotImplementedError("max_samples with a row-sharded rollout: the cap is defined on the global row order "
                                      "(model_sampler.py:327-332); run the capped rollout on one device")
        # 'iv_gae' only: perm_fn(step) -> the forward_shuffle permutation FakeEnv.step draws per call
        # (fake_env.py:309-316, global NumPy RNG upstream), injected as a host launch argument (a captured graph then
        # replays the permutations given at capture time).  None: the env draws it on the device, keyed by rng_base, so
        # every replay of a captured rollout shuffles afresh.
        self.perm_fn = perm_fn
        # Optional critics in the loop: value_fn(obs, n_rows_dev, path_ids) -> (v, vc), CUDA float32
        # tensors shaped [capacity] ('iv_gae': [E, capacity]) holding V(obs), VC(obs) of the live rows -- what
        # policy.get_action_outs / get_v / get_vc return after the sampler's v_inds gather (model_sampler.py:275-294,
        # :497-548).  Called once per state: on the start states at reset and on next_obs every step; survivors'
        # values are compacted along with their observations.  With it the rollout fills val / cval pools and the
        # finish records (finish_ptr, last_val, last_cval) that ``finish()`` hands to cmbpo_finish_paths.
        self.value_fn = value_fn
        # pool_alloc(nbytes) -> uint8 CUDA tensor: where the pool arena lives (default: torch.zeros).  The multi-GPU
        # gather passes a symmetric-memory allocator so that peers can pull the pool over NVLink.
        self.pool_alloc = pool_alloc
        self.device = env.device
        self.B = 0
        # Philox stream base of this rollout, on the device: every reset() advances it by T, so a captured graph draws
        # fresh elite members (and Gaussian noise in sample_mode 1) on every replay instead of the counters baked in at
        # capture time (the per-step host counter FakeEnv._rng_offset is a launch argument and stays frozen in a graph).
        # Reproducible runs / shard-vs-single comparisons: ``rng_base.zero_()`` before ``reset``.
        self.rng_base = torch.zeros(1, dtype=torch.int64, device=self.device)
        self._lib = _lib.load()

    def _allocate(self, B, OD):
        dev, T, AD = self.device, self.T, self.env.act_dim
        z = lambda *s, dt=torch.float32: torch.zeros(*s, dtype=dt, device=dev)
        self.B = B
        # 'iv_gae': one trajectory per ensemble member -- leading E on the per-transition arrays
        # (model_sampler.py:225-226, modelbuffer.py:48-51); dyn_error / populated stay [B, T] (:59,:110)
        E = (self.env.num_networks,) if self.mode == "iv_gae" else ()
        # The tensors are the reference's logical [B, T, .] arrays (views); their STORAGE is time-major
        # [T, B, .] so that one sampler step's stores are contiguous (include/cmbpo_b200.h, pool pointers).
        n = len(E)
        spec = [("obs", E + (T, B, OD), torch.float32), ("act", E + (T, B, AD), torch.float32),
                ("next_obs", E + (T, B, OD), torch.float32), ("rew", E + (T, B), torch.float32),
                ("cost", E + (T, B), torch.float32), ("term", E + (T, B), torch.uint8),
                ("dyn_error", (T, B), torch.float32), ("populated", (T, B), torch.uint8)]
        if self.value_fn is not None:
            spec += [("val", E + (T, B), torch.float32), ("cval", E + (T, B), torch.float32)]
        # one arena for all pool arrays (256-byte aligned fields): a single fill clears it, and the multi-GPU gather moves
        # a rank's whole pool with one peer copy (sharding.DensePoolGatherer; ``pool_alloc`` then hands out symmetric memory)
        self.pool_layout, off = {}, 0
        for name, shape, dt in spec:
            nbytes = int(np.prod(shape)) * (4 if dt == torch.float32 else 1)
            self.pool_layout[name] = (off, shape, dt)
            off += (nbytes + 255) // 256 * 256
        if self.pool_alloc is not None:
            self.pool_arena = self.pool_alloc(off)
            self.pool_arena.zero_()
        else:
            self.pool_arena = torch.zeros(off, dtype=torch.uint8, device=dev)
        self.pool = {}
        for name, (o, shape, dt) in self.pool_layout.items():
            nbytes = int(np.prod(shape)) * (4 if dt == torch.float32 else 1)
            st = self.pool_arena[o:o + nbytes].view(dt).view(shape)
            m = len(shape) - 2 - (1 if name in ("obs", "act", "next_obs") else 0)     # index of the T axis
            self.pool[name] = st.transpose(m, m + 1)
        self.cur = [torch.empty(*E, B, OD, dtype=torch.float32, device=dev) for _ in range(2)]
        if self.value_fn is not None:
            self.vcur = [[torch.zeros(*E, B, dtype=torch.float32, device=dev) for _ in range(2)] for _ in range(2)]
            self.finish_ptr = torch.zeros(B, dtype=torch.int32, device=dev)
            self.last_val, self.last_cval = z(*E, B), z(*E, B)
        self.pid = [torch.empty(B, dtype=torch.int32, device=dev) for _ in range(2)]
        self.dkl_accum = z(B, dt=torch.float64)
        self.state = torch.zeros(12, dtype=torch.int32, device=dev)     # cmbpo_rollout_state (48 bytes)
        self._state_init = torch.zeros(12, dtype=torch.int32, device=dev)
        self._state_init[0] = B
        self._arange = torch.arange(B, dtype=torch.int32, device=dev)
        self.alive_hist = torch.zeros(max(T, 1), dtype=torch.int32, device=dev)
        self.scratch = torch.empty(int(self._lib.cmbpo_rollout_scratch_bytes(B)) + 64, dtype=torch.uint8, device=dev)
        self._row_ids = torch.empty(B, dtype=torch.int32, device=dev)
        self._step_out = {}
        self._graph = None

    def reset(self, start_obs):
        """model_sampler.py:220-263 + modelbuffer.py:45-114.  Allocation happens only when the
        number of start states changes; otherwise the buffers are cleared in place (which is what
        makes reset + run capturable in one CUDA graph)."""
        dev = self.device
        if not torch.is_tensor(start_obs):
            start_obs = torch.as_tensor(np.asarray(start_obs, dtype=np.float32))
        B, OD = start_obs.shape
        if B != self.B:
            self._allocate(B, OD)
        else:
            self.pool_arena.zero_()
            self.dkl_accum.zero_()
        self.alive_hist.zero_()
        self.cur[0].copy_(start_obs, non_blocking=True)   # 'iv_gae': broadcast = np.tile(obs[None], (E,1,1))
        self.pid[0].copy_(self._arange)
        self.state.copy_(self._state_init)
        self.rng_base.add_(self.T)
        self.flip = 0
        self.steps_done = 0
        self.gae = None
        if self.value_fn is not None:
            self.finish_ptr.zero_(); self.last_val.zero_(); self.last_cval.zero_()
            v, vc = self.value_fn(self.cur[0], self.state[0:1], self.pid[0])
            self.vcur[0][0].copy_(v); self.vcur[0][1].copy_(vc)

    def step(self, policy):
        """One ``ModelSampler.sample`` (model_sampler.py:265-474) for all alive rows."""
        env = self.env
        cur, pid = self.cur[self.flip], self.pid[self.flip]
        n_dev = self.state[0:1]
        row_ids = pid
        if self.row_offset:
            torch.add(pid, self.row_offset, out=self._row_ids)
            row_ids = self._row_ids
        if getattr(policy, "takes_rollout_context", False):     # device actor: noise keyed by global row id + stream word
            act = policy(cur, n_dev, row_ids=row_ids, rng_offset_dev=self.rng_base)
        else:
            act = policy(cur, n_dev)
        perm = None
        if self.mode == "iv_gae":       # no perm_fn: drawn on the device per step, fresh on every replay of a capture
            perm = self.perm_fn(self.steps_done) if self.perm_fn is not None else \
                ("device" if getattr(env, "supports_rng_offset_dev", False) else None)
        kw = {"rng_offset_dev": self.rng_base} if getattr(env, "supports_rng_offset_dev", False) else {}
        o = env.step_device(cur, act, n_rows_dev=n_dev, row_ids=row_ids, want_ep_var=False, out=self._step_out,
                            perm=perm, **kw)
        io = _lib.cmbpo_rollout_io()
        io.n_paths, io.max_path_length, io.mode = self.B, self.T, MODES[self.mode]
        io.dkl_lim, io.max_samples, io.step_index = self.dkl_lim, self.max_samples, self.steps_done
        io.obs_dim, io.act_dim = env.obs_dim, env.act_dim
        io.cur_obs_dev, io.act_dev, io.next_obs_dev = cur.data_ptr(), act.data_ptr(), o["next_obs"].data_ptr()
        io.rew_dev, io.term_dev, io.cost_dev = o["rew"].data_ptr(), o["term"].data_ptr(), o["cost"].data_ptr()
        io.dkl_path_dev, io.ep_var_mean_dev = o["dkl_path"].data_ptr(), o["ep_var_mean"].data_ptr()
        io.path_id_dev, io.state_dev, io.dkl_accum_dev = pid.data_ptr(), self.state.data_ptr(), self.dkl_accum.data_ptr()
        p = self.pool
        io.pool_obs_dev, io.pool_act_dev, io.pool_next_obs_dev = p["obs"].data_ptr(), p["act"].data_ptr(), p["next_obs"].data_ptr()
        io.pool_rew_dev, io.pool_cost_dev, io.pool_term_dev = p["rew"].data_ptr(), p["cost"].data_ptr(), p["term"].data_ptr()
        io.pool_dyn_error_dev, io.pool_populated_dev = p["dyn_error"].data_ptr(), p["populated"].data_ptr()
        io.new_obs_dev, io.new_path_id_dev = self.cur[self.flip ^ 1].data_ptr(), self.pid[self.flip ^ 1].data_ptr()
        io.scratch_dev, io.scratch_bytes = self.scratch.data_ptr(), self.scratch.numel()
        io.alive_hist_dev = self.alive_hist.data_ptr()
        io.n_members = env.num_networks if self.mode == "iv_gae" else 0
        if self.value_fn is not None:
            v_next, vc_next = self.value_fn(o["next_obs"], n_dev, pid)
            self._v_next = (v_next.contiguous(), vc_next.contiguous())   # keep alive until the launch is enqueued
            vc_, vn_ = self.vcur[self.flip], self.vcur[self.flip ^ 1]
            io.val_cur_dev, io.cval_cur_dev = vc_[0].data_ptr(), vc_[1].data_ptr()
            io.val_next_dev, io.cval_next_dev = self._v_next[0].data_ptr(), self._v_next[1].data_ptr()
            io.new_val_dev, io.new_cval_dev = vn_[0].data_ptr(), vn_[1].data_ptr()
            io.pool_val_dev, io.pool_cval_dev = p["val"].data_ptr(), p["cval"].data_ptr()
            io.finish_ptr_dev = self.finish_ptr.data_ptr()
            io.last_val_dev, io.last_cval_dev = self.last_val.data_ptr(), self.last_cval.data_ptr()
        stream = C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        _lib.check(self._lib.cmbpo_rollout_advance(C.byref(io), stream))
        self.flip ^= 1
        self.steps_done += 1

    def run(self, policy, max_steps=None):
        """Roll until every path has ended.  A path can store at most ``T-1`` transitions
        (model_sampler.py:437, quirk Q8), so ``T-1`` sampler steps always suffice; no host
        synchronisation happens inside the loop."""
        n = self.T - 1 if max_steps is None else min(max_steps, self.T - 1)
        for _ in range(max(n, 1)):
            self.step(policy)
        return self.pool

    def run_until(self, policy, min_alive_ratio=0.1, sample_budget=None, samples_added=0, max_steps=None):
        """The rollout loop as ``CMBPO._train`` runs it (mbpo/algorithms/cmbpo.py:339-348): ``sample`` until the stored
        transitions reach 99 % of ``sample_budget`` (``approx_model_batch``; returns ``keep_rolling = False`` then), or
        ``alive_ratio <= min_alive_ratio``, or nothing is alive.  One host read of the device counters per sampler step
        (the reference reads ``info['alive_ratio']`` every step as well); paths still alive afterwards are bootstrapped by
        ``finish_all_paths`` / ``finish``.  Returns ``(steps_run, keep_rolling)``."""
        n = self.T - 1 if max_steps is None else min(max_steps, self.T - 1)
        keep_rolling = True
        steps = 0
        for _ in range(max(n, 1)):
            self.step(policy)
            steps += 1
            n_alive, _, total, finished = self.counters()
            if sample_budget is not None and total + samples_added >= 0.99 * sample_budget:
                keep_rolling = False
                break
            if finished or n_alive / max(self.B, 1) <= min_alive_ratio:
                break
        return steps, keep_rolling

    @property
    def alive_ratio(self):
        """``info['alive_ratio']`` of the last ``sample`` (model_sampler.py:471) -- synchronises."""
        return self.counters()[0] / max(self.B, 1)

    def finish_all_paths(self):
        """``ModelSampler.finish_all_paths`` (model_sampler.py:550-570): every path still alive is finished at the
        current pool column with V / VC of its current state as bootstrap values.  No host synchronisation; a no-op when
        nothing is alive (which is the case after a full ``run``)."""
        if self.value_fn is None:
            raise RuntimeError("finish_all_paths() needs the critic values: construct ModelRollout with value_fn=...")
        io = _lib.cmbpo_rollout_io()
        io.n_paths, io.max_path_length, io.mode = self.B, self.T, MODES[self.mode]
        io.n_members = self.env.num_networks if self.mode == "iv_gae" else 0
        io.state_dev, io.path_id_dev = self.state.data_ptr(), self.pid[self.flip].data_ptr()
        vc_ = self.vcur[self.flip]
        io.val_cur_dev, io.cval_cur_dev = vc_[0].data_ptr(), vc_[1].data_ptr()
        io.finish_ptr_dev = self.finish_ptr.data_ptr()
        io.last_val_dev, io.last_cval_dev = self.last_val.data_ptr(), self.last_cval.data_ptr()
        stream = C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        _lib.check(self._lib.cmbpo_rollout_finish_all(C.byref(io), stream))

    def capture(self, start_obs, policy):
        """Record ``reset(start_obs)`` + ``run(policy)`` into one CUDA graph.  ``start_obs`` must be a
        CUDA tensor that stays alive (refill it in place before ``replay``); ``policy`` must be a
        device policy without host synchronisation.  One eager rollout runs first so every lazy
        allocation (workspace, output buffers, kernel attributes) happens outside the capture."""
        assert torch.is_tensor(start_obs) and start_obs.is_cuda
        self.reset(start_obs)
        self.run(policy)
        torch.cuda.synchronize(self.device)
        # No garbage collection while the stream is capturing: collecting an unreferenced model (BNN.__del__ ->
        # cmbpo_destroy -> cudaFree) from the cyclic GC in the middle of the capture would invalidate it.
        import gc
        gc.collect()
        was_enabled = gc.isenabled()
        gc.disable()
        try:
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                self.reset(start_obs)
                self.run(policy)
        finally:
            if was_enabled:
                gc.enable()
        self._graph = g
        # the graph holds raw pointers into the start-state tensor and into whatever the policy owns (its action buffer,
        # its networks' workspaces): keep both alive as long as the graph can be replayed
        self._captured_refs = (start_obs, policy)
        return g

    def replay(self):
        self._graph.replay()

    # ---- results ----------------------------------------------------------------------------
    def counters(self):
        """(n_alive, ptr, total_samples, finished) -- synchronises."""
        s = self.state.cpu().numpy()
        total = int(np.frombuffer(s[2:4].tobytes(), dtype=np.int64)[0])
        return int(s[0]), int(s[1]), total, int(s[5])

    def transitions(self, host_out=None):
        """``ModelBuffer.get`` minus GAE (modelbuffer.py:348-405): the populated transitions,
        flattened in [path, t] order.  Returns CUDA tensors, or fills / returns pinned host
        buffers when ``host_out`` is a dict (device->host copy of exactly the populated rows)."""
        mask = self.pool["populated"].bool()
        # 'iv_gae' pools carry a leading member axis: flatten to [n, E, ...]
        per_member = () if self.mode != "iv_gae" else ("obs", "act", "next_obs", "rew", "cost", "term", "val", "cval")
        out = {k: (v.movedim(0, 2)[mask] if k in per_member else v[mask])
               for k, v in self.pool.items() if k != "populated"}
        if host_out is None:
            return out
        n = out["rew"].shape[0]
        for k, v in out.items():
            h = host_out.get(k)
            if h is None or h.shape[0] < n:
                h = torch.empty((max(n, 1),) + tuple(v.shape[1:]), dtype=v.dtype, pin_memory=True)
                host_out[k] = h
            h[:n].copy_(v, non_blocking=True)
        torch.cuda.current_stream(self.device).synchronize()
        host_out["_n"] = n
        return host_out

    def finish(self, gamma=0.99, lam=0.95, cost_gamma=0.99, cost_lam=0.95, model_ind=0, cares_about_cost=True,
               max_uncertainty_r=5.5, max_uncertainty_c=5.5):
        """``ModelBuffer.finish_path_multiple`` for every path of the finished rollout (modelbuffer.py:176-346):
        adds adv / ret / cadv / cret (+ ret_var, cret_var, roll_lengths, croll_lengths and the horizon-trimmed
        ``populated`` in 'iv_gae' mode) to the pool.  Needs ``value_fn``."""
        if self.value_fn is None:
            raise RuntimeError("finish() needs the critic values: construct ModelRollout with value_fn=...")
        from .gae import finish_paths
        self.finish_all_paths()      # paths the rollout loop left alive (early exit): bootstrap them first
        p = self.pool
        res = finish_paths(p["rew"], p["val"], p["cost"], p["cval"], self.finish_ptr, self.last_val, self.last_cval,
                           self.mode, gamma=gamma, lam=lam, cost_gamma=cost_gamma, cost_lam=cost_lam,
                           model_ind=model_ind, cares_about_cost=cares_about_cost,
                           max_uncertainty_r=max_uncertainty_r, max_uncertainty_c=max_uncertainty_c,
                           populated=p["populated"])
        if "populated" in res:                    # pool["populated"] stays what store_multiple populated
            res["populated_trimmed"] = res.pop("populated")
        self.gae = res
        return res

    def get_batch(self, model_ind=0, group=None, **finish_kw):
        """``ModelBuffer.get`` (modelbuffer.py:348-405) for the finished rollout: ``finish()`` (GAE / iv-GAE + horizon
        trim), the advantage normalisation over the populated entries, then the populated-mask selection in
        [path, t] order.  'iv_gae': per-member arrays are taken from member ``model_ind`` (:376-381).  Returns
        ``(batch dict of CUDA tensors, stats)`` with stats = [adv_mean, adv_std, cadv_mean, ret_mean, cret_mean, n]."""
        from .gae import normalize_advantages
        res = self.gae if self.gae is not None else self.finish(model_ind=model_ind, **finish_kw)
        pop = res.get("populated_trimmed", self.pool["populated"])
        stats = normalize_advantages(res["adv"], res["cadv"], pop, ret=res["ret"], cret=res["cret"], group=group)
        mask = pop.bool()
        iv = self.mode == "iv_gae"
        pick = lambda v: (v[model_ind] if iv else v)[mask]
        batch = {k: pick(self.pool[k]) for k in ("obs", "act", "val", "cval", "cost", "rew")}
        batch.update({k: res[k][mask] for k in res if k != "populated_trimmed"})
        return batch, stats

    def get(self, numpy=False):
        out = dict(self.pool)
        out.update(getattr(self, "gae", None) or {})
        if self.value_fn is not None:
            out.update(finish_ptr=self.finish_ptr, last_val=self.last_val, last_cval=self.last_cval)
        out["dkl_accum"] = self.dkl_accum
        if numpy:
            out = {k: v.cpu().numpy() for k, v in out.items()}
            out["term"] = out["term"].astype(bool)
            out["populated"] = out["populated"].astype(bool)
            if "populated_trimmed" in out:
                out["populated_trimmed"] = out["populated_trimmed"].astype(bool)
        return out
