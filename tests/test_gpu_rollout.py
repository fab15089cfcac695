"""Device-resident rollout bookkeeping (model_sampler.py:311-332,:386-474; modelbuffer.py:130-173):
integer / index work, so the bar is BIT-EXACT against the oracle.

Free-running rollouts diverge between bf16 and fp64 arithmetic (errors compound over the horizon and
flip cutoff decisions), so the bookkeeping is checked teacher-forced: the step outputs the CUDA path
produced are recorded per sampler step and replayed through the oracle's sampler loop; every pool
array, counter and accumulator must then match exactly.  The arithmetic of the step itself is covered
by test_gpu_step.py."""
import numpy as np
import pytest

from tests.helpers import build_cuda_env, make_inputs, make_params, orc

pytestmark = pytest.mark.gpu
HIDDEN = (200, 200, 200, 200)


def _recorded_rollout(env, B, T, mode, dkl_lim=None, max_samples=None, cares_about_cost=True, seed=0):
    import torch
    import constrainedmbpo_b200 as cm
    p = make_params(env, HIDDEN, parity_scalers=False, bias=False)  # unfitted scalers: states stay O(1)
    obs, _ = make_inputs(env, B)
    fe = build_cuda_env(p, env, HIDDEN, cares_about_cost=cares_about_cost, seed=seed)
    pol = cm.StubPolicy(fe.obs_dim, fe.act_dim, device="cuda:0")
    ro = cm.ModelRollout(fe, T, rollout_mode=mode, dkl_lim=dkl_lim, max_samples=max_samples)
    ro.reset(obs)
    rec = []
    for t in range(T - 1):
        n = int(ro.state[0].item())
        if n == 0:
            break
        cur = ro.cur[ro.flip][:n].cpu().numpy().copy()
        ro.step(pol)
        o = ro._step_out
        rec.append(dict(cur=cur, act=pol._act[:n].cpu().numpy().copy(),
                        next_obs=o["next_obs"][:n].cpu().numpy().copy(), rew=o["rew"][:n].cpu().numpy().copy(),
                        term=o["term"][:n].cpu().numpy().astype(bool), cost=o["cost"][:n].cpu().numpy().copy(),
                        dkl=o["dkl_path"][:n].cpu().numpy().copy(), epv=o["ep_var_mean"][:n].cpu().numpy().copy()))
    # a few extra steps after everything finished must be no-ops
    for _ in range(2):
        ro.step(pol)
    return p, obs, ro, rec


def _replay_through_oracle(p, env, obs, rec, T, mode, dkl_lim, max_samples):
    k = {"i": 0}

    def policy(cur):
        r = rec[k["i"]]
        assert np.array_equal(cur.astype(np.float32), r["cur"]), "compacted current observations differ"
        return r["act"]

    def step_fn(cur, a, step):
        r = rec[k["i"]]
        k["i"] += 1
        D = cur.shape[1] + 1
        info = {"cost": r["cost"], "ensemble_dkl_path": r["dkl"],
                # the sampler stores mean_d(ep_var) (model_sampler.py:428); replay it as a constant row
                "ensemble_ep_var": np.repeat(r["epv"][:, None], D, axis=1)}
        return r["next_obs"], r["rew"], r["term"], info
    return orc.rollout(p.astype(np.float32), env, obs, policy, T, mode, dkl_lim=dkl_lim, max_samples=max_samples,
                       step_fn=step_fn)


def _assert_pools_equal(ro, ref):
    got = ro.get(numpy=True)
    for k in ("obs", "act", "next_obs", "rew", "cost", "term", "populated"):
        assert np.array_equal(got[k], ref[k]), f"pool[{k}] differs"
    assert np.allclose(got["dyn_error"], ref["dyn_error"], rtol=1e-6, atol=0)  # mean of a repeated row
    n_alive, ptr, total, finished = ro.counters()
    assert total == ref["total_samples"] == int(ref["populated"].sum())
    assert np.array_equal(got["dkl_accum"], ref["dkl_accum"])  # float64 accumulate, like np.zeros upstream
    return total


def test_schedule_mode_bookkeeping_bit_exact_and_q8_length():
    """'schedule' mode: every path ends when its length reaches max_path_length-1 (quirk Q8)."""
    env, B, T = "halfcheetah", 1500, 6
    p, obs, ro, rec = _recorded_rollout(env, B, T, "schedule", cares_about_cost=False)
    ref = _replay_through_oracle(p, env, obs, rec, T, "schedule", None, None)
    total = _assert_pools_equal(ro, ref)
    assert total == B * (T - 1)                         # halfcheetah never terminates
    got = ro.get(numpy=True)
    assert got["populated"][:, : T - 1].all() and not got["populated"][:, T - 1:].any()
    assert ro.counters()[3] == 1


def test_schedule_mode_with_terminals():
    env, B, T = "antsafe", 2000, 8
    p, obs, ro, rec = _recorded_rollout(env, B, T, "schedule")
    ref = _replay_through_oracle(p, env, obs, rec, T, "schedule", None, None)
    _assert_pools_equal(ro, ref)
    got = ro.get(numpy=True)
    # a path is populated on a prefix of columns and stops right after its terminal transition
    pop = got["populated"]
    assert (np.diff(pop.astype(int), axis=1) <= 0).all()
    last = pop.sum(1) - 1
    rows = np.where((last >= 0) & (last < T - 2))[0]
    assert got["term"][rows, last[rows]].all()
    assert len(rec) > 2 and rec[-1]["cur"].shape[0] < B   # some paths did terminate early


def test_uncertainty_mode_bookkeeping_bit_exact():
    """'uncertainty' mode: accumulated ensemble KL vs dkl_lim (model_sampler.py:311-313); the step that
    crosses the limit is NOT stored."""
    import constrainedmbpo_b200 as cm
    env, B, T = "antsafe", 3000, 12
    p0 = make_params(env, HIDDEN, parity_scalers=False, bias=False)
    obs0, _ = make_inputs(env, 1000, seed=9)
    fe0 = build_cuda_env(p0, env, HIDDEN)
    lim = cm.compute_dynamics_dkl(fe0, cm.StubPolicy(29, 8), obs0, depth=3)
    assert lim > 0
    p, obs, ro, rec = _recorded_rollout(env, B, T, "uncertainty", dkl_lim=lim)
    ref = _replay_through_oracle(p, env, obs, rec, T, "uncertainty", lim, None)
    total = _assert_pools_equal(ro, ref)
    got = ro.get(numpy=True)
    assert 0 < total < B * (T - 1)
    assert (got["dkl_accum"] < lim).all()               # accumulators only absorb stored steps
    lens = got["populated"].sum(1)
    assert lens.min() < lens.max()                      # the cutoff is per path


def test_max_samples_early_termination():
    """model_sampler.py:327-332: the first n still-certain rows are cut so total_samples lands on max_samples."""
    env, B, T = "halfcheetah", 1000, 6
    cap = 2300
    p, obs, ro, rec = _recorded_rollout(env, B, T, "schedule", max_samples=cap, cares_about_cost=False)
    ref = _replay_through_oracle(p, env, obs, rec, T, "schedule", None, cap)
    total = _assert_pools_equal(ro, ref)
    assert total == cap
    got = ro.get(numpy=True)
    # step 3 (0-based 2) had room for only 300 rows: the FIRST 700 alive rows were cut
    assert not got["populated"][:700, 2].any() and got["populated"][700:, 2].all()


def test_cuda_graph_replay_matches_eager():
    import torch
    import constrainedmbpo_b200 as cm
    env, B, T = "antsafe", 5000, 10
    p = make_params(env, HIDDEN, parity_scalers=False, bias=False)
    obs, _ = make_inputs(env, B)
    pol = cm.StubPolicy(29, 8)
    res = []
    for graph in (False, True):
        fe = build_cuda_env(p, env, HIDDEN, cares_about_cost=True, seed=3)
        ro = cm.ModelRollout(fe, T, rollout_mode="uncertainty", dkl_lim=0.5)
        start = torch.from_numpy(obs).cuda()
        if graph:
            fe._rng_offset = 0
            ro.capture(start, pol)   # eager warm-up rollout + capture: offsets 0..T-2 twice
            ro.replay(); ro.replay()
        else:
            ro.reset(start); ro.run(pol)
        torch.cuda.synchronize()
        res.append((ro.get(numpy=True), ro.counters()))
    (a, ca), (b, cb) = res
    # Philox stream ids differ between the eager run (offsets 0..) and the captured one (offsets T-1..):
    # compare structure-independent invariants plus exact equality of the draw-independent first step
    assert np.array_equal(a["obs"][:, 0], b["obs"][:, 0]) and np.array_equal(a["act"][:, 0], b["act"][:, 0])
    assert ca[3] == cb[3] == 1
    assert cb[2] == int(b["populated"].sum())


def test_full_size_schedule_rollout_properties():
    """BASELINE configs[2] shape at full size (1M HalfCheetah start states, 'schedule'): size-independent
    properties -- every path stores exactly T-1 transitions, chained (next_obs[t] == obs[t+1]), pool row b
    column 0 holds start state b, rewards finite."""
    import torch
    import constrainedmbpo_b200 as cm
    env, B, T = "halfcheetah", 1_000_000, 5
    p = make_params(env, HIDDEN, parity_scalers=False, bias=False)
    fe = build_cuda_env(p, env, HIDDEN)
    g = torch.Generator(device="cuda").manual_seed(1)
    start = torch.randn(B, 17, device="cuda", generator=g)
    ro = cm.ModelRollout(fe, T, rollout_mode="schedule")
    ro.reset(start)
    ro.run(cm.StubPolicy(17, 6))
    n_alive, ptr, total, finished = ro.counters()
    assert (n_alive, ptr, total, finished) == (0, T - 1, B * (T - 1), 1)
    pool = ro.pool
    assert bool(pool["populated"][:, : T - 1].all()) and not bool(pool["populated"][:, T - 1].any())
    assert torch.equal(pool["obs"][:, 0], start)
    assert torch.equal(pool["next_obs"][:, : T - 2], pool["obs"][:, 1: T - 1])
    assert bool(torch.isfinite(pool["rew"]).all())
    hist = ro.alive_hist.cpu().numpy()
    assert list(hist[: T - 1]) == [B] * (T - 1)


def _recorded_rollout_iv(env, B, T, seed=0, cares_about_cost=True):
    """'iv_gae' mode: [E, n, .] states; records every sampler step's inputs / outputs and permutations."""
    import constrainedmbpo_b200 as cm
    p = make_params(env, HIDDEN, parity_scalers=False, bias=False)
    obs, _ = make_inputs(env, B)
    fe = build_cuda_env(p, env, HIDDEN, cares_about_cost=cares_about_cost, seed=seed)
    pol = cm.StubPolicy(fe.obs_dim, fe.act_dim, device="cuda:0")
    prng = np.random.default_rng(11)
    perms = [prng.permutation(p.num_nets) for _ in range(T + 2)]
    ro = cm.ModelRollout(fe, T, rollout_mode="iv_gae", perm_fn=lambda t: perms[t])
    ro.reset(obs)
    rec = []
    for t in range(T - 1):
        n = int(ro.state[0].item())
        if n == 0:
            break
        cur = ro.cur[ro.flip][:, :n].cpu().numpy().copy()
        ro.step(pol)
        o = ro._step_out
        g = lambda x: x[:, :n].cpu().numpy().copy()
        rec.append(dict(cur=cur, act=g(pol._act), next_obs=g(o["next_obs"]), rew=g(o["rew"]),
                        term=g(o["term"]).astype(bool), cost=g(o["cost"]),
                        dkl=o["dkl_path"][:n].cpu().numpy().copy(), epv=o["ep_var_mean"][:n].cpu().numpy().copy()))
    for _ in range(2):
        ro.step(pol)
    return p, obs, ro, rec, perms


def test_iv_gae_mode_bookkeeping_bit_exact():
    """'iv_gae' (model_sampler.py:225-226,:444-445; modelbuffer.py:48-51,:134-150): every member rolls its own
    state; pools carry a leading E; a path ends when ANY member terminates.  Teacher-forced, bit-exact."""
    env, B, T = "antsafe", 1200, 7
    p, obs, ro, rec, perms = _recorded_rollout_iv(env, B, T)
    k = {"i": 0}

    def policy(cur):
        r = rec[k["i"]]
        assert np.array_equal(cur.astype(np.float32), r["cur"]), "compacted member states differ"
        return r["act"]

    def step_fn(cur, a, step):
        r = rec[k["i"]]
        k["i"] += 1
        D = cur.shape[-1] + 1
        info = {"cost": r["cost"], "ensemble_dkl_path": r["dkl"],
                "ensemble_ep_var": np.repeat(r["epv"][:, None], D, axis=1)}
        return r["next_obs"], r["rew"], r["term"], info
    ref = orc.rollout_iv_gae(p.astype(np.float32), env, obs, policy, T, step_fn=step_fn)
    got = ro.get(numpy=True)
    for key in ("obs", "act", "next_obs", "rew", "cost", "term", "populated"):
        assert got[key].shape == ref[key].shape, key
        assert np.array_equal(got[key], ref[key]), f"pool[{key}] differs"
    assert np.allclose(got["dyn_error"], ref["dyn_error"], rtol=1e-6, atol=0)
    assert np.array_equal(got["dkl_accum"], ref["dkl_accum"])
    n_alive, ptr, total, finished = ro.counters()
    assert total == ref["total_samples"] == int(ref["populated"].sum()) and finished == 1
    assert len(rec) > 2 and rec[-1]["cur"].shape[1] < B            # some paths ended through a member's terminal
    assert np.array_equal(got["obs"][:, :, 0], np.tile(obs[None], (7, 1, 1)))  # np.tile start states (:226)
    # members diverge after the first step: trajectories are per member
    assert not np.array_equal(got["next_obs"][0, :, 0], got["next_obs"][1, :, 0])
    tr = ro.transitions()
    assert tr["obs"].shape == (total, 7, 29) and tr["dyn_error"].shape == (total,)


def test_iv_gae_step_matches_oracle_on_compacted_states():
    """Arithmetic of the 3-D step inside the rollout (live count < capacity: member stride stays at
    capacity): step outputs of sampler step 2 against the fp64 oracle on the same member states."""
    from tests.helpers import assert_means_close, out_scale
    env, B, T = "antsafe", 900, 5
    p, obs, ro, rec, perms = _recorded_rollout_iv(env, B, T, cares_about_cost=True)
    t = 2
    r = rec[t]
    assert r["cur"].shape[1] < B
    nobs, rew, term, info = orc.fake_env_step(p, env, r["cur"], r["act"], perm=perms[t], cares_about_cost=True)
    s = out_scale(p)
    assert_means_close(r["next_obs"], nobs, s[:-1], what="iv next_obs")
    assert_means_close(r["rew"], rew, s[-1:], what="iv reward")
    assert np.allclose(r["dkl"], info["ensemble_dkl_path"], rtol=0.1, atol=1e-7)


@pytest.mark.parametrize("mode", ["uncertainty", "iv_gae"])
def test_pack_records_kernel_matches_mask_selection(mode):
    """cmbpo_pack_records == populated-mask selection in [path, t] order (modelbuffer.py:348-405), bit-exact,
    for a ragged 2-D pool and an 'iv_gae' pool; coordinates carry the global path id."""
    import torch
    import constrainedmbpo_b200 as cm
    from constrainedmbpo_b200 import sharding, _lib
    env, B, T = "antsafe", 1777, 9
    p = make_params(env, HIDDEN, parity_scalers=False, bias=False)
    obs, _ = make_inputs(env, B)
    fe = build_cuda_env(p, env, HIDDEN, cares_about_cost=True, seed=1)
    ro = cm.ModelRollout(fe, T, rollout_mode=mode, dkl_lim=2e-5 if mode == "uncertainty" else None, row_offset=5000)
    ro.reset(obs)
    ro.run(cm.StubPolicy(29, 8))
    E = 7 if mode == "iv_gae" else 1
    W = sharding.record_width(29, 8, E)
    assert W == _lib.load().cmbpo_record_width(29, 8, E)
    n_ref = int(ro.pool["populated"].sum())
    assert 0 < n_ref < B * (T - 1)
    out = torch.full((n_ref + 3, W), -7.0, device="cuda")
    count = torch.zeros(1, dtype=torch.int64, device="cuda")
    scratch = torch.empty(int(_lib.load().cmbpo_pack_scratch_bytes(B)) + 64, dtype=torch.uint8, device="cuda")
    sharding.pack_records_cuda(ro.pool, ro.row_offset, 29, 8, E, out, count, scratch)
    torch.cuda.synchronize()
    assert int(count) == n_ref
    rec, coords = sharding.pack_records(ro.pool, ro.row_offset)
    assert rec.shape == (n_ref, W - 2)
    assert torch.equal(out[:n_ref, : W - 2], rec)
    assert torch.equal(out[:n_ref, W - 2:].contiguous().view(torch.int32).long(), coords)
    assert bool((out[n_ref:] == -7.0).all())                      # nothing written past the count
    un = sharding.unpack_gathered(out[:n_ref], 29, 8, E)
    assert torch.equal(un["coords"], coords) and un["term"].dtype == torch.bool
    if E > 1:
        assert un["obs"].shape == (n_ref, 7, 29) and un["rew"].shape == (n_ref, 7)
        assert torch.equal(un["obs"], ro.transitions()["obs"])
    # too-small capacity: count still reports the need, the buffer is not overrun
    small = torch.full((10, W), -7.0, device="cuda")
    sharding.pack_records_cuda(ro.pool, ro.row_offset, 29, 8, E, small, count, scratch)
    torch.cuda.synchronize()
    assert int(count) == n_ref


class _ReplayEnv:
    """Stands in for FakeEnv inside ModelRollout: step_device hands back the step outputs the REFERENCE's
    sampler received (tests/golden/sampler_*.npz), so cmbpo_rollout_advance is checked against the pools the
    reference's real ModelSampler + ModelBuffer produced from exactly those arrays."""
    def __init__(self, steps, B, E, iv):
        import torch
        self.steps, self.B, self.iv = steps, B, iv
        self.device = torch.device("cuda:0")
        self.obs_dim, self.act_dim, self.num_networks = 29, 8, E
        self.i = 0
        self.seen_cur = []

    def _cap(self, a, dtype, members=None):
        """Recorded [n, ...] (or [E, n, ...]) array -> capacity-shaped CUDA tensor, live rows first."""
        import torch
        members = self.iv if members is None else members
        a = np.ascontiguousarray(a)
        lead = (self.num_networks,) if members else ()
        n = a.shape[len(lead)]
        out = torch.zeros(lead + (self.B,) + tuple(a.shape[len(lead) + 1:]), dtype=dtype, device=self.device)
        src = torch.from_numpy(a).to(self.device, dtype)
        if members:
            out[:, :n] = src
        else:
            out[:n] = src
        return out

    def _record(self):
        if self.i < len(self.steps):
            return self.steps[self.i]
        lead = (self.num_networks,) if self.iv else ()   # every path has ended: zero live rows
        z = lambda *tail, dt=np.float32: np.zeros(lead + (0,) + tail, dt)
        return dict(cur=z(29), act=z(8), next_obs=z(29), rew=z(1), term=z(1, dt=bool), cost=z(1),
                    dkl=np.zeros(0, np.float32), ep_var=np.zeros((0, 30), np.float32))

    def policy(self, cur, n_dev):
        import torch
        n = int(n_dev.item())
        r = self._record()
        got = (cur[:, :n] if self.iv else cur[:n]).cpu().numpy()
        assert got.shape == r["cur"].shape and np.array_equal(got, r["cur"]), \
            f"compacted observations differ from the reference's at step {self.i}"
        return self._cap(r["act"], torch.float32)

    def step_device(self, cur, act, n_rows_dev=None, row_ids=None, want_ep_var=False, out=None, perm=None):
        import torch
        r = self._record()
        self.i += 1
        f32 = torch.float32
        out.update({
            "next_obs": self._cap(r["next_obs"], f32), "rew": self._cap(r["rew"], f32),
            "term": self._cap(r["term"].astype(np.uint8), torch.uint8), "cost": self._cap(r["cost"], f32),
            "dkl_path": self._cap(r["dkl"].astype(np.float32), f32, members=False),
            # the sampler stores np.mean(dyn_ep_var, axis=-1) (model_sampler.py:428)
            "ep_var_mean": self._cap(np.mean(r["ep_var"], axis=-1).astype(np.float32), f32, members=False),
        })
        return out


@pytest.mark.parametrize("name", ["schedule", "uncertainty", "iv_gae"])
def test_rollout_bookkeeping_matches_reference_modelsampler_golden(name):
    """cmbpo_rollout_advance (through ModelRollout) teacher-forced with the step outputs the reference's REAL
    ModelSampler received, against the pools its REAL ModelBuffer ended with: bit-exact (integer / index / copy
    work), all three rollout modes, max_samples early termination included."""
    import os
    import constrainedmbpo_b200 as cm
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", f"sampler_{name}.npz"), allow_pickle=False)
    steps = [{k: g[f"step{i}_{k}"] for k in ("cur", "act", "next_obs", "rew", "term", "cost", "dkl", "ep_var")}
             for i in range(int(g["n_steps"]))]
    B, T, E = g["start_obs"].shape[0], int(g["T"]), int(g["E"])
    env = _ReplayEnv(steps, B, E, name == "iv_gae")
    ro = cm.ModelRollout(env, T, rollout_mode=name, dkl_lim=float(g["dkl_lim"]) or None,
                         max_samples=int(g["max_samples"]) or None)
    ro.reset(g["start_obs"])
    for _ in range(T - 1):
        ro.step(lambda cur, n_dev: env.policy(cur, n_dev))
    assert env.i >= len(steps)
    got = ro.get(numpy=True)
    for ours, ref in (("obs", "obs_buf"), ("act", "act_buf"), ("next_obs", "nextobs_buf"), ("rew", "rew_buf"),
                      ("cost", "cost_buf"), ("term", "term_buf"), ("dyn_error", "dyn_error_buf")):
        assert got[ours].shape == g[ref].shape and np.array_equal(got[ours], g[ref]), ours
    if name == "iv_gae":   # the reference trims populated_mask by the iv horizon afterwards
        assert np.array_equal(got["populated"], np.arange(T)[None, :] < g["path_length"][:, None])
    else:
        assert np.array_equal(got["populated"], g["populated_mask"])
    n_alive, ptr, total, finished = ro.counters()
    assert (n_alive, total, finished) == (0, int(g["total_samples"]), 1)
    assert np.array_equal(got["dkl_accum"], g["dyn_dkl_path"])
    hist = ro.alive_hist.cpu().numpy()
    assert list(hist[: len(g["alive_hist"])]) == list(g["alive_hist"])


@pytest.mark.parametrize("B,T", [(1, 2), (1, 5), (257, 3), (256, 2), (511, 4)])
def test_rollout_edge_sizes_bit_exact(B, T):
    """Ragged / minimal sizes: one path, one stored column (T=2 -> a path stores exactly T-1 = 1 transition, Q8),
    row counts around the 256-row block boundary."""
    env = "antsafe"
    p, obs, ro, rec = _recorded_rollout(env, B, T, "schedule")
    ref = _replay_through_oracle(p, env, obs, rec, T, "schedule", None, None)
    total = _assert_pools_equal(ro, ref)
    assert 0 < total <= B * (T - 1)
    from constrainedmbpo_b200 import sharding
    rec_t, coords = sharding.pack_records(ro.pool, 0)
    assert rec_t.shape[0] == total


def test_pack_records_empty_pool():
    """A pool nothing was stored in (every path cut at the first step by a zero KL limit) packs to zero records."""
    import torch
    import constrainedmbpo_b200 as cm
    from constrainedmbpo_b200 import sharding, _lib
    env, B, T = "antsafe", 300, 5
    p = make_params(env, HIDDEN, parity_scalers=False, bias=False)
    obs, _ = make_inputs(env, B)
    fe = build_cuda_env(p, env, HIDDEN, cares_about_cost=True)
    ro = cm.ModelRollout(fe, T, rollout_mode="uncertainty", dkl_lim=0.0)
    ro.reset(obs)
    ro.run(cm.StubPolicy(29, 8))
    assert ro.counters() == (0, 0, 0, 1) and int(ro.pool["populated"].sum()) == 0
    W = sharding.record_width(29, 8, 1)
    out = torch.full((4, W), -7.0, device="cuda")
    count = torch.full((1,), 99, dtype=torch.int64, device="cuda")
    scratch = torch.empty(int(_lib.load().cmbpo_pack_scratch_bytes(B)) + 64, dtype=torch.uint8, device="cuda")
    sharding.pack_records_cuda(ro.pool, 0, 29, 8, 1, out, count, scratch)
    torch.cuda.synchronize()
    assert int(count) == 0 and bool((out == -7.0).all())


def _golden_policy_values(seed_cfg):
    """The FakePolicy of tests/golden/make_golden_sampler.py (deterministic functions of the observation) and the
    per-path v_inds / vc_inds the reference sampler drew at reset (same seeded generator, same call order)."""
    r = np.random.default_rng(seed_cfg + 10)
    r.standard_normal((29, 8))                          # P (actions come from the recording)
    wv = (r.standard_normal((7, 29)) / np.sqrt(29)).astype(np.float32)
    wc = (r.standard_normal((7, 29)) / np.sqrt(29)).astype(np.float32)
    return wv, wc, np.random.default_rng(seed_cfg + 11)


@pytest.mark.parametrize("name,seed", [("schedule", 21), ("uncertainty", 22), ("iv_gae", 23)])
def test_rollout_values_and_finish_records_match_reference_golden(name, seed):
    """Critic values in the loop (SURVEY §8f N1 plumbing) + finish_path_multiple (N2), end to end against the
    reference's real sampler / pool: val_buf / cval_buf, the column and bootstrap values of every finished path
    (model_sampler.py:497-548), then adv / ret / cadv / cret (and the iv outputs + horizon trim) from finish()."""
    import os
    import torch
    import constrainedmbpo_b200 as cm
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", f"sampler_{name}.npz"), allow_pickle=False)
    steps = [{k: g[f"step{i}_{k}"] for k in ("cur", "act", "next_obs", "rew", "term", "cost", "dkl", "ep_var")}
             for i in range(int(g["n_steps"]))]
    B, T, E = g["start_obs"].shape[0], int(g["T"]), int(g["E"])
    iv = name == "iv_gae"
    wv, wc, prng = _golden_policy_values(seed)
    v_inds, vc_inds = prng.integers(0, 3, B), prng.integers(0, 3, B)       # policy.random_v_inds / random_vc_inds at reset
    env = _ReplayEnv(steps, B, E, iv)

    def value_fn(obs, n_dev, pid):
        n = int(n_dev.item())
        o = (obs[:, :n] if iv else obs[:n]).cpu().numpy()
        ids = pid[:n].cpu().numpy()
        lead = (E,) if iv else ()
        v, vc = np.zeros(lead + (B,), np.float32), np.zeros(lead + (B,), np.float32)
        if n:
            if iv:   # get_v on [E, n, obs] -> [E, n]
                v[:, :n] = np.sin(np.einsum("ebd,ed->eb", o, wv)).astype(np.float32)
                vc[:, :n] = (0.5 * np.sin(np.einsum("ebd,ed->eb", o, wc)).astype(np.float32) + 0.5)
            else:    # get_v(...)[v_inds, arange]: 3 critics, one drawn per path at reset
                allv = np.sin(o @ wv[:3].T).T.astype(np.float32)
                allc = (0.5 * np.sin(o @ wc[:3].T).T.astype(np.float32) + 0.5)
                v[:n], vc[:n] = allv[v_inds[ids], np.arange(n)], allc[vc_inds[ids], np.arange(n)]
        return torch.from_numpy(v).cuda(), torch.from_numpy(vc).cuda()

    ro = cm.ModelRollout(env, T, rollout_mode=name, dkl_lim=float(g["dkl_lim"]) or None,
                         max_samples=int(g["max_samples"]) or None, value_fn=value_fn)
    ro.reset(g["start_obs"])
    for _ in range(T - 1):
        ro.step(lambda cur, n_dev: env.policy(cur, n_dev))
    got = ro.get(numpy=True)
    # the fixture's policy evaluated the bootstrap values on row SUBSETS (get_v(obs[term_mask]), model_sampler.py:521),
    # this value_fn on all live rows: a float32 BLAS product may round differently for different batch shapes, so
    # the value arrays are compared to 1 ulp-ish; WHICH value lands where (and the finish column) is exact
    close = lambda a, b: a.shape == b.shape and np.allclose(a, b, rtol=1e-6, atol=1e-7)
    assert close(got["val"], g["val_buf"]) and close(got["cval"], g["cval_buf"])
    assert np.array_equal(got["val"] != 0, g["val_buf"] != 0)
    assert np.array_equal(got["finish_ptr"], g["finish_ptr"])
    assert close(got["last_val"], g["last_val"]) and close(got["last_cval"], g["last_cval"])
    assert np.array_equal(got["last_val"] == 0, g["last_val"] == 0)      # env terminals bootstrap with exactly 0
    res = ro.finish(gamma=float(g["gamma"]), lam=float(g["lam"]), cost_gamma=float(g["cost_gamma"]), cost_lam=float(g["cost_lam"]),
                    model_ind=int(g["model_ind"]), cares_about_cost=True,
                    max_uncertainty_r=float(g["max_uncertainty"][0]), max_uncertainty_c=float(g["max_uncertainty"][1]))
    keys = [("adv", "adv_buf"), ("ret", "ret_buf"), ("cadv", "cadv_buf"), ("cret", "cret_buf")]
    if iv:
        keys += [("ret_var", "ret_var_buf"), ("cret_var", "cret_var_buf"), ("roll_lengths", "roll_lengths_buf")]
    for ours, ref in keys:
        assert np.allclose(res[ours].cpu().numpy(), g[ref], rtol=2e-5, atol=2e-6), ours
    if iv:
        assert np.array_equal(res["populated_trimmed"].cpu().numpy().astype(bool), g["populated_mask"])
    # ModelBuffer.get: normalised advantages + populated-mask selection, against what the reference's get() returned
    batch, stats = ro.get_batch(model_ind=int(g["model_ind"]))
    b = {k: v.cpu().numpy() for k, v in batch.items()}
    assert b["obs"].shape == g["get_obs"].shape and np.array_equal(b["obs"], g["get_obs"]) and np.array_equal(b["act"], g["get_act"])
    for ours, ref in (("adv", "get_adv"), ("cadv", "get_cadv"), ("ret", "get_ret"), ("cret", "get_cret"), ("val", "get_val"),
                      ("cost", "get_cost")):
        assert np.allclose(b[ours], g[ref], rtol=5e-5, atol=5e-6), ours
    assert int(stats[5].item()) == int(g["diag_poolm_batch_size"])


def test_rollout_with_device_critics_end_to_end():
    """N1 wiring: V / VC as device ensembles (3 nets, 128x128 swish, MSE -> non-probabilistic predict) feeding the
    rollout through CriticValues; stored values equal BNN.predict of the stored observations with the per-path
    member, bootstrap rules hold, and finish() equals the oracle's GAE on the same pools."""
    import torch
    import constrainedmbpo_b200 as cm
    env, B, T = "antsafe", 700, 7
    p = make_params(env, HIDDEN, parity_scalers=False, bias=False)
    obs, _ = make_inputs(env, B)
    fe = build_cuda_env(p, env, HIDDEN, cares_about_cost=True, seed=2)
    critics = []
    for k, name in enumerate(("V", "VC")):
        pc = orc.init_ensemble(np.random.default_rng(40 + k), 29, 1, (128, 128), 3, probabilistic=False,
                               use_scaler_in=False, use_scaler_out=False, elites=[0, 1], dtype=np.float32)
        m = cm.construct_model(in_dim=29, out_dim=1, name=name, hidden_dims=(128, 128), num_networks=3, num_elites=2,
                               loss="MSE", use_scaler_in=False, use_scaler_out=False, device="cuda:0")
        m.load_state(pc.weights, pc.biases, elites=pc.elites)
        critics.append(m)
    vf = cm.CriticValues(critics[0], critics[1], B, seed=3)
    ro = cm.ModelRollout(fe, T, rollout_mode="schedule", value_fn=vf)
    ro.reset(obs)
    ro.run(cm.StubPolicy(29, 8))
    got = ro.get(numpy=True)
    pop = got["populated"]
    assert pop.sum() > B and got["finish_ptr"].min() >= 1 and (got["finish_ptr"] == pop.sum(1)).all()
    # stored values = the drawn member's prediction for the stored observation
    for key, m, inds in (("val", critics[0], vf.v_inds), ("cval", critics[1], vf.vc_inds)):
        pred = m.predict(got["obs"][pop], factored=True, inc_var=False)[..., 0]          # [3, n]
        member = inds.cpu().numpy()[np.where(pop)[0]]
        assert np.allclose(got[key][pop], pred[member, np.arange(pred.shape[1])], rtol=1e-5, atol=1e-6)
    # environment terminals bootstrap the value with 0, everything else with V(next state)
    last = pop.sum(1) - 1
    ended_by_term = got["term"][np.arange(B), last] & (last < T - 2)
    assert ended_by_term.any() and (got["last_val"][ended_by_term] == 0).all()
    assert (got["last_val"][~ended_by_term] != 0).all() and (got["last_cval"] != 0).all()
    res = ro.finish(gamma=0.99, lam=0.95, cost_gamma=0.97, cost_lam=0.9)
    ref = orc.finish_paths(got["rew"], got["val"], got["cost"], got["cval"], got["finish_ptr"], got["last_val"],
                           got["last_cval"], "schedule", 0.99, 0.95, 0.97, 0.9)
    for k in ("adv", "ret", "cadv", "cret"):
        assert np.allclose(res[k].cpu().numpy(), ref[k], rtol=2e-5, atol=2e-6), k


def test_host_drain_pipeline_delivers_every_rollout():
    """HostDrain (pack kernel + one pinned D2H per rollout on a side stream, one rollout behind): each rollout's records
    arrive on the host complete and in [path, t] order, also when consecutive rollouts differ in size."""
    import torch
    import constrainedmbpo_b200 as cm
    from constrainedmbpo_b200 import sharding
    env, T = "antsafe", 8
    p = make_params(env, HIDDEN, parity_scalers=False, bias=False)
    fe = build_cuda_env(p, env, HIDDEN, cares_about_cost=True, seed=4)
    pol = cm.StubPolicy(29, 8)
    B = 3000
    ro = cm.ModelRollout(fe, T, rollout_mode="uncertainty", dkl_lim=3e-5)
    drain = sharding.HostDrain(ro, compact=False)
    want = []
    got = []
    for k in range(4):
        obs, _ = make_inputs(env, B, seed=20 + k)
        ro.reset(obs)
        ro.run(pol)
        rec, coords = sharding.pack_records(ro.pool, 0)
        want.append(torch.cat([rec, coords.to(torch.int32).view(torch.float32)], dim=1).cpu())
        drain.launch()
        if k > 0:
            h, cnt = drain.last                      # rollout k-1: its copy was just enqueued on the side stream
            drain._side.synchronize()
            got.append(h.clone())
    h, cnt = drain.wait()
    got.append(h.clone())
    assert len({w.shape[0] for w in want}) > 1       # the rollouts really differ in size
    for w, g_ in zip(want, got):
        assert w.shape == g_.shape and torch.equal(w.view(torch.int32), g_.view(torch.int32))


def test_compute_dynamics_dkl_matches_cpu_port():
    """A13, ``ModelSampler.compute_dynamics_dkl`` (model_sampler.py:160-176): same 5 000 calibration states, same
    stub policy matrix, the ``random_inds`` draws of every ``env.step`` injected into both sides.  The value sets
    ``dkl_lim`` -- every 'uncertainty' cutoff downstream -- so it is compared with the CPU port of the reference
    (fp32 GEMMs + the NumPy post-processing) and with the fp64 oracle loop; the accumulators persist across calls
    like the sampler's ``_total_dkl`` / ``_total_samples``."""
    import torch
    import constrainedmbpo_b200 as cm
    from oracle import cpu_baseline as cb
    env, B, depth = "antsafe", 5000, 3
    p = make_params(env, HIDDEN)
    obs, _ = make_inputs(env, B, seed=21)
    obs[:, 0] = np.random.default_rng(4).uniform(0.3, 0.9, B)     # most rows survive the first step
    fe = build_cuda_env(p, env, HIDDEN)
    pol = cm.StubPolicy(29, 8, seed=4)
    P = pol.P.cpu().numpy()
    policy_np = lambda o: np.tanh(np.asarray(o, np.float32) @ P)
    draws = [np.random.default_rng(100 + k).choice(p.elites, size=B) for k in range(2 * depth)]
    acc = cm.DklAccumulators()
    got = cm.compute_dynamics_dkl(fe, pol, obs, depth=depth, accum=acc, elite_fn=lambda k, n: draws[k][:n])
    ens = cb.TorchEnsemble(p)
    totals = [0.0, 0]
    ref = cb.compute_dynamics_dkl(ens, env, obs, policy_np, depth, elite_fn=lambda k, n: draws[k][:n], totals=totals)
    # rows near a termination threshold may drop out on one side only: the row counts must agree to within those
    assert abs(acc.total_samples - totals[1]) <= 0.01 * totals[1], (acc.total_samples, totals[1])
    assert acc.total_samples > 2 * B            # the calibration really rolled `depth` steps
    assert abs(got - ref) <= 2e-2 * ref, (got, ref)
    # a second call keeps accumulating (model_sampler.py:168-169): running mean over both calls
    obs2, _ = make_inputs(env, 1000, seed=22)
    got2 = cm.compute_dynamics_dkl(fe, pol, obs2, depth=1, accum=acc, elite_fn=lambda k, n: draws[depth + k][:n])
    ref2 = cb.compute_dynamics_dkl(ens, env, obs2, policy_np, 1, elite_fn=lambda k, n: draws[depth + k][:n], totals=totals)
    assert acc.total_samples == pytest.approx(totals[1], rel=0.01)
    assert abs(got2 - ref2) <= 2e-2 * ref2
    assert got2 == pytest.approx(acc.total_dkl / (acc.total_samples + 1e-8))     # depth 1: the running mean itself
    # a fresh accumulator (= right after ModelSampler.reset) gives the per-call mean
    fresh = cm.compute_dynamics_dkl(fe, pol, obs2, depth=1, elite_fn=lambda k, n: draws[depth + k][:n])
    assert fresh != pytest.approx(got2, rel=1e-6)


@pytest.mark.parametrize("mode", ["uncertainty", "schedule", "iv_gae"])
def test_compact_records_rebuild_the_full_records(mode):
    """Compact records (no next_obs block; next_obs[t] = obs[t+1] inside a path, model_sampler.py:463-467, plus the
    next_obs of every path's last transition): rebuilt on the host they equal the full records bit for bit, through
    the pack kernel directly and through the HostDrain pipeline (what bench.py's e2e leg times)."""
    import torch
    import constrainedmbpo_b200 as cm
    from constrainedmbpo_b200 import sharding, _lib
    env, B, T = "antsafe", 2100, 7
    p = make_params(env, HIDDEN, parity_scalers=False, bias=False)
    obs, _ = make_inputs(env, B)
    fe = build_cuda_env(p, env, HIDDEN, cares_about_cost=True, seed=2)
    off = 7000
    ro = cm.ModelRollout(fe, T, rollout_mode=mode, dkl_lim=2e-5 if mode == "uncertainty" else None, row_offset=off)
    ro.reset(obs)
    ro.run(cm.StubPolicy(29, 8))
    E = 7 if mode == "iv_gae" else 1
    W, Wc = sharding.record_width(29, 8, E), sharding.record_width(29, 8, E, compact=True)
    assert Wc == _lib.load().cmbpo_record_width_compact(29, 8, E) == W - E * 29
    n = int(ro.pool["populated"].sum())
    assert 0 < n
    count = torch.zeros(1, dtype=torch.int64, device="cuda")
    scratch = torch.empty(int(_lib.load().cmbpo_pack_scratch_bytes(B)) + 64, dtype=torch.uint8, device="cuda")
    full = torch.empty(n, W, device="cuda")
    sharding.pack_records_cuda(ro.pool, off, 29, 8, E, full, count, scratch)
    comp = torch.full((n + 2, Wc), -7.0, device="cuda")
    last = torch.full((B, E * 29), -3.0, device="cuda")
    sharding.pack_records_cuda(ro.pool, off, 29, 8, E, comp, count, scratch, last_next_obs=last)
    torch.cuda.synchronize()
    assert int(count) == n and bool((comp[n:] == -7.0).all())
    a = sharding.unpack_gathered(full, 29, 8, E)
    b = sharding.unpack_gathered(comp[:n], 29, 8, E, last_next_obs=last, row_offset=off)
    for k in a:
        assert torch.equal(a[k], b[k]), k
    lens = ro.pool["populated"].sum(1)
    assert bool((last[lens == 0] == 0).all())        # paths without a stored transition
    # NumPy twin of the reconstruction (host consumers)
    nx = sharding.reconstruct_next_obs(b["obs"].cpu().numpy(), b["coords"][:, 0].cpu().numpy(), last.cpu().numpy(), off)
    assert np.array_equal(nx, a["next_obs"].cpu().numpy())
    # through the drain pipeline
    drain = sharding.HostDrain(ro)
    drain.launch()
    h, cnt = drain.wait()
    assert cnt == [n] and h.shape == (n, Wc) and torch.equal(h, comp[:n].cpu())
    assert torch.equal(drain.last_next, last.cpu())
    assert drain.bytes_per_rollout(n) == n * Wc * 4 + B * E * 29 * 4


def test_early_exit_then_finish_all_paths_matches_reference_golden():
    """The rollout loop left after 3 of 9 sampler steps with 48 of 56 paths alive (CMBPO._train breaks on alive_ratio or
    the sample budget, cmbpo.py:339-348), then ``finish_all_paths`` (model_sampler.py:550-570): finish columns, bootstrap
    values V / VC(current state), GAE outputs and ``get()`` against the reference's REAL sampler + pool
    (tests/golden/sampler_earlybreak.npz)."""
    import os
    import torch
    import constrainedmbpo_b200 as cm
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "sampler_earlybreak.npz"), allow_pickle=False)
    n_steps = int(g["n_steps"])
    steps = [{k: g[f"step{i}_{k}"] for k in ("cur", "act", "next_obs", "rew", "term", "cost", "dkl", "ep_var")}
             for i in range(n_steps)]
    B, T, E = g["start_obs"].shape[0], int(g["T"]), int(g["E"])
    assert n_steps == 3 and g["alive_at_break"].sum() > 0.8 * B
    wv, wc, prng = _golden_policy_values(24)
    v_inds, vc_inds = prng.integers(0, 3, B), prng.integers(0, 3, B)
    env = _ReplayEnv(steps, B, E, False)

    def value_fn(obs, n_dev, pid):
        n = int(n_dev.item())
        o, ids = obs[:n].cpu().numpy(), pid[:n].cpu().numpy()
        v, vc = np.zeros(B, np.float32), np.zeros(B, np.float32)
        if n:
            allv = np.sin(o @ wv[:3].T).T.astype(np.float32)
            allc = (0.5 * np.sin(o @ wc[:3].T).T.astype(np.float32) + 0.5)
            v[:n], vc[:n] = allv[v_inds[ids], np.arange(n)], allc[vc_inds[ids], np.arange(n)]
        return torch.from_numpy(v).cuda(), torch.from_numpy(vc).cuda()

    ro = cm.ModelRollout(env, T, rollout_mode="uncertainty", dkl_lim=float(g["dkl_lim"]), value_fn=value_fn)
    ro.reset(g["start_obs"])
    done, keep = ro.run_until(lambda cur, n_dev: env.policy(cur, n_dev), min_alive_ratio=0.0, max_steps=n_steps)
    assert done == n_steps and keep
    n_alive, ptr, total, finished = ro.counters()
    assert n_alive == int(g["alive_at_break"].sum()) and not finished and total == int(g["total_samples"])
    assert abs(ro.alive_ratio - n_alive / B) < 1e-12
    res = ro.finish(gamma=float(g["gamma"]), lam=float(g["lam"]), cost_gamma=float(g["cost_gamma"]), cost_lam=float(g["cost_lam"]),
                    cares_about_cost=True)
    assert ro.counters()[0] == 0 and ro.counters()[3] == 1          # finish_all_paths left nothing alive
    got = ro.get(numpy=True)
    close = lambda a, b: a.shape == b.shape and np.allclose(a, b, rtol=1e-6, atol=1e-7)
    assert np.array_equal(got["finish_ptr"], g["finish_ptr"])
    assert close(got["last_val"], g["last_val"]) and close(got["last_cval"], g["last_cval"])
    alive = g["alive_at_break"]
    assert (got["finish_ptr"][alive] == n_steps).all() and (got["last_val"][alive] != 0).all()
    assert np.array_equal(got["populated"], g["populated_mask"])
    for ours, ref in (("adv", "adv_buf"), ("ret", "ret_buf"), ("cadv", "cadv_buf"), ("cret", "cret_buf")):
        assert np.allclose(res[ours].cpu().numpy(), g[ref], rtol=2e-5, atol=2e-6), ours
    batch, stats = ro.get_batch()
    b = {k: v.cpu().numpy() for k, v in batch.items()}
    assert np.array_equal(b["obs"], g["get_obs"])
    for ours, ref in (("adv", "get_adv"), ("cadv", "get_cadv"), ("ret", "get_ret"), ("cret", "get_cret")):
        assert np.allclose(b[ours], g[ref], rtol=5e-5, atol=5e-6), ours
    # finish_all_paths on a rollout that ran to the end changes nothing
    fp = ro.finish_ptr.clone()
    ro.finish_all_paths()
    assert torch.equal(fp, ro.finish_ptr)


def test_graph_replays_draw_fresh_philox_streams():
    """The Philox stream base lives in a device word that every reset() advances (ModelRollout.rng_base): replays of one
    captured graph draw different elite members per (row, step) -- the host-side launch counters are frozen at capture
    time -- and restoring the word reproduces a replay bit for bit."""
    import torch
    import constrainedmbpo_b200 as cm
    env, B, T = "halfcheetah", 4000, 5
    p = make_params(env, HIDDEN, parity_scalers=False, bias=False)
    obs, _ = make_inputs(env, B)
    fe = build_cuda_env(p, env, HIDDEN, seed=3)
    ro = cm.ModelRollout(fe, T, rollout_mode="schedule")
    start = torch.from_numpy(obs).cuda()
    ro.capture(start, cm.StubPolicy(17, 6))
    ro.replay(); torch.cuda.synchronize()
    a = ro.pool["next_obs"].clone()
    base = ro.rng_base.clone()
    ro.replay(); torch.cuda.synchronize()
    b = ro.pool["next_obs"].clone()
    assert int(ro.rng_base) == int(base) + T
    changed = (a[:, 0] != b[:, 0]).any(dim=-1).float().mean().item()
    assert 0.6 < changed < 0.95, changed          # 5 elites: a row keeps its member with probability 1/5
    assert torch.equal(ro.pool["obs"][:, 0], torch.from_numpy(obs).cuda())      # same start states
    ro.rng_base.copy_(base)
    ro.replay(); torch.cuda.synchronize()
    assert torch.equal(ro.pool["next_obs"], b)
    # (the policy above is a temporary: capture() must keep it -- and its action buffer -- alive for the replays)
    assert ro._captured_refs[1] is not None


def test_iv_gae_graph_replays_shuffle_afresh():
    """'iv_gae' without an injected perm_fn: the forward_shuffle permutation is drawn on the device per step (keyed by
    the rollout's rng_base word), so replays of ONE captured graph use different permutations -- upstream draws
    np.random.permutation on every FakeEnv.step call (fake_env.py:314) -- and restoring the word reproduces a replay."""
    import torch
    import constrainedmbpo_b200 as cm
    env, B, T = "humanoidsafe", 1500, 4
    p = make_params(env, HIDDEN, parity_scalers=False, bias=False)
    obs, _ = make_inputs(env, B)
    fe = build_cuda_env(p, env, HIDDEN, cares_about_cost=True, seed=5)
    ro = cm.ModelRollout(fe, T, rollout_mode="iv_gae")
    pol = cm.StubPolicy(fe.obs_dim, fe.act_dim, device="cuda:0")
    ro.capture(torch.from_numpy(obs).cuda(), pol)
    ro.replay(); torch.cuda.synchronize()
    a = ro.pool["next_obs"].clone()
    base = ro.rng_base.clone()
    ro.replay(); torch.cuda.synchronize()
    b = ro.pool["next_obs"].clone()
    # slot j holds f_{p[j]}(x[p[p[j]]]): with another permutation (almost) every slot's first transition changes
    assert not torch.equal(a[:, :, 0], b[:, :, 0])
    ro.rng_base.copy_(base)
    ro.replay(); torch.cuda.synchronize()
    assert torch.equal(ro.pool["next_obs"], b)
