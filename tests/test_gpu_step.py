"""GPU parity of ``FakeEnv.step`` (mbpo/models/fake_env.py:131-222) through the C ABI vs the fp64 oracle,
with the RNG draws injected (elite indices, shuffle permutation, noise)."""
import os

import numpy as np
import pytest

from tests.helpers import (ATOL_SCALE_BF16, ENV_DIMS, RTOL_BF16, assert_means_close, build_cuda_env, make_inputs,
                           make_params, near_threshold_rows, orc, out_scale)

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
HIDDEN = (200, 200, 200, 200)


STATIC_COST = {"antsafe", "halfcheetahsafe", "humanoidsafe", "hoppersafe"}


def _check_step(env, B=3000, cares_about_cost=None, elites=(0, 1, 2, 3, 4)):
    if cares_about_cost is None:   # the CostNN branch (no static cost_f) is covered by test_step_costnn_path
        cares_about_cost = env in STATIC_COST
    p = make_params(env, HIDDEN, elites=elites)
    obs, act = make_inputs(env, B)
    fe = build_cuda_env(p, env, HIDDEN, cares_about_cost=cares_about_cost)
    idx = np.random.default_rng(7).choice(p.elites, size=B)
    no, rw, tm, info = fe.step(obs, act, elite_idx=idx)
    rno, rrw, rtm, rinfo = orc.fake_env_step(p, env, obs.astype(np.float64), act.astype(np.float64), elite_idx=idx,
                                             cares_about_cost=cares_about_cost)
    s = out_scale(p)
    # reference shapes / dtypes (fake_env.py:212-222)
    assert no.shape == (B, ENV_DIMS[env][0]) and rw.shape == (B, 1) and tm.shape == (B, 1)
    assert tm.dtype == np.bool_ and no.dtype == np.float32
    assert info["ensemble_dkl_path"].shape == (B,) and info["ensemble_ep_var"].shape == (B, ENV_DIMS[env][0] + 1)
    assert_means_close(no, rno, s[:-1], what="next_obs")
    assert_means_close(rw, rrw, s[-1:], what="reward")
    # decisions: must match except rows within the propagated tolerance of a threshold
    tol = RTOL_BF16 * np.abs(rno) + ATOL_SCALE_BF16 * s[:-1]
    near = near_threshold_rows(env, rno, tol)
    mism_t = (tm != rtm)[:, 0]
    assert not (mism_t & ~near).any(), f"{int((mism_t & ~near).sum())} termination mismatches away from thresholds"
    if cares_about_cost:
        mism_c = (info["cost"] != rinfo["cost"])[:, 0]
        assert not (mism_c & ~near).any()
    # ensemble statistics
    rel = np.abs(info["ensemble_dkl_path"] - rinfo["ensemble_dkl_path"]) / (np.abs(rinfo["ensemble_dkl_path"]) + 1e-6)
    assert np.median(rel) < 2e-2 and rel.max() < 0.15, f"dkl_path rel err median {np.median(rel):.3e} max {rel.max():.3e}"
    ev, rev = info["ensemble_ep_var"], rinfo["ensemble_ep_var"]
    dm = ATOL_SCALE_BF16 * s  # per-member mean error bound => bound on var_E(mean)
    assert (np.abs(ev - rev) <= 2 * np.sqrt(rev) * dm + dm ** 2).all()
    assert abs(info["ensemble_dkl_mean"] - rinfo["ensemble_dkl_mean"]) < 2e-2 * rinfo["ensemble_dkl_mean"]
    assert abs(info["ensemble_mean_var"] - rinfo["ensemble_mean_var"]) < 2e-2 * rinfo["ensemble_mean_var"]
    assert abs(info["rew_mean"] - rinfo["rew_mean"]) < 1e-2 * (abs(rinfo["rew_mean"]) + s[-1])
    # scalars are means of what was returned (checksum of the fused reduction)
    assert abs(info["ensemble_dkl_mean"] - info["ensemble_dkl_path"].astype(np.float64).mean()) < 1e-5 * (1 + abs(info["ensemble_dkl_mean"]))
    assert abs(info["rew_mean"] - rw.astype(np.float64).mean()) < 1e-5 * (1 + abs(rw).mean())
    assert abs(info["cost_mean"] - info["cost"].astype(np.float64).mean()) < 1e-6
    return int(near.sum()), int(mism_t.sum())


@pytest.mark.parametrize("env", ["antsafe", "ant", "halfcheetah", "halfcheetahsafe", "humanoid", "humanoidsafe",
                                 "hopper", "hoppersafe", "walker2d"])
def test_step_parity(env):
    near, mism = _check_step(env)
    print(f"{env}: rows near a threshold {near}, termination flips among them {mism}")


def test_step_duplicate_elites_injected():
    """Quirk Q7: the initial elite list may hold duplicates (bnn.py:98-102)."""
    _check_step("halfcheetah", B=1000, elites=(3, 3, 6, 0, 3))


def _zero_weight_env(env):
    cares_about_cost = env in STATIC_COST
    """Ensemble with zero weights / biases and identity-like output scaler: next_obs == obs exactly, so
    the fused static functions see precisely the golden inputs."""
    p = make_params(env, HIDDEN, parity_scalers=False, bias=False)
    for w in p.weights:
        w[...] = 0
    return p, build_cuda_env(p, env, HIDDEN, cares_about_cost=cares_about_cost)


@pytest.mark.parametrize("env", ["antsafe", "ant", "halfcheetah", "halfcheetahsafe", "humanoid", "humanoidsafe",
                                 "hopper", "hoppersafe", "walker2d"])
def test_static_fns_bit_exact_vs_reference_golden(env):
    """termination_fn / cost_f fused in the step kernel vs outputs of the REFERENCE's own
    mbpo/static/<env>.py (tests/golden/static_fns.npz), bit for bit, incl. quirks Q1 / Q2.  Rows with
    non-finite inputs are covered by test_step_nonfinite_rows (a NaN input poisons the whole row of
    a zero-weight GEMM, in the reference too)."""
    g = np.load(os.path.join(GOLDEN, "static_fns.npz"))
    nobs = g[f"{env}/next_obs"]
    act = g[f"{env}/act"]
    finite = np.isfinite(nobs).all(-1)
    p, fe = _zero_weight_env(env)
    idx = np.zeros(nobs.shape[0], np.int32)
    no, rw, tm, info = fe.step(nobs, act, elite_idx=idx)
    assert np.array_equal(no[finite], nobs[finite])          # mean = 0*1+0 (+obs) exactly
    assert np.array_equal(tm[finite], g[f"{env}/term"][finite])
    if f"{env}/cost" in g:
        assert np.array_equal(info["cost"][finite], g[f"{env}/cost"][finite].astype(np.float32))
    else:
        assert not info["cost"].any()
    # KAT: identical members => dkl = 0, ep_var = 0, var = s_out^2 = 1
    assert not info["ensemble_dkl_path"][finite].any()
    # var_E of 7 identical fp32 values: only the rounding of sum/7 survives
    assert (info["ensemble_ep_var"][finite] <= 1e-12 * (1 + np.c_[nobs, np.zeros(len(nobs))][finite] ** 2)).all()


def test_step_nonfinite_rows():
    """NaN / Inf propagate silently and surface through isfinite in the termination fns."""
    env = "antsafe"
    p = make_params(env, HIDDEN)
    obs, act = make_inputs(env, 64)
    obs[3, 7] = np.nan
    obs[5, 0] = np.inf
    act[9, 2] = -np.inf
    fe = build_cuda_env(p, env, HIDDEN, cares_about_cost=True)
    idx = np.random.default_rng(1).choice(p.elites, size=64)
    no, rw, tm, info = fe.step(obs, act, elite_idx=idx)
    with np.errstate(all="ignore"):
        rno, rrw, rtm, rinfo = orc.fake_env_step(p, env, obs.astype(np.float64), act.astype(np.float64),
                                                 elite_idx=idx, cares_about_cost=True)
    for r in (3, 5, 9):
        assert not np.isfinite(no[r]).all()
        assert tm[r, 0] == rtm[r, 0] and info["cost"][r, 0] == rinfo["cost"][r, 0]
    ok = np.ones(64, bool); ok[[3, 5, 9]] = False
    assert_means_close(no[ok], rno[ok], out_scale(p)[:-1])
    assert np.isnan(info["ensemble_dkl_path"][3]) == np.isnan(rinfo["ensemble_dkl_path"][3])


def test_step_kl_pair_clip_takes_the_exact_path():
    """The post kernel sums the pair KLs in O(E) form, which cannot apply the reference's per-pair clip to
    [0, 1e10] (utils.py:24); columns where a pair can reach the clip must fall back to the exact E x E path.
    One column gets log-variances of about +26 (member 0) and -64 (member 1): finite, inside the kernel's
    'moderate' range, and several pairs clip -- the unclipped sum would be ~1e21 instead of ~7e10."""
    env = "antsafe"
    p = make_params(env, HIDDEN)
    D = ENV_DIMS[env][0] + 1
    p.biases[-1][0, ..., D + 4] += 30
    p.biases[-1][1, ..., D + 4] -= 60
    obs, act = make_inputs(env, 300)
    fe = build_cuda_env(p, env, HIDDEN, cares_about_cost=True)
    idx = np.random.default_rng(2).choice(p.elites, size=300)
    no, rw, tm, info = fe.step(obs, act, elite_idx=idx)
    with np.errstate(all="ignore"):
        rno, rrw, rtm, rinfo = orc.fake_env_step(p, env, obs.astype(np.float64), act.astype(np.float64),
                                                 elite_idx=idx, cares_about_cost=True)
    ref = rinfo["ensemble_dkl_path"]
    assert ref.min() > 1e7   # the clipped pairs dominate
    rel = np.abs(info["ensemble_dkl_path"] - ref) / ref
    assert rel.max() < 2e-2, rel.max()
    assert_means_close(no, rno, out_scale(p)[:-1])


def test_step_3d_with_injected_permutation_quirk_q4():
    """3-D inputs (iv_gae rollouts): forward shuffle, and the 'inverse' shuffle that re-applies the same
    permutation (fake_env.py:309-324): slot j holds f_{p[j]}(x[p[p[j]]])."""
    env = "humanoidsafe"
    p = make_params(env, HIDDEN)
    E, B = 7, 500
    obs, act = make_inputs(env, B, E=E)
    perm = np.array([2, 0, 3, 1, 6, 4, 5])
    fe = build_cuda_env(p, env, HIDDEN, cares_about_cost=True)
    no, rw, tm, info = fe.step(obs, act, perm=perm)
    rno, rrw, rtm, rinfo = orc.fake_env_step(p, env, obs.astype(np.float64), act.astype(np.float64), perm=perm,
                                             cares_about_cost=True)
    s = out_scale(p)
    assert no.shape == (E, B, 45) and rw.shape == (E, B, 1) and tm.shape == (E, B, 1) and info["cost"].shape == (E, B, 1)
    assert_means_close(no, rno, s[:-1], what="next_obs 3d")
    assert_means_close(rw, rrw, s[-1:], what="reward 3d")
    tol = RTOL_BF16 * np.abs(rno) + ATOL_SCALE_BF16 * s[:-1]
    near = near_threshold_rows(env, rno, tol)
    assert not ((tm != rtm)[..., 0] & ~near).any()
    rel = np.abs(info["ensemble_dkl_path"] - rinfo["ensemble_dkl_path"]) / (np.abs(rinfo["ensemble_dkl_path"]) + 1e-6)
    assert np.median(rel) < 2e-2
    # identity permutation == no shuffle
    no_id, *_ = fe.step(obs, act, perm=np.arange(E))
    rno_id, *_ = orc.fake_env_step(p, env, obs.astype(np.float64), act.astype(np.float64), perm=np.arange(E))
    assert_means_close(no_id, rno_id, s[:-1])


def test_step_noise_injection_mode():
    """north_star's optional Gaussian sampling (off in the live reference, fake_env.py:164-170):
    sample = mean + noise * std of the drawn elite."""
    env = "halfcheetah"
    p = make_params(env, HIDDEN)
    B = 800
    obs, act = make_inputs(env, B)
    fe = build_cuda_env(p, env, HIDDEN)
    idx = np.random.default_rng(2).choice(p.elites, size=B)
    noise = np.random.default_rng(3).standard_normal((B, 18)).astype(np.float32)
    no, rw, tm, info = fe.step(obs, act, elite_idx=idx, noise=noise)
    rno, rrw, _, rinfo = orc.fake_env_step(p, env, obs.astype(np.float64), act.astype(np.float64), elite_idx=idx,
                                           noise=noise.astype(np.float64))
    s = out_scale(p)
    std_sel = np.sqrt(rinfo["_vars"][idx, np.arange(B)])
    bound = RTOL_BF16 * np.abs(rno) + ATOL_SCALE_BF16 * s[:-1] + 3e-2 * np.abs(noise[:, :-1]) * std_sel[:, :-1]
    assert (np.abs(no - rno) <= bound).all()
    zero, *_ = fe.step(obs, act, elite_idx=idx, noise=np.zeros_like(noise))
    plain, *_ = fe.step(obs, act, elite_idx=idx)
    assert np.array_equal(zero, plain)  # injected zeros == reference behaviour


def test_step_costnn_path():
    """cares_about_cost without a static cost_f: CostNN ensemble-mean prediction on concat(obs, act,
    next_obs) (fake_env.py:191-196)."""
    env = "halfcheetah"
    p = make_params(env, HIDDEN)
    B = 600
    obs, act = make_inputs(env, B)
    fe = build_cuda_env(p, env, HIDDEN, cares_about_cost=True)
    assert fe._cost_model is not None
    rng = np.random.default_rng(12)
    cp = orc.init_ensemble(rng, 2 * 17 + 6, 1, (64, 64), 7, probabilistic=False, dtype=np.float64)
    for b in cp.biases:
        b += rng.normal(0, 0.1, b.shape)
    fe._cost_model.load_state([w.astype(np.float32) for w in cp.weights], [b.astype(np.float32) for b in cp.biases])
    idx = np.random.default_rng(2).choice(p.elites, size=B)
    no, rw, tm, info = fe.step(obs, act, elite_idx=idx)
    # teacher-forced: the oracle's cost model sees the CUDA next_obs
    ref_cost = orc.predict(cp, np.concatenate([obs, act, no], -1).astype(np.float64), factored=False, inc_var=False)
    assert info["cost"].shape == (B, 1)
    assert_means_close(info["cost"], ref_cost, np.ones(1))
    assert abs(info["cost_mean"] - info["cost"].mean()) < 1e-5


def test_step_philox_elites_respect_list_and_sharding():
    """In-kernel Philox draw (no injected indices): every row's sample is the mean of a member of the
    elite list (with duplicate weighting, Q7), and the draw depends only on the global row id, so
    two half-batches with row offsets reproduce the full batch bit for bit."""
    import torch
    env = "halfcheetah"
    elites = (1, 1, 4, 6, 1)
    p = make_params(env, HIDDEN, elites=elites)
    B = 4000
    obs, act = make_inputs(env, B)
    fe = build_cuda_env(p, env, HIDDEN, seed=11)
    od, ad = torch.from_numpy(obs).cuda(), torch.from_numpy(act).cuda()
    fe._rng_offset = 5
    full = {k: v.clone() for k, v in fe.step_device(od, ad).items()}
    means, _ = fe._model.predict(np.concatenate([obs, act], -1), factored=True)
    means[:, :, :-1] += obs
    no = full["next_obs"].cpu().numpy()
    which = np.full(B, -1)
    for e in range(7):
        hit = np.all(no == means[e][:, :-1], axis=-1)
        which[hit & (which < 0)] = e
    assert (which >= 0).all() and set(np.unique(which)) <= {1, 4, 6}
    frac1 = (which == 1).mean()
    assert 0.55 < frac1 < 0.65, frac1   # 3 of the 5 list entries are member 1
    # sharding invariance through row_ids (global ids), as ModelRollout passes them
    half = B // 2
    outs = []
    for lo in (0, half):
        fe._rng_offset = 5
        ids = torch.arange(lo, lo + half, dtype=torch.int32, device="cuda")
        o = fe.step_device(od[lo:lo + half].contiguous(), ad[lo:lo + half].contiguous(), row_ids=ids)
        outs.append({k: v.clone() for k, v in o.items()})
    for k in ("next_obs", "rew", "term", "dkl_path"):
        cat = torch.cat([outs[0][k], outs[1][k]], dim=0)
        assert torch.equal(cat, full[k]), k


def test_step_rejects_1d_like_reference():
    env = "halfcheetah"
    p = make_params(env, HIDDEN)
    fe = build_cuda_env(p, env, HIDDEN)
    with pytest.raises(NameError):  # quirk Q3: fake_env.py:204-210
        fe.step(np.zeros(17, np.float32), np.zeros(6, np.float32))


@pytest.mark.parametrize("tag,env", [("antsafe2d", "antsafe"), ("humanoidsafe3d", "humanoidsafe"), ("hopper2d", "hopper"),
                                     ("halfcheetahsafe2d", "halfcheetahsafe"), ("ant2d_costnn", "ant"),
                                     ("humanoid3d_costnn", "humanoid")])
def test_step_against_reference_fake_env_golden(tag, env):
    """The CUDA FakeEnv.step against what the reference's REAL FakeEnv.step returned (tests/golden/fakeenv_step.npz:
    real fake_env.py + real StaticFns, model outputs from the fp32 restatement of the TF graph) for the same ensemble,
    inputs and injected draws: means within the bf16 tolerance, decisions equal away from thresholds."""
    from tests.test_oracle_golden import load_fakeenv_golden
    g = load_fakeenv_golden(tag)
    hidden = (24, 24)
    p = make_params(env, hidden, parity_scalers=True, bias=True, dtype=np.float32)
    cost = bool(g["cares_about_cost"])
    fe = build_cuda_env(p, env, hidden, cares_about_cost=cost)
    three_d = g["obs"].ndim == 3
    costnn = tag.endswith("costnn")
    if costnn:   # the learned cost head of the fixture (fake_env.py:76-94): same seeded 64x64 MSE ensemble
        od, ad = g["obs"].shape[-1], g["act"].shape[-1]
        pc = orc.init_ensemble(np.random.default_rng(17), 2 * od + ad, 1, (64, 64), 7, probabilistic=False,
                               use_scaler_in=False, use_scaler_out=False, elites=list(range(5)), dtype=np.float32)
        fe._cost_model.load_state(pc.weights, pc.biases, elites=pc.elites)
    no, rw, tm, info = fe.step(g["obs"], g["act"], elite_idx=None if three_d else g["idx"], perm=g["perm"] if three_d else None)
    s = out_scale(p)
    assert no.shape == g["next_obs"].shape and rw.shape == g["rew"].shape and tm.shape == g["term"].shape
    # 24-wide nets fed inputs at the scaler floor (two input dims have var 1e-6 => normalised magnitudes ~1e2, whose
    # bf16 rounding is ~0.5 absolute) see less averaging than the 200-wide nets of test_step_parity: the absolute
    # companion of rtol 1e-2 is 4e-2 * s_out here (observed worst 3.5e-2, 1 element in 20160 above 2e-2)
    atol = 4e-2
    assert_means_close(no, g["next_obs"], s[:-1], atol_scale=atol, what="next_obs vs reference FakeEnv.step")
    assert_means_close(rw, g["rew"], s[-1:], atol_scale=atol, what="reward vs reference FakeEnv.step")
    tol = RTOL_BF16 * np.abs(g["next_obs"].astype(np.float64)) + atol * s[:-1]
    near = near_threshold_rows(env, g["next_obs"], tol)
    mism = (tm != g["term"])[..., 0]
    assert not (mism & ~near).any(), f"{int((mism & ~near).sum())} termination mismatches away from thresholds"
    if cost and costnn:    # continuous output of the cost ensemble's mean (fake_env.py:196), unit scale
        assert info["cost"].shape == g["cost"].shape
        assert (np.abs(info["cost"] - g["cost"]) <= 1e-2 * np.abs(g["cost"]) + 2e-2).all()
    elif cost:
        mc = (info["cost"] != g["cost"])[..., 0]
        assert not (mc & ~near).any()
    rel = np.abs(info["ensemble_dkl_path"] - g["dkl_path"]) / (np.abs(g["dkl_path"]) + 1e-6)
    assert np.median(rel) < 2e-2, f"dkl_path median rel err {np.median(rel):.3e}"
    assert abs(info["ensemble_mean_var"] - g["mean_var"]) < 2e-2 * g["mean_var"]
    assert abs(info["rew_mean"] - g["rew_mean"]) < 1e-2 * (abs(g["rew_mean"]) + s[-1])


def test_step_philox_gaussian_sampling_mode():
    """sample_mode 1 (north_star's fused Philox Gaussian sampling; off in the live reference): the kernel's draws are
    reproduced in NumPy (Philox4x32-10 + Box-Muller keyed by global row id and column) and injected into the oracle;
    draws do not depend on where a row sits in the batch."""
    from tests.helpers import philox_normals
    env, B = "halfcheetah", 4096
    p = make_params(env, HIDDEN)
    obs, act = make_inputs(env, B)
    import constrainedmbpo_b200 as cm
    fe = cm.FakeEnv(cm.SimpleEnv(17, 6), env, num_networks=7, num_elites=5, hidden_dims=HIDDEN, seed=1234, sample_noise=True)
    fe._model.load_state([w.astype(np.float32) for w in p.weights], [b.astype(np.float32) for b in p.biases],
                         scaler_in=(p.mu_in, p.var_in), scaler_out=(p.mu_out, p.var_out), elites=p.elites)
    idx = np.random.default_rng(7).choice(p.elites, size=B)
    stream = fe._rng_offset
    no, rw, tm, info = fe.step(obs, act, elite_idx=idx)
    z = philox_normals(np.arange(B), 18, stream, fe._seed)
    assert abs(z.mean()) < 0.02 and abs(z.std() - 1) < 0.02
    rno, rrw, _, rinfo = orc.fake_env_step(p, env, obs.astype(np.float64), act.astype(np.float64), elite_idx=idx, noise=z)
    s = out_scale(p)
    std_sel = np.sqrt(rinfo["_vars"][idx, np.arange(B)])
    bound = RTOL_BF16 * np.abs(rno) + ATOL_SCALE_BF16 * s[:-1] + 3e-2 * np.abs(z[:, :-1]) * std_sel[:, :-1]
    assert (np.abs(no - rno) <= bound).all(), float((np.abs(no - rno) - bound).max())
    assert np.abs(rw - rrw).max() <= RTOL_BF16 * np.abs(rrw).max() + ATOL_SCALE_BF16 * s[-1] + 3e-2 * np.abs(z[:, -1:] * std_sel[:, -1:]).max()
    # the noise really is there (the deterministic step differs by about one std)
    fe2 = build_cuda_env(p, env, HIDDEN)
    det, *_ = fe2.step(obs, act, elite_idx=idx)
    ratio = np.abs(no - det) / std_sel[:, :-1]
    assert 0.6 < np.median(ratio) / 0.6745 < 1.4          # median |N(0,1)| = 0.6745
    # shard invariance: rows [1000, 2000) stepped alone with their global ids draw the same normals
    import torch
    fe._rng_offset = stream
    o = fe.step_device(torch.from_numpy(obs[1000:2000]).cuda(), torch.from_numpy(act[1000:2000]).cuda(),
                       elite_idx=torch.from_numpy(idx[1000:2000].astype(np.int32)),
                       row_ids=torch.arange(1000, 2000, dtype=torch.int32, device="cuda"))
    assert np.array_equal(o["next_obs"].cpu().numpy(), no[1000:2000])


def test_step_logvar_soft_bounding_option():
    """bound_logvar (PETS soft bounds, commented out upstream at bnn.py:888-889): variances and everything derived from
    them follow the oracle's bounded form; means are untouched."""
    env, B = "antsafe", 2000
    p = make_params(env, HIDDEN)
    obs, act = make_inputs(env, B)
    import constrainedmbpo_b200 as cm
    bounds = (0.5, -10.0)                      # FakeEnv's max_logvar / min_logvar (fake_env.py:61-75)
    fe = cm.FakeEnv(cm.SimpleEnv(29, 8), env, num_networks=7, num_elites=5, hidden_dims=HIDDEN, seed=0, logvar_bound=bounds)
    fe._model.load_state([w.astype(np.float32) for w in p.weights], [b.astype(np.float32) for b in p.biases],
                         scaler_in=(p.mu_in, p.var_in), scaler_out=(p.mu_out, p.var_out), elites=p.elites)
    idx = np.random.default_rng(7).choice(p.elites, size=B)
    no, rw, tm, info = fe.step(obs, act, elite_idx=idx)
    rno, _, _, rb = orc.fake_env_step(p, env, obs.astype(np.float64), act.astype(np.float64), elite_idx=idx, logvar_bound=bounds)
    _, _, _, ru = orc.fake_env_step(p, env, obs.astype(np.float64), act.astype(np.float64), elite_idx=idx)
    assert_means_close(no, rno, out_scale(p)[:-1], what="next_obs (bounded logvar)")
    assert abs(info["ensemble_mean_var"] - rb["ensemble_mean_var"]) < 2e-2 * rb["ensemble_mean_var"]
    assert abs(rb["ensemble_mean_var"] - ru["ensemble_mean_var"]) > 0.2 * ru["ensemble_mean_var"]   # the bound bites
    rel = np.abs(info["ensemble_dkl_path"] - rb["ensemble_dkl_path"]) / (np.abs(rb["ensemble_dkl_path"]) + 1e-6)
    assert np.median(rel) < 2e-2


@pytest.mark.parametrize("hidden", [HIDDEN, (256, 256), (512, 512)])
def test_step_3d_device_drawn_permutation(hidden):
    """perm="device": np.random.permutation(E) of forward_shuffle (fake_env.py:314) drawn by cmbpo_draw_permutation --
    the kernel's Fisher-Yates over Philox words restated in tests/helpers.philox_permutation -- and consumed by the
    forward kernels from device memory; the step then equals the one with the same permutation injected from the host.
    All three forward kernels' slot maps: ensemble_fwd_kernel, fused_wide_kernel with two row-tile streams and with one."""
    import torch
    from tests.helpers import philox_permutation
    env = "humanoidsafe"
    p = make_params(env, hidden)
    E, B = 7, 700
    obs, act = make_inputs(env, B, E=E)
    fe = build_cuda_env(p, env, hidden, cares_about_cost=True, seed=11)
    o_d, a_d = torch.from_numpy(obs).cuda(), torch.from_numpy(act).cuda()
    seen = set()
    for call in range(6):
        stream = fe._rng_offset
        out = fe.step_device(o_d, a_d, perm="device", want_ep_var=False)
        got = {k: v.clone() for k, v in out.items() if v is not None}
        perm = fe._perm_dev.cpu().numpy()
        assert sorted(perm.tolist()) == list(range(E))
        assert np.array_equal(perm, philox_permutation(E, fe._seed, stream))
        seen.add(tuple(perm.tolist()))
        ref = fe.step_device(o_d, a_d, perm=perm, want_ep_var=False)
        for k in ("next_obs", "rew", "term", "cost", "dkl_path"):
            assert torch.equal(got[k], ref[k]), (call, k)
        # a CUDA tensor permutation takes the same path
        ref2 = fe.step_device(o_d, a_d, perm=torch.from_numpy(perm).cuda(), want_ep_var=False)
        assert torch.equal(got["next_obs"], ref2["next_obs"])
    assert len(seen) >= 5                                    # 7! = 5040 permutations: six draws practically never collide
    # the device word is added to the stream id (what a captured rollout advances per replay)
    base = torch.tensor([1000], dtype=torch.int64, device="cuda")
    stream = fe._rng_offset
    fe.step_device(o_d, a_d, perm="device", want_ep_var=False, rng_offset_dev=base)
    assert np.array_equal(fe._perm_dev.cpu().numpy(), philox_permutation(E, fe._seed, stream + 1000))


def test_draw_permutation_is_uniform():
    import ctypes as C
    import torch
    from constrainedmbpo_b200 import _lib
    lib = _lib.load()
    n, draws = 4, 6000
    buf = torch.empty((draws, n), dtype=torch.int32, device="cuda")
    s = torch.cuda.current_stream().cuda_stream
    for i in range(draws):
        _lib.check(lib.cmbpo_draw_permutation(n, 99, i, None, 0, buf[i].data_ptr(), s))
    perms = buf.cpu().numpy()
    assert (np.sort(perms, 1) == np.arange(n)).all()
    codes = (perms * (n ** np.arange(n))).sum(1)
    _, counts = np.unique(codes, return_counts=True)
    assert counts.size == 24
    expect = draws / 24
    chi2 = ((counts - expect) ** 2 / expect).sum()
    assert chi2 < 60, chi2                                   # 23 dof: P(chi2 > 60) ~ 3e-5
