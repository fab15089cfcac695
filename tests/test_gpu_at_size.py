"""GPU parity AT THE BENCHMARKED SIZES (BASELINE.json configs: 10 k / 100 k / 1 M rows).

The small-batch tests (<= 5 000 rows) only ever drive the forward kernel through its "G = 7 member groups,
one tile (pair) per virtual cluster" split (ensemble_fwd.cu, work split).  What ``bench.py`` times is a different
walk through the same kernel: 100 k rows -> G = 7 with ~20 tile pairs per cluster (weight-ring wrap, barrier phase
wrap, XU token hand-over across pairs); 1 M rows -> G = 1, every cluster visits all 7 members of ~27 pairs (A0 tile
reused across members).  Each test here therefore checks, at the benchmarked row count,

  (a) ~4-5 k sampled rows -- first / last / ragged-tail tiles and random rows out of every cluster's tile list, for
      both the 74-cluster and the 10-virtual-cluster partition -- against the fp64 oracle with the standard tolerance;
  (b) ALL rows bit for bit against the same rows evaluated in 2 000-row batches (the kernels are row-independent and
      deterministic; the small-batch walk is the one the rest of the suite pins to the oracle).

All comparisons of (b) run on the device; only the sampled rows travel to the host.
"""
import numpy as np
import pytest

from tests.helpers import (ATOL_SCALE_BF16, ENV_DIMS, RTOL_BF16, VAR_RTOL_BF16, assert_means_close, build_cuda_env,
                           make_inputs, make_params, near_threshold_rows, orc, out_scale)
from tests.test_gpu_predict import _cuda_model

pytestmark = pytest.mark.gpu
HIDDEN = (200, 200, 200, 200)
CHUNK = 2000


def sample_rows(B, per_list=40, seed=0):
    """Row sample that touches every cluster's tile list (256-row tiles; cluster c owns tiles c, c+C, c+2C, ...;
    C = 74 CTA pairs at G = 1, 10 virtual clusters at G = 7), plus whole first / last tiles and the ragged tail."""
    rng = np.random.default_rng(seed)
    n_tiles = (B + 255) // 256
    rows = [np.arange(min(B, 256)), np.arange(max(0, B - 300), B)]
    for C in (74, 10):
        for c in range(min(C, n_tiles)):
            tiles = np.arange(c, n_tiles, C)
            # the first pair, the last pair (a lone X tile when the list is odd) and two random ones
            pick = np.unique(np.r_[tiles[:2], tiles[-2:], rng.choice(tiles, size=min(2, len(tiles)))])
            for t in pick:
                r = t * 256 + rng.choice(256, size=max(2, per_list // len(pick)), replace=False)
                rows.append(r[r < B])
    return np.unique(np.concatenate(rows))


def _all_rows_equal_chunked(full, chunk_fn, B, row_axis):
    """torch.equal of every CHUNK-row slice of the full-batch outputs with the outputs of the same rows alone."""
    import torch
    bad = []
    for lo in range(0, B, CHUNK):
        hi = min(B, lo + CHUNK)
        part = chunk_fn(lo, hi)
        for k, v in part.items():
            ref = full[k].narrow(row_axis[k], lo, hi - lo)
            if not torch.equal(ref, v):
                bad.append((k, lo, int((ref != v).sum().item())))
        if len(bad) > 8:
            break
    assert not bad, f"full-batch rows differ from the same rows in {CHUNK}-row batches: {bad[:8]}"


@pytest.mark.parametrize("env,B", [("antsafe", 10_000), ("antsafe", 100_000), ("halfcheetah", 1_000_000),
                                   ("antsafe", 1_000_003)])
def test_predict_at_benchmark_size(env, B):
    """BNN.predict (bnn.py:711-717) at configs[0] (10 k), configs[1] (100 k) and configs[2] (1 M) row counts."""
    import torch
    p = make_params(env, HIDDEN)
    obs, act = make_inputs(env, B)
    x = torch.from_numpy(np.concatenate([obs, act], -1)).cuda()
    m = _cuda_model(p, HIDDEN, "MSPE")
    mean, var = m.predict_device([x], 2, 0)
    torch.cuda.synchronize()
    full = {"mean": mean.clone(), "var": var.clone()}

    def chunk(lo, hi):
        a, b = m.predict_device([x[lo:hi].contiguous()], 2, 0)
        return {"mean": a, "var": b}
    _all_rows_equal_chunked(full, chunk, B, {"mean": 1, "var": 1})
    rows = sample_rows(B)
    ridx = torch.from_numpy(rows).cuda()
    got_m, got_v = full["mean"][:, ridx].cpu().numpy(), full["var"][:, ridx].cpu().numpy()
    rm, rv = orc.predict(p, x[ridx].cpu().numpy().astype(np.float64), factored=True)
    assert_means_close(got_m, rm, out_scale(p), what=f"mean @ {B} rows ({len(rows)} sampled)")
    rel = np.abs(got_v - rv) / rv
    assert rel.max() < VAR_RTOL_BF16, f"variance rel err max {rel.max():.3e}"
    assert m._lib.cmbpo_status(m._handle, None) == 0


@pytest.mark.parametrize("env,B", [("antsafe", 100_000), ("halfcheetah", 1_000_000), ("antsafe", 1_000_003)])
def test_step_at_benchmark_size(env, B):
    """FakeEnv.step (fake_env.py:131-222) at the configs[1] / configs[2] row counts: forward + output kernel."""
    import torch
    cost = env == "antsafe"
    p = make_params(env, HIDDEN)
    obs, act = make_inputs(env, B)
    fe = build_cuda_env(p, env, HIDDEN, cares_about_cost=cost)
    idx = np.random.default_rng(7).choice(p.elites, size=B).astype(np.int32)
    od, ad, idd = torch.from_numpy(obs).cuda(), torch.from_numpy(act).cuda(), torch.from_numpy(idx).cuda()
    keys = ("next_obs", "rew", "term", "cost", "dkl_path", "ep_var_mean", "ep_var")
    o = fe.step_device(od, ad, elite_idx=idd)
    torch.cuda.synchronize()
    full = {k: o[k].clone() for k in keys}
    scal = o["scalars"].cpu().numpy().astype(np.float64)

    def chunk(lo, hi):
        c = fe.step_device(od[lo:hi].contiguous(), ad[lo:hi].contiguous(), elite_idx=idd[lo:hi].contiguous())
        return {k: c[k] for k in keys}
    _all_rows_equal_chunked(full, chunk, B, {k: 0 for k in keys})
    # the fused scalar diagnostics are means of what was written (fake_env.py:213-220)
    assert abs(scal[0] - full["dkl_path"].double().mean().item()) < 1e-5 * (1 + abs(scal[0]))
    assert abs(scal[2] - full["rew"].double().mean().item()) < 1e-5 * (1 + abs(scal[2]))
    assert abs(scal[3] - full["cost"].double().mean().item()) < 1e-6
    # sampled rows vs the fp64 oracle
    rows = sample_rows(B)
    ridx = torch.from_numpy(rows).cuda()
    rno, rrw, rtm, rinfo = orc.fake_env_step(p, env, obs[rows].astype(np.float64), act[rows].astype(np.float64),
                                             elite_idx=idx[rows], cares_about_cost=cost)
    s = out_scale(p)
    no = full["next_obs"][ridx].cpu().numpy()
    assert_means_close(no, rno, s[:-1], what=f"next_obs @ {B} rows")
    assert_means_close(full["rew"][ridx].cpu().numpy(), rrw, s[-1:], what=f"reward @ {B} rows")
    tol = RTOL_BF16 * np.abs(rno) + ATOL_SCALE_BF16 * s[:-1]
    near = near_threshold_rows(env, rno, tol)
    mism = (full["term"][ridx].cpu().numpy().astype(bool) != rtm)[:, 0]
    assert not (mism & ~near).any(), f"{int((mism & ~near).sum())} termination mismatches away from thresholds"
    if cost:
        mc = (full["cost"][ridx].cpu().numpy() != rinfo["cost"])[:, 0]
        assert not (mc & ~near).any()
    dk = full["dkl_path"][ridx].cpu().numpy()
    rel = np.abs(dk - rinfo["ensemble_dkl_path"]) / (np.abs(rinfo["ensemble_dkl_path"]) + 1e-6)
    assert np.median(rel) < 2e-2 and rel.max() < 0.15, f"dkl_path rel err median {np.median(rel):.3e} max {rel.max():.3e}"
    assert fe._model._lib.cmbpo_status(fe._model._handle, None) == 0


def test_step_3d_iv_gae_at_benchmark_size():
    """configs[3]: Humanoid-shaped 3-D ('iv_gae') step at 125 k rows per member (1 M over 8 GPUs): per-slot inputs
    (a fresh A0 tile per member instead of one reused across the member loop) and step_post3d_kernel."""
    import torch
    env, E, B = "humanoidsafe", 7, 125_000
    p = make_params(env, HIDDEN)
    obs, act = make_inputs(env, B, E=E)
    fe = build_cuda_env(p, env, HIDDEN, cares_about_cost=True)
    perm = np.array([2, 0, 3, 1, 6, 4, 5])
    od, ad = torch.from_numpy(obs).cuda(), torch.from_numpy(act).cuda()
    keys = ("next_obs", "rew", "term", "cost", "dkl_path", "ep_var_mean")
    axis = {"next_obs": 1, "rew": 1, "term": 1, "cost": 1, "dkl_path": 0, "ep_var_mean": 0}
    o = fe.step_device(od, ad, perm=perm, want_ep_var=False)
    torch.cuda.synchronize()
    full = {k: o[k].clone() for k in keys}

    def chunk(lo, hi):
        c = fe.step_device(od[:, lo:hi].contiguous(), ad[:, lo:hi].contiguous(), perm=perm, want_ep_var=False)
        return {k: c[k] for k in keys}
    _all_rows_equal_chunked(full, chunk, B, axis)
    rows = sample_rows(B, per_list=16)
    ridx = torch.from_numpy(rows).cuda()
    rno, rrw, rtm, rinfo = orc.fake_env_step(p, env, obs[:, rows].astype(np.float64), act[:, rows].astype(np.float64),
                                             perm=perm, cares_about_cost=True)
    s = out_scale(p)
    assert_means_close(full["next_obs"][:, ridx].cpu().numpy(), rno, s[:-1], what="next_obs 3d @ 125k rows")
    assert_means_close(full["rew"][:, ridx].cpu().numpy(), rrw, s[-1:], what="reward 3d @ 125k rows")
    tol = RTOL_BF16 * np.abs(rno) + ATOL_SCALE_BF16 * s[:-1]
    near = near_threshold_rows(env, rno, tol)
    assert not ((full["term"][:, ridx].cpu().numpy().astype(bool) != rtm)[..., 0] & ~near).any()
    dk = full["dkl_path"][ridx].cpu().numpy()
    rel = np.abs(dk - rinfo["ensemble_dkl_path"]) / (np.abs(rinfo["ensemble_dkl_path"]) + 1e-6)
    assert np.median(rel) < 2e-2


@pytest.mark.parametrize("hidden,B", [((512, 512), 150_000), ((256, 256, 256, 256), 140_000)])
def test_predict_wide_beyond_one_slab(hidden, B):
    """The layer-at-a-time kernel above its 131 072-row slab size with E = 7 (the shipped (512, 512) nets,
    examples/config/antsafe/cmbpo_uc.py:36, and the 256-wide point of the configs[4] sweep)."""
    import torch
    from tests.test_gpu_wide import _atol
    env = "antsafe"
    p = make_params(env, hidden)
    obs, act = make_inputs(env, B)
    x = torch.from_numpy(np.concatenate([obs, act], -1)).cuda()
    m = _cuda_model(p, hidden, "MSPE")
    mean, var = m.predict_device([x], 2, 0)
    torch.cuda.synchronize()
    full = {"mean": mean.clone(), "var": var.clone()}

    def chunk(lo, hi):
        a, b = m.predict_device([x[lo:hi].contiguous()], 2, 0)
        return {"mean": a, "var": b}
    _all_rows_equal_chunked(full, chunk, B, {"mean": 1, "var": 1})
    rows = sample_rows(B, per_list=12)
    # the slab boundary itself
    rows = np.unique(np.r_[rows, np.arange(131072 - 130, 131072 + 130)])
    ridx = torch.from_numpy(rows).cuda()
    rm, rv = orc.predict(p, x[ridx].cpu().numpy().astype(np.float64), factored=True)
    assert_means_close(full["mean"][:, ridx].cpu().numpy(), rm, out_scale(p), atol_scale=_atol(hidden),
                       what=f"wide {hidden} mean @ {B} rows")
    rel = np.abs(full["var"][:, ridx].cpu().numpy() - rv) / rv
    # two-hidden-layer nets pass the rounding of the floor-scaled inputs almost directly to the log-variance head
    # (see tests/test_gpu_wide.py _atol): 5e-2 there, the standard bound for the 4-layer nets
    assert rel.max() < (5e-2 if len(hidden) < 3 else VAR_RTOL_BF16), f"variance rel err max {rel.max():.3e}"


@pytest.mark.parametrize("hidden,B,T,n_shards", [(HIDDEN, 100_000, 30, 10), ((512, 512), 30_000, 12, 3)])
def test_full_rollout_at_benchmark_size_matches_chunked_rollouts(hidden, B, T, n_shards):
    """The captured-graph rollout bench.py times (configs[1]: 100 k start states, 'uncertainty', H = 30) against the
    same rollout run as 10 independent shards of 10 k rows with their global row ids: every pool array must be
    bit-identical (row independence of the whole sampler step, shrinking live row counts included).  Second case: the
    shipped (512, 512) shape on the fused wide kernel (device-side live row counts, CTA pairs that run out of work)."""
    import torch
    import constrainedmbpo_b200 as cm
    env = "antsafe"
    p = make_params(env, hidden, parity_scalers=False, bias=False)
    fe = build_cuda_env(p, env, hidden, cares_about_cost=True, seed=5)
    obs, _ = make_inputs(env, B)
    obs[:, 0] = np.random.default_rng(3).uniform(0.3, 0.9, B)
    start = torch.from_numpy(obs).cuda()
    policy = cm.StubPolicy(29, 8, seed=4)
    cal = start[:5000].contiguous()
    fe._rng_offset = 0
    dkl_lim = cm.compute_dynamics_dkl(fe, policy, cal, depth=5)
    fe._rng_offset = 100
    ro = cm.ModelRollout(fe, T, rollout_mode="uncertainty", dkl_lim=dkl_lim)
    ro.reset(start)
    ro.run(policy)
    torch.cuda.synchronize()
    n_alive, ptr, total, finished = ro.counters()
    assert finished == 1 and total > B
    full = {k: v.clone() for k, v in ro.pool.items()}
    shard = B // n_shards
    tot = 0
    for r in range(n_shards):
        fe._rng_offset = 100
        rs = cm.ModelRollout(fe, T, rollout_mode="uncertainty", dkl_lim=dkl_lim, row_offset=r * shard)
        rs.reset(start[r * shard:(r + 1) * shard].contiguous())
        rs.run(policy)
        tot += rs.counters()[2]
        for k, v in rs.pool.items():
            assert torch.equal(full[k][r * shard:(r + 1) * shard], v), (k, r)   # logical [B, T, .] views
    assert tot == total
