"""cmbpo_finish_paths (GAE / inverse-variance GAE) against the buffers the reference's REAL
ModelBuffer.finish_path_multiple wrote (tests/golden/sampler_*.npz) and against the oracle on larger seeded pools."""
import os

import numpy as np
import pytest

from tests.helpers import orc

pytestmark = pytest.mark.gpu
RTOL, ATOL = 2e-5, 2e-6   # fp64 arithmetic rounded to float32 buffers; reassociation only


def _golden(name):
    return np.load(os.path.join(os.path.dirname(__file__), "golden", f"sampler_{name}.npz"), allow_pickle=False)


def _run_cuda(rew, val, cost, cval, fin, lv, lc, mode, populated=None, **kw):
    import torch
    import constrainedmbpo_b200 as cm
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    out = cm.finish_paths(t(rew), t(val), t(cost), t(cval), t(fin.astype(np.int32)), t(lv), t(lc), mode,
                          populated=None if populated is None else t(populated.astype(np.uint8)), **kw)
    torch.cuda.synchronize()
    return {k: v.cpu().numpy() for k, v in out.items()}


@pytest.mark.parametrize("name", ["schedule", "uncertainty", "iv_gae"])
def test_finish_paths_matches_reference_modelbuffer_golden(name):
    g = _golden(name)
    T = int(g["T"])
    stored = np.arange(T)[None, :] < g["path_length"][:, None]
    out = _run_cuda(g["rew_buf"], g["val_buf"], g["cost_buf"], g["cval_buf"], g["finish_ptr"], g["last_val"], g["last_cval"],
                    name, populated=stored, gamma=float(g["gamma"]), lam=float(g["lam"]), cost_gamma=float(g["cost_gamma"]),
                    cost_lam=float(g["cost_lam"]), model_ind=int(g["model_ind"]), cares_about_cost=True,
                    max_uncertainty_r=float(g["max_uncertainty"][0]), max_uncertainty_c=float(g["max_uncertainty"][1]))
    keys = [("adv", "adv_buf"), ("ret", "ret_buf"), ("cadv", "cadv_buf"), ("cret", "cret_buf")]
    if name == "iv_gae":
        keys += [("ret_var", "ret_var_buf"), ("cret_var", "cret_var_buf"), ("roll_lengths", "roll_lengths_buf"),
                 ("croll_lengths", "croll_lengths_buf")]
    for ours, ref in keys:
        assert out[ours].shape == g[ref].shape
        assert np.allclose(out[ours], g[ref], rtol=RTOL, atol=ATOL), (ours, float(np.abs(out[ours] - g[ref]).max()))
    if name == "iv_gae":
        assert np.array_equal(out["populated"].astype(bool), g["populated_mask"])   # horizon trim incl. the argmax quirk


@pytest.mark.parametrize("mode,E,B,T", [("uncertainty", 1, 5000, 30), ("iv_gae", 7, 3000, 12), ("iv_gae", 3, 700, 33)])
def test_finish_paths_matches_oracle_on_seeded_pools(mode, E, B, T):
    """Ragged path lengths (incl. 0 and T-1), zero and non-zero bootstrap values, against oracle.finish_paths."""
    rng = np.random.default_rng(5)
    lead = (E,) if mode == "iv_gae" else ()
    f = lambda scale=1.0: (rng.standard_normal(lead + (B, T)) * scale).astype(np.float32)
    rew, val, cost, cval = f(), f(2.0), np.abs(f(0.3)), np.abs(f())
    fin = rng.integers(0, T, B)
    fin[:3] = (0, 1, T - 1)
    lv = (rng.standard_normal(lead + (B,)) * (rng.random(B) < 0.6)).astype(np.float32)
    lc = np.abs(rng.standard_normal(lead + (B,))).astype(np.float32)
    stored = np.arange(T)[None, :] < fin[:, None]
    kw = dict(gamma=0.99, lam=0.95, cost_gamma=0.97, cost_lam=0.9, model_ind=E // 2, cares_about_cost=True,
              max_uncertainty_r=1.5, max_uncertainty_c=1.8)
    n_or = min(B, 400)   # the oracle is a pure-Python loop: check a prefix of the paths
    ref = orc.finish_paths(rew[..., :n_or, :], val[..., :n_or, :], cost[..., :n_or, :], cval[..., :n_or, :], fin[:n_or],
                           lv[..., :n_or], lc[..., :n_or], mode, kw["gamma"], kw["lam"], kw["cost_gamma"], kw["cost_lam"],
                           model_ind=kw["model_ind"], cares_about_cost=True, max_uncertainty_r=1.5, max_uncertainty_c=1.8,
                           populated=stored[:n_or])
    out = _run_cuda(rew, val, cost, cval, fin, lv, lc, mode, populated=stored, **kw)
    names = ["adv", "ret", "cadv", "cret"] + (["ret_var", "cret_var", "roll_lengths", "croll_lengths"] if mode == "iv_gae" else [])
    for k in names:
        assert np.allclose(out[k][:n_or], ref[k], rtol=RTOL, atol=ATOL), (k, float(np.abs(out[k][:n_or] - ref[k]).max()))
        assert np.isfinite(out[k]).all()
        assert (out[k][~stored] == 0).all()                  # nothing written outside a path's slice
    if mode == "iv_gae":
        got, want = out["populated"][:n_or].astype(bool), ref["populated"]
        # the trim compares float32 ratios with a threshold: rows whose ratio sits within float32 rounding of it may flip
        assert (got != want).any(axis=1).sum() <= 1
        assert (out["populated"].astype(bool) <= stored).all()


@pytest.mark.parametrize("name", ["schedule", "uncertainty", "iv_gae"])
def test_normalize_advantages_matches_reference_get(name):
    """cmbpo_normalize_advantages on the reference's own adv / cadv buffers against what its real ModelBuffer.get
    returned (float32 statistics: summation order differs, so a few ulp)."""
    import torch
    import constrainedmbpo_b200 as cm
    g = _golden(name)
    pop = g["populated_mask"]
    t = lambda a, dt: torch.from_numpy(np.ascontiguousarray(a)).to("cuda", dt)
    adv, cadv = t(g["adv_buf"], torch.float32), t(g["cadv_buf"], torch.float32)
    stats = cm.normalize_advantages(adv, cadv, t(pop.astype(np.uint8), torch.uint8), ret=t(g["ret_buf"], torch.float32),
                                    cret=t(g["cret_buf"], torch.float32)).cpu().numpy()
    a, c = adv.cpu().numpy(), cadv.cpu().numpy()
    assert np.allclose(a[pop], g["get_adv"], rtol=1e-5, atol=1e-6) and np.allclose(c[pop], g["get_cadv"], rtol=1e-5, atol=1e-6)
    assert np.array_equal(a[~pop], g["adv_buf"][~pop]) and np.array_equal(c[~pop], g["cadv_buf"][~pop])   # untouched
    assert int(stats[5]) == int(g["diag_poolm_batch_size"])
    assert np.isclose(stats[3], float(g["diag_poolm_ret_mean"]), rtol=1e-5, atol=1e-7)
    assert np.isclose(stats[4], float(g["diag_poolm_cret_mean"]), rtol=1e-5, atol=1e-7)
    assert abs(a[pop].mean()) < 1e-5 and abs(a[pop].std() - 1) < 1e-4


def test_normalize_advantages_large_and_empty():
    import torch
    import constrainedmbpo_b200 as cm
    rng = np.random.default_rng(2)
    B, T = 200_000, 10
    adv = (rng.standard_normal((B, T)) * 3 + 5).astype(np.float32)
    cadv = (rng.standard_normal((B, T)) + 2).astype(np.float32)
    pop = rng.random((B, T)) < 0.4
    ra, rc, am, asd, cmn = orc.normalize_advantages(adv, cadv, pop)
    a, c = torch.from_numpy(adv).cuda().t().contiguous().t(), torch.from_numpy(cadv).cuda().t().contiguous().t()   # time-major views
    p = torch.from_numpy(pop.astype(np.uint8)).cuda().t().contiguous().t()
    stats = cm.normalize_advantages(a, c, p).cpu().numpy()
    assert np.allclose(a.cpu().numpy(), ra, rtol=2e-5, atol=2e-5) and np.allclose(c.cpu().numpy(), rc, rtol=2e-5, atol=2e-5)
    assert np.isclose(stats[0], am, rtol=1e-5) and np.isclose(stats[1], asd, rtol=1e-5) and np.isclose(stats[2], cmn, rtol=1e-5)
    z = torch.zeros(4, 3, device="cuda")
    s0 = cm.normalize_advantages(z.clone(), z.clone(), torch.zeros(4, 3, dtype=torch.uint8, device="cuda")).cpu().numpy()
    assert s0[5] == 0 and s0[1] == 0
