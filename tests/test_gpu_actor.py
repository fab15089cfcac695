"""The CPO actor and critics inside the device rollout loop: Gaussian policy head
(softlearning/models/ac_network.py:105-149) on a 256 x 256 tanh MLP, V / VC for 'iv_gae' states, and a captured rollout
whose graph contains no stub -- actor, critics, dynamics step and bookkeeping are all kernels of libcmbpo_b200.so."""
import numpy as np
import pytest

from tests.helpers import build_cuda_env, make_inputs, make_params, orc

pytestmark = pytest.mark.gpu
HIDDEN = (200, 200, 200, 200)


def _actor_params(obs_dim, act_dim, hidden, seed):
    r = np.random.default_rng(seed)
    dims = [obs_dim] + list(hidden) + [act_dim]
    W = [(r.standard_normal((a, b)) / np.sqrt(a)).astype(np.float32) for a, b in zip(dims[:-1], dims[1:])]
    b = [(0.1 * r.standard_normal(n)).astype(np.float32) for n in dims[1:]]
    log_std = r.uniform(-1.0, 0.0, act_dim).astype(np.float32)
    return W, b, log_std


@pytest.mark.parametrize("hidden", [(256, 256), (64, 64)])
def test_actor_outs_match_oracle_with_injected_noise(hidden):
    import constrainedmbpo_b200 as cm
    obs_dim, act_dim, B = 29, 8, 3000
    W, b, log_std = _actor_params(obs_dim, act_dim, hidden, 3)
    actor = cm.GaussianActor(obs_dim, act_dim, hidden_sizes=hidden, seed=1)
    actor.load_state(W, b, log_std)
    obs, _ = make_inputs("antsafe", B, seed=5)
    noise = np.random.default_rng(6).standard_normal((B, act_dim)).astype(np.float32)
    out = actor.get_action_outs(obs, noise=noise)
    pi, logp, mu = orc.actor_outs(W, b, log_std, obs.astype(np.float64), noise.astype(np.float64))
    assert out["pi"].shape == (B, act_dim) and out["logp_pi"].shape == (B,) and out["pi_info"]["log_std"].shape == (B, act_dim)
    # bf16 operands: mu within 1e-2 relative + 1e-2 absolute (outputs are O(1)); the head itself is fp32
    assert (np.abs(out["pi_info"]["mu"] - mu) <= 1e-2 * np.abs(mu) + 2e-2).all()
    assert (np.abs(out["pi"] - pi) <= 1e-2 * np.abs(mu) + 2e-2).all()
    # logp_pi depends on (pi - mu) only: exact arithmetic of the injected noise, independent of the GEMM precision
    assert np.allclose(out["logp_pi"], logp, rtol=1e-5, atol=1e-4)
    assert np.array_equal(out["pi_info"]["log_std"][7], log_std)
    # deterministic policy: pi = mu
    actor.deterministic = True
    det = actor.get_action_outs(obs)
    assert np.array_equal(det["pi"], det["pi_info"]["mu"])


def test_actor_3d_states_share_one_draw_per_row_and_philox_is_shard_invariant():
    import torch
    import constrainedmbpo_b200 as cm
    obs_dim, act_dim, E, B = 45, 17, 7, 1200
    W, b, log_std = _actor_params(obs_dim, act_dim, (256, 256), 8)
    actor = cm.GaussianActor(obs_dim, act_dim, seed=11)
    actor.load_state(W, b, log_std)
    obs, _ = make_inputs("humanoidsafe", B, E=E)
    noise = np.random.default_rng(2).standard_normal((B, act_dim)).astype(np.float32)
    out = actor.get_action_outs(obs, noise=noise)
    pi, logp, mu = orc.actor_outs(W, b, log_std, obs.astype(np.float64), noise.astype(np.float64))
    assert out["pi"].shape == (E, B, act_dim) and out["logp_pi"].shape == (E, B)
    assert (np.abs(out["pi"] - pi) <= 1e-2 * np.abs(mu) + 2e-2).all()
    z = (out["pi"] - out["pi_info"]["mu"]) / np.exp(log_std)
    assert np.allclose(z, np.broadcast_to(noise, z.shape), atol=2e-5)            # the same draw in every slot (ac_network.py:135-138)
    # in-kernel Philox normals: N(0,1), keyed by the global row id -> rows [400, 800) alone draw the same numbers
    o2 = torch.from_numpy(obs[0]).cuda()
    actor._rng_offset = 9
    full = {k: (v.clone() if torch.is_tensor(v) else v) for k, v in actor.get_action_outs_device(o2).items()}
    zz = ((full["pi"] - full["pi_info"]["mu"]) / torch.exp(actor.log_std)).cpu().numpy()
    assert abs(zz.mean()) < 0.03 and abs(zz.std() - 1) < 0.03
    actor._rng_offset = 9
    part = actor.get_action_outs_device(o2[400:800].contiguous(), row_ids=torch.arange(400, 800, dtype=torch.int32, device="cuda"))
    assert torch.equal(part["pi"], full["pi"][400:800]) and torch.equal(part["logp_pi"], full["logp_pi"][400:800])


def test_iv_gae_rollout_with_device_actor_and_critics_in_one_graph():
    """configs[3] end to end with REAL networks: the captured graph holds the actor (tanh 256 x 256 + Gaussian head), the V /
    VC ensembles (7 members, one per dynamics member: 3-D predict), the 3-D dynamics step and the 'iv_gae' bookkeeping;
    stored values equal BNN.predict of the stored states, and finish() equals the oracle's iv-GAE on the same pools."""
    import torch
    import constrainedmbpo_b200 as cm
    env, E, B, T = "humanoidsafe", 7, 600, 6
    p = make_params(env, HIDDEN, parity_scalers=False, bias=False)
    fe = build_cuda_env(p, env, HIDDEN, cares_about_cost=True, seed=2)
    W, b, log_std = _actor_params(45, 17, (256, 256), 4)
    actor = cm.GaussianActor(45, 17, seed=5)
    actor.load_state(W, b, log_std)
    critics = []
    for k, name in enumerate(("V", "VC")):
        pc = orc.init_ensemble(np.random.default_rng(50 + k), 45, 1, (128, 128), E, probabilistic=False,
                               use_scaler_in=False, use_scaler_out=False, dtype=np.float32)
        m = cm.construct_model(in_dim=45, out_dim=1, name=name, hidden_dims=(128, 128), num_networks=E, num_elites=E,
                               loss="MSE", use_scaler_in=False, use_scaler_out=False)
        m.load_state(pc.weights, pc.biases)
        critics.append(m)
    vf = cm.CriticValues(critics[0], critics[1], B, seed=3)
    perms = [np.random.default_rng(100 + t).permutation(E) for t in range(T + 1)]
    ro = cm.ModelRollout(fe, T, rollout_mode="iv_gae", value_fn=vf, perm_fn=lambda t: perms[t])
    obs, _ = make_inputs(env, B)
    obs[:, 0] = np.random.default_rng(1).uniform(1.2, 1.8, B)
    start = torch.from_numpy(obs).cuda()
    ro.capture(start, actor)
    ro.replay()
    torch.cuda.synchronize()
    got = ro.get(numpy=True)
    pop = got["populated"]
    assert pop.sum() > B and got["obs"].shape == (E, B, T, 45)
    # stored values = member e's critic on member e's stored state
    for key, m in (("val", critics[0]), ("cval", critics[1])):
        x = got["obs"][:, pop]                                                   # [E, n, 45]
        pred = m.predict(x, factored=True, inc_var=False)[..., 0]               # 3-D input: [E, n]
        assert np.allclose(got[key][:, pop], pred, rtol=1e-5, atol=1e-6), key
    # stored actions are the actor's: one noise draw per row shared by the members => (act - mu) / std equal across e
    mu = actor.get_action_outs(got["obs"][:, :, 0], noise=np.zeros((B, 17), np.float32))["pi"]      # [E, B, 17], first column
    z = (got["act"][:, :, 0] - mu) / np.exp(log_std)
    assert np.allclose(z, z[0:1], atol=3e-4) and 0.8 < z[0].std() < 1.2
    res = ro.finish(gamma=0.99, lam=0.95, cost_gamma=0.97, cost_lam=0.9, model_ind=2, max_uncertainty_r=5.5, max_uncertainty_c=5.5)
    ref = orc.finish_paths(got["rew"], got["val"], got["cost"], got["cval"], got["finish_ptr"], got["last_val"],
                           got["last_cval"], "iv_gae", 0.99, 0.95, 0.97, 0.9, model_ind=2, cares_about_cost=True,
                           max_uncertainty_r=5.5, max_uncertainty_c=5.5, populated=pop)
    for k in ("adv", "ret", "cadv", "cret", "ret_var"):
        assert np.allclose(res[k].cpu().numpy(), ref[k], rtol=5e-5, atol=5e-6), k
    # a second replay draws fresh actions (device Philox base) and fresh elites... the start states stay
    a0 = ro.pool["act"].clone()
    ro.replay(); torch.cuda.synchronize()
    assert not torch.equal(a0, ro.pool["act"]) and torch.equal(ro.pool["obs"][:, :, 0], torch.from_numpy(got["obs"][:, :, 0]).cuda())
