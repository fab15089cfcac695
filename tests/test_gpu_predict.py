"""GPU parity of ``BNN.predict`` (mbpo/models/bnn.py:671-776) through the C ABI vs the fp64 oracle."""
import numpy as np
import pytest

from tests.helpers import (ENV_DIMS, VAR_RTOL_BF16, assert_means_close, make_inputs, make_params, orc, out_scale)

pytestmark = pytest.mark.gpu


def _cuda_model(p, hidden, loss, activation="swish", use_scalers=True, seed=0, operand_dtype="bf16"):
    import constrainedmbpo_b200 as cm
    m = cm.construct_model(in_dim=p.in_dim, out_dim=p.out_dim, name="T", hidden_dims=tuple(hidden),
                           num_networks=p.num_nets, num_elites=len(p.elites), loss=loss, activation=activation,
                           use_scaler_in=use_scalers, use_scaler_out=use_scalers, seed=seed, operand_dtype=operand_dtype)
    m.load_state([w.astype(np.float32) for w in p.weights], [b.astype(np.float32) for b in p.biases],
                 scaler_in=(p.mu_in, p.var_in) if use_scalers else None,
                 scaler_out=(p.mu_out, p.var_out) if use_scalers else None, elites=p.elites)
    return m


def _generic(in_dim, out_dim, hidden, E, prob, activation="swish", seed=0, scalers=True):
    rng = np.random.default_rng(seed)
    p = orc.init_ensemble(rng, in_dim, out_dim, hidden, E, probabilistic=prob, activation=activation,
                          use_scaler_in=scalers, use_scaler_out=scalers, elites=list(range(max(1, E - 1))),
                          dtype=np.float64)
    for b in p.biases:
        b += rng.normal(0, 0.1, b.shape)
    if scalers:
        p.mu_in = rng.uniform(-1, 1, p.mu_in.shape)
        p.var_in = rng.uniform(0.25, 4, p.var_in.shape)
        p.mu_out = rng.uniform(-1, 1, p.mu_out.shape)
        p.var_out = rng.uniform(0.25, 4, p.var_out.shape)
    return p


@pytest.mark.parametrize("env", ["antsafe", "halfcheetah", "humanoid", "hopper"])
def test_predict_factored_dynamics(env):
    """2-D factored probabilistic predict (bnn.py:711-717) at the BASELINE ensemble shape."""
    hidden = (200, 200, 200, 200)
    p = make_params(env, hidden)
    obs, act = make_inputs(env, 2000)
    x = np.concatenate([obs, act], -1)
    m = _cuda_model(p, hidden, "MSPE")
    mean, var = m.predict(x, factored=True)
    rm, rv = orc.predict(p, x.astype(np.float64), factored=True)
    assert mean.shape == rm.shape and var.shape == rv.shape and mean.dtype == np.float32
    assert_means_close(mean, rm, out_scale(p))
    rel = np.abs(var - rv) / rv
    assert rel.max() < VAR_RTOL_BF16, f"variance rel err max {rel.max():.3e}"
    assert np.median(rel) < 1e-2


@pytest.mark.parametrize("B", [1, 7, 31, 127, 128, 129, 255, 256, 257, 1023])
def test_predict_ragged_batches_are_row_independent(B):
    """Ragged tile tails: every row's result must be bit-identical to the same row inside a large
    batch (tiles of 256 rows, tails masked), and match the oracle."""
    env, hidden = "antsafe", (200, 200, 200, 200)
    p = make_params(env, hidden)
    obs, act = make_inputs(env, 1300)
    x = np.concatenate([obs, act], -1)
    m = _cuda_model(p, hidden, "MSPE")
    big_m, big_v = m.predict(x, factored=True)
    mean, var = m.predict(x[:B], factored=True)
    assert np.array_equal(mean, big_m[:, :B]) and np.array_equal(var, big_v[:, :B])
    rm, _ = orc.predict(p, x[:B].astype(np.float64), factored=True)
    assert_means_close(mean, rm, out_scale(p))


def test_predict_empty_batch():
    env, hidden = "halfcheetah", (200, 200, 200, 200)
    p = make_params(env, hidden)
    m = _cuda_model(p, hidden, "MSPE")
    mean, var = m.predict(np.zeros((0, p.in_dim), np.float32), factored=True)
    assert mean.shape == (7, 0, p.out_dim) and var.shape == (7, 0, p.out_dim)


def test_predict_non_factored_probabilistic():
    """bnn.py:350-352, :731-735: ensemble mean, mean_E(var) + mean_E((mean-mean_E)^2)."""
    env, hidden = "halfcheetah", (200, 200, 200, 200)
    p = make_params(env, hidden)
    obs, act = make_inputs(env, 777)
    x = np.concatenate([obs, act], -1)
    m = _cuda_model(p, hidden, "MSPE")
    mean, var = m.predict(x, factored=False)
    rm, rv = orc.predict(p, x.astype(np.float64), factored=False)
    assert mean.shape == rm.shape == (777, p.out_dim)
    assert_means_close(mean, rm, out_scale(p))
    assert (np.abs(var - rv) / rv).max() < VAR_RTOL_BF16
    only_mean = m.predict(x, factored=False, inc_var=False)
    assert np.array_equal(only_mean, mean)


@pytest.mark.parametrize("factored", [True, False])
def test_predict_critic_shape_non_probabilistic(factored):
    """The V / VC critics (cpo_policy.py:487-506): 3 nets, 128x128, MSE => no variance head;
    inc_var returns var_E(means) (repeated over E when factored, bnn.py:718-724, :736-743)."""
    p = _generic(29, 1, (128, 128), 3, prob=False, seed=2, scalers=False)
    m = _cuda_model(p, (128, 128), "MSE", use_scalers=False)
    x = np.random.default_rng(3).standard_normal((500, 29)).astype(np.float32)
    mean, var = m.predict(x, factored=factored, inc_var=True)
    rm, rv = orc.predict(p, x.astype(np.float64), factored=factored, inc_var=True)
    assert mean.shape == rm.shape and var.shape == rv.shape
    assert_means_close(mean, rm, np.ones(1))
    assert np.allclose(var, rv, rtol=0.25, atol=1e-4)
    m2 = m.predict(x, factored=factored, inc_var=False)
    assert np.array_equal(m2, mean)


def test_predict_costnn_shape():
    """CostNN (fake_env.py:81-94): in = 2*obs+act, 64x64, MSE, factored=False, inc_var=False."""
    p = _generic(2 * 17 + 6, 1, (64, 64), 7, prob=False, seed=4, scalers=False)
    m = _cuda_model(p, (64, 64), "MSE", use_scalers=False)
    x = np.random.default_rng(5).standard_normal((1000, 40)).astype(np.float32)
    got = m.predict(x, factored=False, inc_var=False)
    ref = orc.predict(p, x.astype(np.float64), factored=False, inc_var=False)
    assert got.shape == ref.shape == (1000, 1)
    assert_means_close(got, ref, np.ones(1))


def test_predict_3d_inputs():
    """3-D inputs: member e consumes inputs[e] (fc.py:107-108, bnn.py:749-754)."""
    env, hidden = "humanoid", (200, 200, 200, 200)
    p = make_params(env, hidden)
    obs, act = make_inputs(env, 300, E=7)
    x = np.concatenate([obs, act], -1)
    m = _cuda_model(p, hidden, "MSPE")
    mean, var = m.predict(x, factored=True)
    rm, rv = orc.predict(p, x.astype(np.float64), factored=True)
    assert mean.shape == (7, 300, p.out_dim)
    assert_means_close(mean, rm, out_scale(p))
    assert (np.abs(var - rv) / rv).max() < VAR_RTOL_BF16


@pytest.mark.parametrize("activation", ["ReLU", "tanh", "sigmoid", "swish"])
def test_predict_activations(activation):
    """FC._activations (fc.py:14-21) on a small 2-layer net."""
    p = _generic(20, 5, (96, 64), 4, prob=True, activation=activation, seed=6)
    m = _cuda_model(p, (96, 64), "NLL", activation=activation)
    x = np.random.default_rng(7).standard_normal((333, 20)).astype(np.float32)
    mean, var = m.predict(x, factored=True)
    rm, rv = orc.predict(p, x.astype(np.float64), factored=True)
    assert_means_close(mean, rm, out_scale(p))
    assert (np.abs(var - rv) / rv).max() < VAR_RTOL_BF16


def test_predict_invalid_dimension_raises():
    p = _generic(20, 5, (64,), 3, prob=True)
    m = _cuda_model(p, (64,), "NLL")
    with pytest.raises(ValueError, match="Invalid input dimension"):  # fc.py:110-111
        m.predict(np.zeros((4, 21), np.float32), factored=True)
    with pytest.raises(ValueError):
        m.predict(np.zeros((2, 4, 20), np.float32), factored=True)  # leading dim != E


def test_zero_weights_known_answer():
    """SURVEY §4 KAT: zero weights and biases => mean = mu_out exactly, var = s_out^2 (utils.py:172,192)."""
    env, hidden = "antsafe", (200, 200, 200, 200)
    p = make_params(env, hidden)
    for w in p.weights:
        w[...] = 0
    for b in p.biases:
        b[...] = 0
    m = _cuda_model(p, hidden, "MSPE")
    obs, act = make_inputs(env, 300)
    mean, var = m.predict(np.concatenate([obs, act], -1), factored=True)
    s = out_scale(p).astype(np.float32)
    assert np.array_equal(mean, np.broadcast_to(p.mu_out.astype(np.float32), mean.shape))
    assert np.allclose(var, np.broadcast_to(s * s, var.shape), rtol=2e-6)


def test_scaler_floor_quirk_q6():
    """var_in -> 0 hits the 1e-2 floor (utils.py:161): dims 5 and obs+1 of make_params have var 1e-6,
    so those inputs are multiplied by 100 -- a model that ignored the floor would be off by 10x."""
    env, hidden = "antsafe", (200, 200, 200, 200)
    p = make_params(env, hidden)
    assert p.var_in[0, 5] == 1e-6
    m = _cuda_model(p, hidden, "MSPE")
    obs, act = make_inputs(env, 256)
    x = np.concatenate([obs, act], -1)
    x[:, 5] = p.mu_in[0, 5] + 0.004 * np.random.default_rng(1).standard_normal(256)
    mean, _ = m.predict(x, factored=True)
    rm, _ = orc.predict(p, x.astype(np.float64), factored=True)
    assert_means_close(mean, rm, out_scale(p))


def test_set_state_after_training_refreshes_weights():
    """After BNN.train() the shim must re-read weights / scalers / elites."""
    env, hidden = "halfcheetah", (200, 200, 200, 200)
    p1, p2 = make_params(env, hidden, seed=0), make_params(env, hidden, seed=9)
    m = _cuda_model(p1, hidden, "MSPE")
    obs, act = make_inputs(env, 200)
    x = np.concatenate([obs, act], -1)
    a, _ = m.predict(x, factored=True)
    m.load_state([w.astype(np.float32) for w in p2.weights], [b.astype(np.float32) for b in p2.biases],
                 scaler_in=(p2.mu_in, p2.var_in), scaler_out=(p2.mu_out, p2.var_out), elites=[6, 6, 2])
    b, _ = m.predict(x, factored=True)
    rb, _ = orc.predict(p2, x.astype(np.float64), factored=True)
    assert not np.array_equal(a, b)
    assert_means_close(b, rb, out_scale(p2))
    assert list(m.elite_inds) == [6, 6, 2]


def test_unsupported_width_fails_loudly():
    import constrainedmbpo_b200 as cm
    with pytest.raises(cm.CmbpoError):
        cm.construct_model(in_dim=10, out_dim=3, hidden_dims=(4096,), num_networks=2, num_elites=1)
