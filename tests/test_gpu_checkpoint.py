"""Checkpoint ingest: the ``{name}_{epoch}.nns`` + ``.mat`` pair the reference's ``BNN.save`` writes
(mbpo/models/bnn.py:799-827) read back the way its constructor + ``finalize`` do (``load_model=True``, bnn.py:90-97,
:374-379, structure reader :829-846), on the CUDA path."""
import os

import numpy as np
import pytest

from tests.helpers import assert_means_close, make_inputs, make_params, orc, out_scale
from tests.test_gpu_predict import _cuda_model, _generic

pytestmark = pytest.mark.gpu
HIDDEN = (200, 200, 200, 200)


def _write_reference_checkpoint(p, hidden, model_dir, name, loss, decay=1e-6, bounds=None):
    """Files exactly as the reference lays them out: one ``repr(FC)`` line per layer with the head at its
    pre-finalize width (bnn.py:811-819) and ``savemat`` of ``nonoptvars + optvars`` under the keys '0', '1', ...:
    [mu_in, var_in, mu_out, var_out, W0, b0, ..., W_L, b_L, (max_logvar, min_logvar for 'NLL')] (bnn.py:214-227,:821-827)."""
    from scipy.io import savemat
    E = p.num_nets
    dims = [p.in_dim] + list(hidden) + [p.out_dim]
    decays = [decay / 4] + [decay / 2] * (len(hidden) - 1) + [decay]      # constructor.py:57-62
    with open(os.path.join(model_dir, name + ".nns"), "w") as f:
        for i in range(len(dims) - 1):
            act = "'swish'" if i < len(dims) - 2 else "None"
            f.write(f"FC(output_dim={dims[i + 1]}, input_dim={dims[i]}, activation={act}, weight_decay={decays[i]!r}, "
                    f"ensemble_size={E})\n")
    vals = []
    if p.mu_in is not None:
        vals += [p.mu_in.astype(np.float32), p.var_in.astype(np.float32)]
    if p.mu_out is not None:
        vals += [p.mu_out.astype(np.float32), p.var_out.astype(np.float32)]
    for w, b in zip(p.weights, p.biases):
        vals += [w.astype(np.float32), b.astype(np.float32).reshape(E, 1, -1)]
    if loss == "NLL":
        vals += [np.full((1, p.out_dim), bounds[0], np.float32), np.full((1, p.out_dim), bounds[1], np.float32)]
    savemat(os.path.join(model_dir, name + ".mat"), {str(i): v for i, v in enumerate(vals)})


def test_load_model_from_reference_layout(tmp_path):
    import constrainedmbpo_b200 as cm
    env = "antsafe"
    p = make_params(env, HIDDEN)
    _write_reference_checkpoint(p, HIDDEN, str(tmp_path), "BNN_40", "MSPE")
    m = cm.BNN({"name": "BNN_40", "model_dir": str(tmp_path), "load_model": True, "num_elites": 5, "loss": "MSPE",
                "use_scaler_in": True, "use_scaler_out": True})
    assert m.num_nets == 7 and [l.get_output_dim() for l in m.layers] == [200, 200, 200, 200, 30]
    m.finalize()
    assert m.layers[-1].get_output_dim() == 60            # head doubled by finalize (bnn.py:190-191)
    assert np.allclose(m.scaler_in.cached_mu, p.mu_in.astype(np.float32))
    assert np.allclose(m.scaler_out.cached_var, p.var_out.astype(np.float32))
    obs, act = make_inputs(env, 1500)
    x = np.concatenate([obs, act], -1)
    mean, var = m.predict(x, factored=True)
    # identical to the same parameters handed over with load_state, and within tolerance of the oracle
    m2 = _cuda_model(p, HIDDEN, "MSPE")
    mean2, var2 = m2.predict(x, factored=True)
    assert np.array_equal(mean, mean2) and np.array_equal(var, var2)
    rm, _ = orc.predict(p, x.astype(np.float64), factored=True)
    assert_means_close(mean, rm, out_scale(p))


def test_save_then_load_round_trip(tmp_path):
    """``save`` writes files a ``load_model=True`` constructor reads back bit for bit (weights, biases, scalers),
    with the '.nns' head line restored to its pre-finalize form."""
    import constrainedmbpo_b200 as cm
    from constrainedmbpo_b200.bnn import parse_nns
    env = "halfcheetah"
    p = make_params(env, HIDDEN)
    m = _cuda_model(p, HIDDEN, "MSPE")
    m.save(str(tmp_path), 3)
    layers = parse_nns(str(tmp_path / "T_3.nns"))
    assert [l.get_output_dim() for l in layers] == [200, 200, 200, 200, 18] and layers[-1].get_activation() is None
    r = cm.BNN({"name": "T_3", "model_dir": str(tmp_path), "load_model": True, "num_elites": 5, "loss": "MSPE",
                "use_scaler_in": True, "use_scaler_out": True})
    r.finalize()
    for a, b in zip(m.weights + m.biases, r.weights + r.biases):
        assert a.shape == b.shape and bool((a == b).all())
    obs, act = make_inputs(env, 700)
    x = np.concatenate([obs, act], -1)
    a, av = m.predict(x, factored=True)
    b, bv = r.predict(x, factored=True)
    assert np.array_equal(a, b) and np.array_equal(av, bv)
    # a second save of the reloaded model is byte-identical in structure
    r.save(str(tmp_path), 4)
    assert (tmp_path / "T_3.nns").read_text() == (tmp_path / "T_3_4.nns").read_text()


def test_nll_checkpoint_carries_logvar_bounds_and_count_is_checked(tmp_path):
    """'NLL' models append max_logvar / min_logvar to the variable list (bnn.py:208-212,:221-222); a .mat whose
    variable count does not fit the structure is refused."""
    import constrainedmbpo_b200 as cm
    p = _generic(20, 5, (96, 64), 4, prob=True, seed=6)
    _write_reference_checkpoint(p, (96, 64), str(tmp_path), "N_1", "NLL", bounds=(0.5, -6.0))
    m = cm.BNN({"name": "N_1", "model_dir": str(tmp_path), "load_model": True, "num_elites": 3, "loss": "NLL",
                "use_scaler_in": True, "use_scaler_out": True})
    m.finalize()
    assert m.max_logvar.shape == (1, 5) and float(m.max_logvar[0, 0]) == 0.5 and float(m.min_logvar[0, 0]) == -6.0
    x = np.random.default_rng(7).standard_normal((300, 20)).astype(np.float32)
    mean, _ = m.predict(x, factored=True)
    rm, _ = orc.predict(p, x.astype(np.float64), factored=True)
    assert_means_close(mean, rm, out_scale(p))
    m.save(str(tmp_path), 2)
    from scipy.io import loadmat
    d = loadmat(str(tmp_path / "N_1_2.mat"))
    assert len([k for k in d if k.isdigit()]) == 4 + 2 * 3 + 2
    # same files read as an MSPE model: two variables too many
    bad = cm.BNN({"name": "N_1", "model_dir": str(tmp_path), "load_model": True, "num_elites": 3, "loss": "MSPE",
                  "use_scaler_in": True, "use_scaler_out": True})
    with pytest.raises(ValueError, match="variables"):
        bad.finalize()
