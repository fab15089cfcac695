"""Host-side mirror of the reference surface (no GPU needed): construction protocol errors, FC repr /
.nns line format, scaler running moments, static_fns registry, bench FLOP accounting."""
import numpy as np
import pytest

import constrainedmbpo_b200 as cm
from constrainedmbpo_b200.bnn import BNN, FC, TensorStandardScaler


def test_bnn_structure_protocol_errors_match_reference():
    m = BNN({"name": "x", "num_networks": 3, "num_elites": 2, "loss": "MSPE"})
    assert m.is_probabilistic and not m.is_tf_model and m.num_nets == 3
    with pytest.raises(ValueError, match="input dimension"):   # bnn.py:135-136
        m.add(FC(8))
    with pytest.raises(RuntimeError, match="empty"):           # bnn.py:155-156
        m.pop()
    with pytest.raises(RuntimeError, match="empty"):           # bnn.py:170-171
        m.finalize()
    m.add(FC(8, input_dim=4, activation="swish"))
    m.add(FC(2))
    assert m.layers[1].get_input_dim() == 8 and m.layers[1].get_ensemble_size() == 3
    with pytest.raises(ValueError, match="name"):
        BNN({"num_elites": 1})
    assert len(m.elite_inds) == 2 and all(0 <= i < 3 for i in m.elite_inds)   # bnn.py:98-102
    assert set(m.random_inds(100)) <= set(int(i) for i in m.elite_inds)       # bnn.py:437-439
    assert not BNN({"name": "c", "num_elites": 1, "loss": "MSE"}).is_probabilistic


def test_fc_repr_is_the_nns_line_format():
    fc = FC(200, input_dim=37, activation="swish", weight_decay=2.5e-07, ensemble_size=7)
    assert repr(fc) == "FC(output_dim=200, input_dim=37, activation='swish', weight_decay=2.5e-07, ensemble_size=7)"
    with pytest.raises(ValueError):
        FC(3, activation="gelu")


def test_scaler_running_moments_match_batch_statistics():
    """utils.py:120-139, :228-239: chunked fit == moments of the concatenated data (population variance)."""
    rng = np.random.default_rng(0)
    data = rng.standard_normal((300, 5)) * 3 + 1
    s = TensorStandardScaler(5)
    for chunk in np.split(data, [50, 120, 121]):
        s.fit(chunk)
    assert np.allclose(s.cached_mu, data.mean(0, keepdims=True))
    assert np.allclose(s.cached_var, data.var(0, keepdims=True))
    assert s.cached_count == 300 and s.fitted


def test_static_registry():
    assert cm.STATIC["antsafe"].has_cost_f and not cm.STATIC["ant"].has_cost_f
    from constrainedmbpo_b200.static import resolve

    class StaticFns:  # the reference's classes are all called StaticFns and identified by module
        __module__ = "mbpo.static.hoppersafe"
    assert resolve(StaticFns).name == "hoppersafe"
    assert resolve("HalfCheetah").env_id == 3
    with pytest.raises(KeyError):
        resolve("safexp-pointgoal2")


def test_rollout_mode_validation():
    class E:  # never touched before validation
        device = "cpu"
    with pytest.raises(ValueError, match="iv_gae"):
        cm.ModelRollout(E(), 10, rollout_mode="bogus")
    assert cm.ModelRollout(E(), 10, rollout_mode="iv_gae").mode == "iv_gae"   # cmbpo.py:89 third mode
    with pytest.raises(ValueError, match="dkl_lim"):
        cm.ModelRollout(E(), 10, rollout_mode="uncertainty")


def test_bench_flop_accounting_matches_survey():
    import bench
    assert bench.flops_per_row(bench.WORKLOADS["antsafe_uncertainty_H30_100k"]) == 1_951_600
    assert bench.flops_per_row(bench.WORKLOADS["halfcheetah_schedule_L10_1M"]) == 1_845_200
    assert bench.bytes_per_transition(bench.WORKLOADS["antsafe_uncertainty_H30_100k"]) == 4 * (29 + 8 + 29 + 3) + 1
