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


def test_bench_sharding_and_workload_table():
    """Strong scaling cuts the SAME global batch into contiguous blocks; weak scaling gives every rank its own rows;
    every BASELINE config has a workload."""
    import bench
    w = bench.WORKLOADS[bench.DEFAULT_WORKLOAD]
    assert bench.DEFAULT_WORKLOAD == "halfcheetah_schedule_L10_1M" and w["scaling"] == "strong"
    for world in (1, 2, 3, 8):
        cuts = [bench.shard_rows(w, world, r) for r in range(world)]
        assert cuts[0][0] == 0 and sum(n for _, n in cuts) == w["rows"]
        assert all(cuts[i][0] + cuts[i][1] == cuts[i + 1][0] for i in range(world - 1))
    small = dict(w, rows=10)
    full = bench.synth_start_states(small, 10, 1)
    lo, n = bench.shard_rows(small, 3, 1)
    assert np.array_equal(bench.synth_start_states(small, n, 1, lo=lo), full[lo:lo + n])
    wk = bench.WORKLOADS["antsafe_uncertainty_H30_100k"]
    assert bench.shard_rows(wk, 4, 2) == (200_000, 100_000)
    for name in ("antsafe_predict_10k", "antsafe_uncertainty_H30_100k", "halfcheetah_schedule_L10_1M",
                 "humanoidsafe_iv_gae_H12_1M", "antsafe_step_4M", "antsafe_step_256k_4x1024"):
        assert name in bench.WORKLOADS
    W, B = bench.draw_weights(dict(w, E=2))
    assert [x.shape for x in W] == [(2, 23, 200), (2, 200, 200), (2, 200, 200), (2, 200, 200), (2, 200, 36)]
    assert all(np.abs(x).max() <= 2 * (1 / (2 * np.sqrt(x.shape[1]))) + 1e-7 for x in W) and not B[0].any()


def test_nns_structure_reader_round_trip(tmp_path):
    """``BNN._load_structure`` (bnn.py:829-846) on a file in the exact line format ``BNN.save`` writes (:811-819):
    the literal lines below are what the reference's ``FC.__repr__`` (fc.py:47-51) prints for the antsafe dynamics
    ensemble of ``FakeEnv`` (fake_env.py:61-75, decay 1e-6)."""
    from constrainedmbpo_b200.bnn import parse_nns
    lines = [
        "FC(output_dim=200, input_dim=37, activation='swish', weight_decay=2.5e-07, ensemble_size=7)",
        "FC(output_dim=200, input_dim=200, activation='swish', weight_decay=5e-07, ensemble_size=7)",
        "FC(output_dim=30, input_dim=200, activation=None, weight_decay=1e-06, ensemble_size=7)",
    ]
    path = tmp_path / "BNN_12.nns"
    path.write_text("\n".join(lines) + "\n")
    layers = parse_nns(str(path))
    assert [repr(l) for l in layers] == lines
    assert layers[0].get_activation() == "swish" and layers[-1].get_activation() is None
    assert layers[1].weight_decay == 5e-07 and layers[0].get_ensemble_size() == 7
    # load_model=True takes structure and ensemble size from the file (bnn.py:90-97); no GPU until finalize()
    m = BNN({"name": "BNN_12", "model_dir": str(tmp_path), "load_model": True, "num_elites": 5, "loss": "MSPE",
             "num_networks": 3})
    assert m.model_loaded and m.num_nets == 7 and len(m.layers) == 3 and len(m.elite_inds) == 5
    with pytest.raises(RuntimeError, match="loaded"):       # bnn.py:139-140
        m.add(FC(4))
    with pytest.raises(ValueError, match="model directory"):
        BNN({"name": "BNN_12", "load_model": True, "num_elites": 5})
    (tmp_path / "bad.nns").write_text("Dense(3)\n")
    with pytest.raises(ValueError, match="structure line"):
        parse_nns(str(tmp_path / "bad.nns"))
