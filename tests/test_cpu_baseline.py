"""The timed CPU port (oracle/cpu_baseline.py: torch BLAS GEMMs + NumPy post-processing, fp32) must agree
with the fp64 NumPy oracle -- it is what bench.py reports as cpu_baseline / --impl reference."""
import json
import os
import subprocess
import sys

import numpy as np

from oracle import cpu_baseline as cb
from tests.helpers import make_inputs, make_params, orc, out_scale

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_port_step_matches_oracle():
    env = "antsafe"
    p = make_params(env)
    obs, act = make_inputs(env, 1500)
    idx = np.random.default_rng(7).choice(p.elites, size=1500)
    ens = cb.TorchEnsemble(p, threads=2)
    no, rw, tm, info = cb.fake_env_step(ens, env, obs, act, idx, cares_about_cost=True)
    rno, rrw, rtm, rinfo = orc.fake_env_step(p, env, obs.astype(np.float64), act.astype(np.float64), elite_idx=idx,
                                             cares_about_cost=True)
    s = out_scale(p)
    assert (np.abs(no - rno) <= 1e-5 * np.abs(rno) + 1e-5 * s[:-1]).all()
    assert (tm != rtm).sum() <= 1 and (info["cost"] != rinfo["cost"]).sum() <= 1
    rel = np.abs(info["ensemble_dkl_path"] - rinfo["ensemble_dkl_path"]) / (rinfo["ensemble_dkl_path"] + 1e-9)
    assert rel.max() < 1e-3
    m3, v3 = ens.predict_factored(np.concatenate(make_inputs(env, 40, E=7), -1))
    r3, rv3 = orc.predict(p, np.concatenate(make_inputs(env, 40, E=7), -1).astype(np.float64), factored=True)
    assert np.allclose(m3, r3, rtol=1e-4, atol=1e-5) and np.allclose(v3, rv3, rtol=1e-3)


def test_reference_arm_prints_contract_line():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1",
                          "--warmup", "0", "--cpu-rows", "512"], check=True, capture_output=True, text=True, cwd=ROOT)
    line = json.loads(out.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["unit"] == "transitions/s" and line["value"] > 0
    assert line["cpu_baseline"]["kind"] == "port" and line["cpu_baseline"]["cores"] >= 1
    # the default workload is BASELINE configs[2] (1 M HalfCheetah start states, strong scaling); both arms print the
    # same `config` for it
    import bench
    assert line["e2e"]["h2d_bytes_per_step"] == 0 and line["config"]["workload"] == "halfcheetah_schedule_L10_1M"
    assert line["config"] == bench.config_of("halfcheetah_schedule_L10_1M", bench.WORKLOADS["halfcheetah_schedule_L10_1M"], 1)
    assert line["scaling"] == "strong" and line["config"]["start_states_total"] == 1_000_000


def test_reference_arm_predict_workload_and_other_ranks_stay_silent():
    """configs[0] (one BNN.predict on 10 k rows) on the CPU port; under torchrun only rank 0 runs the reference arm."""
    cmd = [sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0",
           "--workload", "antsafe_predict_10k", "--cpu-rows", "2000"]
    out = subprocess.run(cmd, check=True, capture_output=True, text=True, cwd=ROOT)
    line = json.loads(out.stdout.strip().splitlines()[-1])
    assert line["config"]["kind"] == "predict" and line["value"] > 0
    out = subprocess.run(cmd, check=True, capture_output=True, text=True, cwd=ROOT, env=dict(os.environ, RANK="1", WORLD_SIZE="2"))
    assert out.stdout.strip() == ""
