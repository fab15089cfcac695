"""Generates tests/golden/*.npz by running the reference's OWN code where it can run here.

Run in the build container only (needs /root/reference); the fixtures are committed because
/root/reference does not exist on the GPU box.

  static_fns.npz  -- mbpo/static/<env>.py termination_fn / cost_f imported from file
                     (pure NumPy) on seeded fp32 inputs that straddle every threshold and
                     contain NaN/Inf rows.
  average_dkl.npz -- mbpo/models/utils.py:7-57 (EPS, gaussian_kl_np, average_dkl) executed from
                     the reference's own source lines (the module's top-level
                     ``import tensorflow`` cannot run here, so the pure-NumPy functions are
                     exec'd by line range) on seeded [E,B,D] means / stds incl. zero stds.
"""
import importlib.util
import os
import sys

import numpy as np

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from tests.helpers import ENV_DIMS, make_inputs  # noqa: E402

ENVS = ["ant", "antsafe", "halfcheetah", "halfcheetahsafe", "humanoid", "humanoidsafe", "hopper",
        "hoppersafe", "walker2d"]


def load_static(name):
    spec = importlib.util.spec_from_file_location(f"ref_static_{name}", f"{REF}/mbpo/static/{name}.py")
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod.StaticFns


def static_inputs(env, B=512, seed=11):
    """next_obs rows around every threshold the static fns test, plus non-finite rows."""
    obs, act = make_inputs(env, B, seed=seed)
    nobs, _ = make_inputs(env, B, seed=seed + 100)
    rng = np.random.default_rng(seed + 7)
    nobs[:, 0] = rng.uniform(-0.2, 2.4, B).astype(np.float32)
    if nobs.shape[1] > 4:
        q = rng.standard_normal((B, 4))
        q[: B // 2] /= np.linalg.norm(q[: B // 2], axis=-1, keepdims=True)
        nobs[:, 1:5] = q.astype(np.float32)
    nobs[:, 1] = np.where(rng.random(B) < 0.5, rng.uniform(-0.4, 0.4, B), nobs[:, 1]).astype(np.float32)
    nobs[:, -1] = rng.uniform(-4, 4, B).astype(np.float32)
    # exact threshold values and specials
    specials = [0.2, 1.0, 0.7, 0.8, 2.0, np.nextafter(np.float32(0.2), np.float32(0)), np.float32(1.0000001)]
    for i, v in enumerate(specials):
        nobs[i, 0] = v
    nobs[20, 3] = np.nan
    nobs[21, -1] = np.inf
    nobs[22, 0] = -np.inf
    nobs[23, 2] = 150.0
    nobs[24, 1] = 100.0
    nobs[25, -1] = np.float32(3.2)
    nobs[26, -1] = np.float32(-3.0)
    nobs[27, -1] = np.float32(0.2)
    nobs[28, -1] = np.float32(1.0)
    return obs, act, nobs


def main():
    out = {}
    for env in ENVS:
        fns = load_static(env)
        obs, act, nobs = static_inputs(env)
        with np.errstate(all="ignore"):
            term = fns.termination_fn(obs, act, nobs)
        out[f"{env}/obs"], out[f"{env}/act"], out[f"{env}/next_obs"] = obs, act, nobs
        out[f"{env}/term"] = np.asarray(term)
        if hasattr(fns, "cost_f"):
            with np.errstate(all="ignore"):
                out[f"{env}/cost"] = np.asarray(fns.cost_f(obs, act, nobs, None))
        # 3-D call ([E,B,obs]) as the iv_gae path makes it
        o3, a3, n3 = (np.stack([x, x[::-1]]) for x in (obs, act, nobs))
        try:
            with np.errstate(all="ignore"):
                out[f"{env}/term3d"] = np.asarray(fns.termination_fn(o3, a3, n3))
        except AssertionError:  # walker2d.py:7 asserts 2-D inputs
            pass
    np.savez_compressed(os.path.join(HERE, "static_fns.npz"), **out)

    # average_dkl from the reference's own source lines
    src = open(f"{REF}/mbpo/models/utils.py").read().split("\n")
    code = "\n".join(["import numpy as np"] + src[6:57])  # lines 7..57: EPS .. end of average_dkl
    ns = {}
    exec(compile(code, "ref_utils_7_57", "exec"), ns)
    rng = np.random.default_rng(5)
    E, B, D = 7, 64, 30
    mu = rng.standard_normal((E, B, D)).astype(np.float32)
    std = np.exp(rng.uniform(-6, 2, (E, B, D))).astype(np.float32)
    std[0, 0, :] = 0.0           # log -> -inf -> clipped to -100
    std[:, 1, :] = std[0, 1, :]  # identical stds
    mu[:, 2, :] = mu[0, 2, :]    # identical members -> dkl == 0 path
    std[:, 2, :] = std[0, 2, :]
    mu[3, 3, 0] = np.nan
    with np.errstate(all="ignore"):
        dkl = ns["average_dkl"](mu.copy(), std.copy())
        mu2, std2 = mu[:2, :, :4].astype(np.float64), std[:2, :, :4].astype(np.float64)
        dkl2 = ns["average_dkl"](mu2.copy(), std2.copy())
    np.savez_compressed(os.path.join(HERE, "average_dkl.npz"), mu=mu, std=std, dkl=dkl, mu2=mu2, std2=std2, dkl2=dkl2)
    print("wrote", os.path.join(HERE, "static_fns.npz"), os.path.join(HERE, "average_dkl.npz"))


if __name__ == "__main__":
    main()
