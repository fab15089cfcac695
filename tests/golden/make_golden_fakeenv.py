#!/usr/bin/env python
"""Golden fixtures from the reference's OWN ``FakeEnv.step`` (mbpo/models/fake_env.py:131-222), run here.

TensorFlow 1.14 is not installable in this container, so the TF graph behind ``BNN.predict`` cannot run; but
everything ``FakeEnv.step`` does AROUND ``self._model.predict`` is NumPy.  The module is loaded by file with
stub modules for its imports (tensorflow, the constructor, priors, logging), a ``FakeEnv`` is created without
``__init__`` and given a model whose ``predict`` returns the oracle's (means, variances) for a seeded ensemble
and whose ``random_inds`` draw is recorded.  The real ``mbpo/static/<env>.py`` StaticFns are used.  Recorded:
inputs, the model outputs, the injected draws, and the 4-tuple + info ``step`` returned (2-D and 3-D inputs).

  python tests/golden/make_golden_fakeenv.py     # writes tests/golden/fakeenv_step.npz
"""
import importlib.util
import os
import sys
import types

import numpy as np

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
if not hasattr(np, "int"):
    np.int = int


def _stub(name, **attrs):
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    m.__path__ = []
    sys.modules[name] = m
    return m


def _load(name, rel):
    spec = importlib.util.spec_from_file_location(name, os.path.join(REF, rel))
    m = importlib.util.module_from_spec(spec)
    sys.modules[name] = m
    spec.loader.exec_module(m)
    return m


def load_fake_env():
    _stub("tensorflow")
    for pkg in ("mbpo", "mbpo.models", "mbpo.utils", "mbpo.static", "softlearning", "softlearning.policies",
                "softlearning.policies.safe_utils"):
        _stub(pkg)
    _stub("mbpo.models.constructor", construct_model=None, format_samples_for_dyn=None, format_samples_for_cost=None)
    _stub("mbpo.models.priors", WEIGHTS_PER_DOMAIN={}, PRIORS_BY_DOMAIN={}, PRIOR_DIMS={}, POSTS_BY_DOMAIN={})
    _stub("mbpo.utils.logging", Progress=None, Silent=None)
    _load("softlearning.policies.safe_utils.utils", "softlearning/policies/safe_utils/utils.py")
    _load("mbpo.models.utils", "mbpo/models/utils.py")       # average_dkl: pure NumPy below `import tensorflow`
    return _load("mbpo.models.fake_env", "mbpo/models/fake_env.py").FakeEnv


from tests.helpers import ENV_DIMS, make_inputs, make_params, orc  # noqa: E402


class _Model:
    def __init__(self, p, rng):
        self.p, self.rng = p, rng
        self.last = {}

    def predict(self, inputs, factored=True, inc_var=True):
        out = orc.predict(self.p, inputs, factored=factored, inc_var=inc_var)
        if not inc_var:                       # the cost head: a single array (bnn.py:744-748, :772-776)
            return np.array(out)
        m, v = out
        self.last["means"], self.last["vars"] = np.array(m), np.array(v)
        return np.array(m), np.array(v)

    def random_inds(self, batch_size):
        idx = self.rng.choice(self.p.elites, size=batch_size)
        self.last["idx"] = idx
        return idx


def main():
    FakeEnv = load_fake_env()
    out = {}
    for tag, env, B, ndim, cost in (("antsafe2d", "antsafe", 96, 2, True), ("humanoidsafe3d", "humanoidsafe", 64, 3, True),
                                    ("hopper2d", "hopper", 80, 2, False), ("halfcheetahsafe2d", "halfcheetahsafe", 50, 2, True),
                                    ("ant2d_costnn", "ant", 72, 2, True), ("humanoid3d_costnn", "humanoid", 40, 3, True)):
        obs_dim, act_dim = ENV_DIMS[env]
        p = make_params(env, (24, 24), parity_scalers=True, bias=True, dtype=np.float32)
        static = _load(f"mbpo.static.{env}", f"mbpo/static/{env}.py").StaticFns
        fe = FakeEnv.__new__(FakeEnv)
        fe.prior_f, fe.static_r, fe.domain, fe.env, fe.rew_dim = None, None, "", None, 1
        fe.static_fns, fe.static_done = static, static.termination_fn
        fe.static_c = hasattr(static, "cost_f")
        fe.cares_about_cost = cost
        fe._model = _Model(p, np.random.default_rng(31))
        if cost and not fe.static_c:
            # the learned cost head (fake_env.py:76-94, :191-196): MSE ensemble on concat(obs, act, next_obs), 64x64
            pc = orc.init_ensemble(np.random.default_rng(17), 2 * obs_dim + act_dim, 1, (64, 64), 7, probabilistic=False,
                                   use_scaler_in=False, use_scaler_out=False, elites=list(range(5)), dtype=np.float32)
            fe._cost_model = _Model(pc, np.random.default_rng(32))
            fe.cost_m_loss = "MSE"
        obs, act = make_inputs(env, B, seed=13, E=7 if ndim == 3 else None)
        np.random.seed(77)                       # forward_shuffle draws np.random.permutation(E) (fake_env.py:314)
        next_obs, rew, term, info = fe.step(obs, act)
        np.random.seed(77)
        perm = np.random.permutation(7)
        d = dict(obs=obs, act=act, next_obs=next_obs, rew=rew, term=term, cost=info["cost"],
                 dkl_path=info["ensemble_dkl_path"], dkl_mean=info["ensemble_dkl_mean"], mean_var=info["ensemble_mean_var"],
                 ep_var=info["ensemble_ep_var"], rew_mean=info["rew_mean"], cost_mean=info["cost_mean"],
                 model_means=fe._model.last["means"], model_vars=fe._model.last["vars"],
                 idx=fe._model.last.get("idx", np.zeros(0, np.int64)), perm=perm, cares_about_cost=cost)
        for k, v in d.items():
            out[f"{tag}__{k}"] = np.asarray(v)
        print(tag, "next_obs", next_obs.shape, next_obs.dtype, "terminals", int(np.sum(term)), "cost_mean", float(info["cost_mean"]))
    path = os.path.join(HERE, "fakeenv_step.npz")
    np.savez_compressed(path, **out)
    print("->", path, os.path.getsize(path) // 1000, "kB")


if __name__ == "__main__":
    main()
