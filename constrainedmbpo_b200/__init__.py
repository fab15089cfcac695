"""constrainedmbpo_b200 -- B200-native model-rollout path of anyboby/ConstrainedMBPO.

Public surface mirrors the reference's own for this path:
``construct_model`` (mbpo/models/constructor.py:9), ``BNN.predict`` (mbpo/models/bnn.py:671),
``FakeEnv.step`` (mbpo/models/fake_env.py:131), plus the device-resident rollout driver
(``ModelRollout``: softlearning/samplers/model_sampler.py:220-474 cutoff logic and
softlearning/replay_pools/modelbuffer.py:45-173 pool layout).  All compute runs in
``libcmbpo_b200.so`` (hand-written sm_100a CUDA behind a C ABI, include/cmbpo_b200.h); there is
no CPU fallback.
"""
from ._lib import CmbpoError, LIB_PATH, load as load_library  # noqa: F401
from .bnn import BNN, FC, TensorStandardScaler  # noqa: F401
from .constructor import construct_model  # noqa: F401
from .fake_env import FakeEnv  # noqa: F401
from .static import STATIC, StaticFns  # noqa: F401
from .spaces import SimpleEnv  # noqa: F401
from .rollout import CriticValues, DklAccumulators, ModelRollout, StubPolicy, compute_dynamics_dkl  # noqa: F401
from .gae import finish_paths, normalize_advantages  # noqa: F401
from .policy import GaussianActor  # noqa: F401

__all__ = ["construct_model", "BNN", "FC", "FakeEnv", "ModelRollout", "STATIC", "StaticFns",
           "SimpleEnv", "StubPolicy", "GaussianActor", "compute_dynamics_dkl", "DklAccumulators", "TensorStandardScaler", "CmbpoError", "load_library"]
