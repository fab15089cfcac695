"""Minimal stand-in for the ``true_environment`` argument of FakeEnv (fake_env.py:20-32): only
``observation_space.shape``, ``action_space.shape``, ``domain`` and ``stacks`` are read."""
from __future__ import annotations


class _Box:
    def __init__(self, dim):
        self.shape = (int(dim),)


class SimpleEnv:
    def __init__(self, obs_dim, act_dim, domain="", stacks=1):
        self.observation_space = _Box(obs_dim)
        self.action_space = _Box(act_dim)
        self.domain = domain
        self.stacks = stacks
