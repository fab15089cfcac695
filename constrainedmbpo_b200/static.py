"""Registry of the static termination / cost functions the fused step kernel implements
(mbpo/static/__init__.py:7-24; device code: csrc/step_post.cu ``termination`` / ``static_cost``)."""
from __future__ import annotations

from . import _lib


class StaticFns:
    """Handle for one ``mbpo/static/<env>.py`` module: the kernel evaluates ``termination_fn``
    (and ``cost_f`` where the module defines one) per row on the device."""

    def __init__(self, name):
        if name not in _lib.ENV_IDS:
            raise KeyError(f"no fused static_fns for {name!r}; known: {sorted(k for k in _lib.ENV_IDS if k)}")
        self.name = name
        self.env_id = _lib.ENV_IDS[name]
        self.has_cost_f = name in _lib.ENVS_WITH_STATIC_COST

    def __repr__(self):
        return f"StaticFns({self.name!r})"


def resolve(static_fns):
    """Accepts an env name, a ``StaticFns`` of this package, or the reference's own
    ``StaticFns`` class (identified by its defining module ``mbpo.static.<env>``)."""
    if isinstance(static_fns, StaticFns):
        return static_fns
    if isinstance(static_fns, str):
        return StaticFns(static_fns.lower())
    mod = getattr(static_fns, "__module__", "") or ""
    name = mod.split(".")[-1].lower()
    if name in _lib.ENV_IDS:
        return StaticFns(name)
    raise KeyError(f"cannot map {static_fns!r} to a fused static_fns module")


STATIC = {name: StaticFns(name) for name in _lib.ENV_IDS if isinstance(name, str) and name != "none"}
