"""Register ``engine_type: b200`` in an unmodified V-Gate checkout at import time.

The maintainer's version of this is two edits: ``"b200"`` in the ``allowed``
tuple of ``vgate/config.py:118`` and a ``b200`` branch in ``vgate/engine.py:52-58``.  This module does
the same from the outside: it widens the ``ModelConfig.engine_type`` validator and wraps
``vgate.engine._create_backend``; precedence stays remote endpoints > ``VGATE_DRY_RUN`` > engine_type
(``vgate/engine.py:47-58``).
"""
import types

import vgate.config as _cfg
import vgate.engine as _eng

_ALLOWED = ("vllm", "sglang", "b200")


def _validate_engine_type(cls, v: str) -> str:
    if v not in _ALLOWED:
        raise ValueError(f"engine_type must be one of {_ALLOWED}, got '{v}'")
    return v


def _patch_config():
    dec = _cfg.ModelConfig.__pydantic_decorators__.field_validators["validate_engine_type"]
    dec.func = types.MethodType(_validate_engine_type, _cfg.ModelConfig)      # bound: pydantic sees (value)
    _cfg.ModelConfig.validate_engine_type = classmethod(_validate_engine_type)
    _cfg.ModelConfig.model_rebuild(force=True)
    for owner in (_cfg.VGateConfig,):            # models that embed ModelConfig must pick up the new schema
        owner.model_rebuild(force=True)


_orig_create = _eng._create_backend


def _create_backend(engine_type, worker_config=None):
    import os
    if worker_config is not None and _eng._is_remote(worker_config) and os.environ.get("VGATE_B200_REMOTE") == "coalesce":
        # gateway side: the coalescing / streaming worker hop instead of the reference's RemoteBackend
        from vgate_b200.remote import CoalescingRemoteBackend
        return CoalescingRemoteBackend(worker_config)
    if engine_type == "b200" and not (worker_config is not None and _eng._is_remote(worker_config)) and not _eng.DRY_RUN:
        from vgate_b200 import B200Backend
        return B200Backend()
    return _orig_create(engine_type, worker_config)


if not getattr(_eng, "_b200_registered", False):
    _patch_config()
    _eng._create_backend = _create_backend
    _eng._b200_registered = True
