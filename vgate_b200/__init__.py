"""b200gate: a B200-native (sm_100a) worker backend for V-Gate's generate path.

Public surface (mirrors ``vgate/backends/vllm_backend.py:35-149``):
    from vgate_b200 import B200Backend
"""
__all__ = ["B200Backend"]


def __getattr__(name):
    if name == "B200Backend":
        from .backend import B200Backend
        return B200Backend
    raise AttributeError(name)
