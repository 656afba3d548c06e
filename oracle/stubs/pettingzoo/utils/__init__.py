"""TEST INFRASTRUCTURE ONLY -- names imported by battle_env.py:11-12; never called by any driver."""
from . import wrappers  # noqa: F401


def parallel_to_aec(env):
    raise NotImplementedError("AEC wrapping is outside the hot path")
