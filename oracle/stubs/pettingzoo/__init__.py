"""TEST INFRASTRUCTURE ONLY -- stand-in for PettingZoo==1.19.0 (reference requirements.txt:5).
Only the ParallelEnv base class is used on the hot path (battle_env.py:61)."""
from . import utils  # noqa: F401


class ParallelEnv:
    pass
