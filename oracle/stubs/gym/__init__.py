"""TEST INFRASTRUCTURE ONLY -- stand-in for gym==0.25.0 (reference requirements.txt:1).
Containers only; no arithmetic happens here (battle_env.py:134,155,160)."""
from . import spaces, utils  # noqa: F401
