"""Import shim: put ``deep-rl-battlespace_b200`` ahead of the reference on ``sys.path`` and the
reference's drivers (``import envs.battle_env as battle_env``, main.py:1, evaluate.py:1,
test_env.py:1) pick up the CUDA env unchanged."""
from battlespace.env import parallel_env, BattleEnv  # noqa: F401
