"""battlespace -- B200-native batched Deep-RL-Battlespace environment step + instinct policy.

Importing this package loads ``libbattlespace.so`` (CUDA, sm_100a); it raises if the library has
not been built.  There is no CPU fallback anywhere on the product path.
"""
from ._lib import BattlespaceError, LIB_PATH, SYMBOLS  # noqa: F401
from .env import BattleEnv, parallel_env, tie_tick  # noqa: F401
from .replay import DeviceReplay  # noqa: F401
from .sharding import allreduce_stats, broadcast_actors, shard_range, summarize  # noqa: F401
from .instinct_team import Team as InstinctTeam  # noqa: F401
from .actors import DeviceOUNoise, TeamActors  # noqa: F401
from .stats import EpisodeLog  # noqa: F401
