"""Import shim for ``import instinct.team as instinct`` (main.py:3): the CUDA-backed team."""
from battlespace.instinct_team import Team  # noqa: F401
