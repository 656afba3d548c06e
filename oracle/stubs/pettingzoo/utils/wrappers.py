"""TEST INFRASTRUCTURE ONLY -- wrapper names referenced by battle_env.py:23-25 (never called)."""


def _identity(env):
    return env


CaptureStdoutWrapper = _identity
AssertOutOfBoundsWrapper = _identity
OrderEnforcingWrapper = _identity
