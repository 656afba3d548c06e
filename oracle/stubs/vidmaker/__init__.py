"""TEST INFRASTRUCTURE ONLY -- stand-in for vidmaker==2.3.0 (battle_env.py:13,570); rendering is off the hot path."""


class Video:
    def __init__(self, *a, **k):
        pass

    def update(self, *a, **k):
        pass

    def export(self, *a, **k):
        pass
