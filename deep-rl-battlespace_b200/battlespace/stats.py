"""Episode statistics / report path: the ``score_dict`` history and the print-interval
report of the reference's driver (main.py:218-268, scores.json) from the device-side episode log that
every kernel appends to when a game ends (``bs_episode_log``)."""
from __future__ import annotations

import ctypes as C
import json

import numpy as np
import torch

from ._lib import check, lib

CODE_RED, CODE_BLUE, CODE_TIE, CODE_BOTH = 1, 2, 3, 6


class EpisodeLog:
    def __init__(self, env, capacity=1 << 20):
        self.env, self.capacity = env, 1 << max(int(capacity) - 1, 0).bit_length()  # the ring is a power of two
        self.buf = torch.zeros((self.capacity, 3), dtype=torch.float32, device=env.device)
        check(lib.bs_episode_log(env._h, C.c_void_p(self.buf.data_ptr()), self.capacity))
        self._read = 0  # games already handed out by new_games()

    def detach(self):
        check(lib.bs_episode_log(self.env._h, None, 0))

    def count(self):
        n = C.c_int64()
        check(lib.bs_episode_log_count(self.env._h, C.byref(n), self.env._stream()))
        return int(n.value)

    def new_games(self):
        """Games logged since the last call, as a float32 array [n,3] (red score, blue score, code).
        Raises if the ring wrapped past unread entries."""
        n = self.count()
        if n - self._read > self.capacity:
            raise RuntimeError("episode log overflowed: read it more often or make it larger")
        idx = (torch.arange(self._read, n, device=self.buf.device) % self.capacity)
        self._read = n
        return self.buf[idx].cpu().numpy()

    @staticmethod
    def score_dict(games):
        """main.py:219-221: {'red': [...], 'blue': [...]} with scores rounded to 3 places."""
        return {"red": [round(float(x), 3) for x in games[:, 0]], "blue": [round(float(x), 3) for x in games[:, 1]]}

    @staticmethod
    def save_scores(path, score_dict):
        with open(path, "w") as f:   # save_dict, main.py:53-55
            json.dump(score_dict, f)

    @staticmethod
    def report(games, total_games, timesteps):
        """The numbers of the print-interval statement (main.py:231-253) over ``games``."""
        n = max(len(games), 1)
        code = games[:, 2].astype(np.int64)
        return {"games": int(total_games), "timesteps": int(timesteps),
                "red_avg_score": round(float(np.mean(games[:, 0])) if len(games) else 0.0, 3),
                "blue_avg_score": round(float(np.mean(games[:, 1])) if len(games) else 0.0, 3),
                "red_winrate": round(float(np.sum(code == CODE_RED)) / n, 3),
                "blue_winrate": round(float(np.sum((code == CODE_BLUE) | (code == CODE_BOTH))) / n, 3),  # winner ends 'blue'
                "tie_rate": round(float(np.sum(code == CODE_TIE)) / n, 3)}
