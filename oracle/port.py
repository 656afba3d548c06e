"""TEST INFRASTRUCTURE ONLY -- numpy/ctypes front end of the CPU oracle (oracle/liboracle.so).

Used by tests/ as the checker, by __graft_entry__.smoke() and by bench.py's
cpu_baseline / --impl reference legs.  Never imported by the product package.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "liboracle.so")

ACT_I32, ACT_F32, ACT_F64, ACT_NONE = 0, 1, 2, 3
POL_EXTERNAL, POL_RANDOM, POL_INSTINCT, POL_INSTINCT_FF = 0, 1, 2, 3
SLOTS = 12

DEFAULT_REWARDS = (100.0, 10.0, -1.0, -5.0, -20.0)        # battle_env.py:73
TRAINED_REWARDS = (1.0, 0.9, -0.02, -0.03, -0.05)         # main.py:33-37, models/completed_model/cf.json


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "battle_oracle.c")
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
        subprocess.run(["make", "-C", _HERE, "-B", "liboracle.so"], check=True, capture_output=True)
    return _LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB_PATH)
        L.bo_create.restype = C.c_void_p
        L.bo_create.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_int64, C.c_int64, C.c_uint64]
        L.bo_destroy.argtypes = [C.c_void_p]
        L.bo_reset.argtypes = [C.c_void_p] * 4
        L.bo_step.argtypes = [C.c_void_p, C.c_void_p, C.c_int] + [C.c_void_p] * 6
        L.bo_instinct.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        L.bo_rollout.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p] + [C.c_void_p] * 5 + [C.c_int64, C.c_int64]
        L.bo_get_state.argtypes = [C.c_void_p] * 12
        L.bo_stats.argtypes = [C.c_void_p, C.c_void_p]
        L.bo_tie_tick.restype = C.c_int
        L.bo_tie_tick.argtypes = [C.c_int]
        _lib = L
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class OracleEnv:
    """Batched CPU env with the same call surface as the C-ABI in include/battlespace.h."""

    def __init__(self, n_agents=1, continuous=False, rewards=DEFAULT_REWARDS, n_envs=1,
                 env_offset=0, seed=0):
        self.N, self.P, self.L = n_agents, 2 * n_agents, 3 * n_agents + 2
        self.cont = bool(continuous)
        self.A = 3 if self.cont else 4
        self.E = int(n_envs)
        rw = np.asarray(rewards, np.float64)
        assert rw.shape == (5,)
        self._h = lib().bo_create(n_agents, int(self.cont), _p(rw), self.E, int(env_offset), int(seed))
        if not self._h:
            raise ValueError("bo_create failed")

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h and _lib is not None:
            _lib.bo_destroy(h)

    def reset(self, mask=None, draws=None):
        obs = np.empty((self.E, self.P, self.L), np.float32)
        if mask is not None:
            mask = np.ascontiguousarray(mask, np.uint8)
        if draws is not None:
            draws = np.ascontiguousarray(draws, np.int32).reshape(self.E, 4 + 6 * self.N)
        lib().bo_reset(self._h, _p(mask), _p(draws), _p(obs))
        return obs

    def step(self, actions, draws=None):
        a = None if actions is None else np.ascontiguousarray(actions)
        if a is None:
            dt = ACT_NONE
        elif a.dtype == np.int32:
            dt = ACT_I32; assert a.size == self.E * self.P and not self.cont
        elif a.dtype == np.float32:
            dt = ACT_F32; assert a.size == self.E * self.P * self.A
        elif a.dtype == np.float64:
            dt = ACT_F64; assert a.size == self.E * self.P * 3 and self.cont
        else:
            raise TypeError(a.dtype)
        if draws is not None:
            draws = np.ascontiguousarray(draws, np.float64).reshape(self.E, self.P, 5)
        obs = np.empty((self.E, self.P, self.L), np.float32)
        rew = np.empty((self.E, self.P), np.float64)
        done = np.empty((self.E, self.P), np.uint8)
        env_done = np.empty(self.E, np.uint8)
        winner = np.empty(self.E, np.int8)
        lib().bo_step(self._h, _p(a), dt, _p(draws), _p(obs), _p(rew), _p(done), _p(env_done), _p(winner))
        return obs, rew, done.astype(bool), env_done.astype(bool), winner

    def instinct(self, team, obs, draws=None):
        obs = np.ascontiguousarray(obs, np.float32).reshape(self.E, self.P, self.L)
        if draws is not None:
            draws = np.ascontiguousarray(draws, np.float64).reshape(self.E, self.P, 5)
        out = (np.empty((self.E, self.N, 3), np.float64) if self.cont
               else np.empty((self.E, self.N), np.int32))
        lib().bo_instinct(self._h, int(team), _p(obs), _p(draws), _p(out))
        return out

    def rollout(self, T, policies, learner=0, ext_actions=None, capacity=0, cursor=0, replay=None):
        """Fused T-tick rollout; ``replay`` is a dict of preallocated numpy arrays or None
        (allocated here when capacity > 0).  Returns the replay dict."""
        pol = np.asarray(policies, np.int32)
        N, L, A = self.N, self.L, self.A
        if capacity and replay is None:
            replay = dict(state=np.zeros((capacity, N * L), np.float32),
                          new_state=np.zeros((capacity, N * L), np.float32),
                          action=np.zeros((capacity, N, A), np.float32),
                          reward=np.zeros((capacity, N), np.float32),
                          done=np.zeros((capacity, N), np.uint8))
        if ext_actions is not None:
            ext_actions = np.ascontiguousarray(ext_actions, np.float32)
            assert ext_actions.size == T * self.E * N * A
        r = replay or {}
        lib().bo_rollout(self._h, int(T), int(learner), _p(pol), _p(ext_actions),
                         _p(r.get("state")), _p(r.get("new_state")), _p(r.get("action")),
                         _p(r.get("reward")), _p(r.get("done")), int(capacity), int(cursor))
        return replay

    def get_state(self):
        E, P = self.E, self.P
        st = dict(px=np.empty((E, P), np.int16), py=np.empty((E, P), np.int16),
                  pdir=np.empty((E, P), np.float64), php=np.empty((E, P), np.int8),
                  base_pos=np.empty((E, 2, 2), np.int16), bhp=np.empty((E, 2), np.int16),
                  bullets=np.empty((E, P, SLOTS, 2), np.int16), tick=np.empty(E, np.int32),
                  env_done=np.empty(E, np.uint8), winner=np.empty(E, np.int8),
                  episode=np.empty(E, np.uint32))
        lib().bo_get_state(self._h, *[_p(st[k]) for k in ("px", "py", "pdir", "php", "base_pos", "bhp",
                                                           "bullets", "tick", "env_done", "winner", "episode")])
        return st

    def stats(self):
        out = np.zeros(8, np.int64)
        lib().bo_stats(self._h, _p(out))
        return out


def tie_tick(n_agents: int) -> int:
    return int(lib().bo_tie_tick(n_agents))
