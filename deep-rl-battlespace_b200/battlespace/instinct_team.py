"""Mirror of the reference's scripted opponent (instinct/team.py:3-14, instinct/agent.py:4-62):
``Team(agent_list, enemy_list, env).choose_actions(observations)`` with the policy evaluated by
the CUDA kernel behind ``bs_instinct_actions``.  Random draws come from the env's Philox streams
(or are injected with ``draws=`` for parity tests)."""
from __future__ import annotations

import ctypes as C

import numpy as np

from ._lib import check, lib


class Team:
    def __init__(self, agent_list, enemy_list, env):
        self.agent_list, self.enemy_list, self.env = list(agent_list), list(enemy_list), env
        self.team = 0 if self.agent_list[0] == env.possible_agents[0] else 1

    def choose_actions(self, observations, *, draws=None):
        env = self.env
        if env.n_envs != 1:
            raise ValueError("dict API is for n_envs == 1; use env.instinct_actions for batches")
        N, P, L = env.n_agents, 2 * env.n_agents, env.obs_size
        obs = np.full((1, P, L), -1.0, np.float32)
        for a in self.agent_list:
            obs[0, env.possible_agents.index(a)] = np.asarray(observations[a], np.float32)
        d = None if draws is None else np.ascontiguousarray(draws, np.float64).reshape(1, P, 5)
        out = np.empty((1, N, 3), np.float64) if env.continuous_actions else np.empty((1, N), np.int32)
        check(lib.bs_instinct_actions_host(env._h, self.team, obs.ctypes.data_as(C.c_void_p),
                                           None if d is None else d.ctypes.data_as(C.c_void_p),
                                           out.ctypes.data_as(C.c_void_p)))
        if env.continuous_actions:
            return {a: out[0, k].copy() for k, a in enumerate(self.agent_list)}
        return {a: int(out[0, k]) for k, a in enumerate(self.agent_list)}
