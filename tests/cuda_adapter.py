"""Adapters that give the CUDA env the same call surface as oracle/port.py:OracleEnv, so the same
replay/compare code drives both.  Everything goes through the C ABI (include/battlespace.h):
mode "device" = device-pointer entry points with torch tensors, mode "host" = *_host entry points."""
import ctypes as C

import numpy as np
import torch

from battlespace import parallel_env
from battlespace._lib import ACT_F32, ACT_F64, ACT_I32, ACT_NONE, check, lib

KEYS = ("hit_base_reward", "hit_plane_reward", "miss_punishment", "die_punishment", "lose_punishment")


def _np(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class CudaEnv:
    def __init__(self, n_agents, continuous, rewards, n_envs, env_offset=0, seed=0, mode="device"):
        self.env = parallel_env(n_agents=n_agents, continuous_actions=continuous, n_envs=n_envs, seed=seed,
                                env_offset=env_offset, **dict(zip(KEYS, [float(r) for r in rewards])))
        self.N, self.P, self.L, self.E = n_agents, 2 * n_agents, 3 * n_agents + 2, n_envs
        self.cont, self.A, self.mode = bool(continuous), 3 if continuous else 4, mode
        self.dev = self.env.device

    def reset(self, mask=None, draws=None):
        if self.mode == "host":
            obs = np.empty((self.E, self.P, self.L), np.float32)
            m = None if mask is None else np.ascontiguousarray(mask, np.uint8)
            d = None if draws is None else np.ascontiguousarray(draws, np.int32)
            check(lib.bs_reset_host(self.env._h, _np(m), _np(d), _np(obs)))
            return obs
        m = None if mask is None else torch.as_tensor(np.asarray(mask, np.uint8), device=self.dev)
        d = None if draws is None else torch.as_tensor(np.asarray(draws, np.int32), device=self.dev)
        return self.env.reset_batched(m, d).cpu().numpy()

    def step(self, actions, draws=None):
        E, P, L = self.E, self.P, self.L
        if self.mode == "host":
            a = None if actions is None else np.ascontiguousarray(actions)
            dt = ACT_NONE if a is None else {np.dtype(np.int32): ACT_I32, np.dtype(np.float32): ACT_F32,
                                             np.dtype(np.float64): ACT_F64}[a.dtype]
            d = None if draws is None else np.ascontiguousarray(draws, np.float64)
            obs = np.empty((E, P, L), np.float32); rew = np.empty((E, P), np.float32)
            done = np.empty((E, P), np.uint8); ed = np.empty(E, np.uint8); win = np.empty(E, np.int8)
            check(lib.bs_step_host(self.env._h, _np(a), dt, _np(d), _np(obs), _np(rew), _np(done), _np(ed), _np(win)))
            return obs, rew, done.astype(bool), ed.astype(bool), win
        a = None if actions is None else torch.as_tensor(np.ascontiguousarray(actions), device=self.dev)
        d = None if draws is None else torch.as_tensor(np.asarray(draws, np.float64), device=self.dev)
        obs, rew, done, ed, win = self.env.step_batched(a, d)
        return (obs.cpu().numpy(), rew.cpu().numpy(), done.cpu().numpy().astype(bool),
                ed.cpu().numpy().astype(bool), win.cpu().numpy())

    def step_team(self, team_actions, learner_team, other_policy, draws=None):
        a = torch.as_tensor(np.ascontiguousarray(team_actions), device=self.dev)
        d = None if draws is None else torch.as_tensor(np.asarray(draws, np.float64), device=self.dev)
        obs, rew, done, ed, win = self.env.step_team_batched(a, learner_team, other_policy, d)
        return (obs.cpu().numpy(), rew.cpu().numpy(), done.cpu().numpy().astype(bool),
                ed.cpu().numpy().astype(bool), win.cpu().numpy())

    def instinct(self, team, obs, draws=None):
        obs = np.ascontiguousarray(obs, np.float32)
        if self.mode == "host":
            d = None if draws is None else np.ascontiguousarray(draws, np.float64)
            out = np.empty((self.E, self.N, 3), np.float64) if self.cont else np.empty((self.E, self.N), np.int32)
            check(lib.bs_instinct_actions_host(self.env._h, team, _np(obs), _np(d), _np(out)))
            return out
        d = None if draws is None else torch.as_tensor(np.asarray(draws, np.float64), device=self.dev)
        return self.env.instinct_actions(team, torch.as_tensor(obs, device=self.dev), d).cpu().numpy()

    def get_state(self):
        st = self.env.get_state()
        return dict(px=st["plane_x"], py=st["plane_y"], pdir=st["plane_dir"], php=st["plane_hp"],
                    base_pos=st["base_pos"], bhp=st["base_hp"], bullets=st["bullet_pos"], tick=st["tick"],
                    env_done=st["env_done"], winner=st["winner"], episode=st["episode"], raw=st)

    def stats(self):
        return self.env.stats()
