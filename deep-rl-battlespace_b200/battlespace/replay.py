"""Device-resident replay ring in the row layout of the reference's ``ReplayBuffer``
(maddpg/buffer.py:3-69): the fused rollout kernel writes the rows, ``sample`` returns the 7-tuple
``Team.learn`` unpacks (maddpg/team.py:30).  float32 on the device instead of float64 numpy."""
from __future__ import annotations

import ctypes as C

import torch

from ._lib import bs_replay, check, lib


class DeviceReplay:
    def __init__(self, mem_size, batch_size, n_agents, obs_size, n_actions, device):
        self.mem_size, self.batch_size, self.n_agents = int(mem_size), int(batch_size), int(n_agents)
        self.obs_size, self.n_actions = int(obs_size), int(n_actions)
        self.mem_cntr = 0
        self.device = torch.device(device)
        M, N, L, A = self.mem_size, self.n_agents, self.obs_size, self.n_actions
        z = lambda *s, dt=torch.float32: torch.zeros(s, dtype=dt, device=self.device)
        self.state_mem = z(M, N * L)          # buffer.py:11 ; actor_states[k] = state_mem[:, k*L:(k+1)*L]
        self.new_state_mem = z(M, N * L)      # buffer.py:12
        self.action_mem = z(M, N, A)          # buffer.py:23 (agent-major there, row-major here)
        self.rew_mem = z(M, N)                # buffer.py:13
        self.done_mem = z(M, N, dt=torch.uint8)  # buffer.py:14

    def as_struct(self):
        return bs_replay(*(C.c_void_p(t.data_ptr()) for t in
                           (self.state_mem, self.new_state_mem, self.action_mem, self.rew_mem, self.done_mem)),
                         self.mem_size)

    def is_ready(self):
        return self.mem_cntr >= self.batch_size

    def sample(self, generator=None, idx=None):
        """Uniform sample with replacement of filled rows (buffer.py:49-67) by one gather launch
        (``bs_replay_sample``).  Returns (actor_states [N,B,L], states [B,N*L], actions [N,B,A],
        rewards [B,N], actor_new_states [N,B,L], states_ [B,N*L], dones [B,N] bool) as device tensors."""
        max_mem = min(self.mem_cntr, self.mem_size)
        if idx is None:
            idx = torch.randint(0, max_mem, (self.batch_size,), device=self.device, generator=generator)
        idx = idx.to(device=self.device, dtype=torch.int64).contiguous()
        B, N, L, A = idx.numel(), self.n_agents, self.obs_size, self.n_actions
        e = lambda *s, dt=torch.float32: torch.empty(s, dtype=dt, device=self.device)
        actor_states, states, actions, rewards = e(N, B, L), e(B, N * L), e(N, B, A), e(B, N)
        actor_new_states, states_, dones = e(N, B, L), e(B, N * L), e(B, N, dt=torch.uint8)
        rp = self.as_struct()
        p = lambda t: C.c_void_p(t.data_ptr())
        check(lib.bs_replay_sample(C.byref(rp), N, L, A, p(idx), B, p(actor_states), p(states), p(actions), p(rewards),
                                   p(actor_new_states), p(states_), p(dones),
                                   C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)))
        return actor_states, states, actions, rewards, actor_new_states, states_, dones.bool()
