"""Device-side mirror of the learner team's actors.

``TeamActors`` holds the parameters of the N ``ActorNetwork``s of one ``maddpg.Team``
(maddpg/networks.py:54-85, one per plane: maddpg/team.py:17-19) as float32 device tensors and hands
them to ``bs_actor_forward`` / ``bs_rollout_actor`` -- the batched ``Team.choose_actions``
(maddpg/team.py:21-28) for every env at once.  ``DeviceOUNoise`` is ``utils/noise.py`` for every
(env, plane).  No torch module is evaluated here: the forward pass is the library's CUDA kernel.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from ._lib import ACTOR_FIELDS, bs_actor, bs_ou_noise

# state_dict names of ActorNetwork (networks.py:58-72) in bs_actor field order
STATE_KEYS = ("fc1.weight", "fc1.bias", "bn1.weight", "bn1.bias", "fc2.weight", "fc2.bias", "bn2.weight", "bn2.bias",
              "pi.weight", "pi.bias")


class TeamActors:
    def __init__(self, params, device):
        """params: list (one per plane of the team) of dicts keyed by ``STATE_KEYS``."""
        self.device = torch.device(device)
        self.n_agents = len(params)
        self.tensors = []
        for p in params:
            row = [torch.as_tensor(np.asarray(p[k]) if not torch.is_tensor(p[k]) else p[k]).detach()
                   .to(device=self.device, dtype=torch.float32).contiguous() for k in STATE_KEYS]
            self.tensors.append(row)
        t0 = self.tensors[0]
        self.hidden, self.obs_size, self.n_actions = t0[0].shape[0], t0[0].shape[1], t0[8].shape[0]
        for row in self.tensors:
            H, L, A = self.hidden, self.obs_size, self.n_actions
            want = [(H, L), (H,), (H,), (H,), (H, H), (H,), (H,), (H,), (A, H), (A,)]
            if [tuple(t.shape) for t in row] != want:
                raise ValueError("actor parameter shapes do not match ActorNetwork(obs_len, n_actions, 64, 64)")
        self._structs = (bs_actor * self.n_agents)()
        self._bind()

    def _bind(self):
        for k, row in enumerate(self.tensors):
            for f, t in zip(ACTOR_FIELDS, row):
                setattr(self._structs[k], f, t.data_ptr())

    @classmethod
    def from_modules(cls, actors, device=None):
        """actors: the team's ``agent.actor`` modules in team order (maddpg/team.py:17-19)."""
        sds = [a.state_dict() for a in actors]
        device = device or next(iter(sds[0].values())).device
        return cls([{k: sd[k] for k in STATE_KEYS} for sd in sds], device)

    @classmethod
    def from_npz(cls, npz, names, device):
        """npz with keys ``<name>.<state key>`` (tests/golden/completed_model_actors.npz)."""
        return cls([{k: npz[f"{n}.{k}"] for k in STATE_KEYS} for n in names], device)

    def load_(self, actors):
        """Copy the current parameters of the torch modules in (after a learner update)."""
        for row, a in zip(self.tensors, actors):
            sd = a.state_dict()
            for t, k in zip(row, STATE_KEYS):
                t.copy_(sd[k])

    def as_array(self):
        return self._structs


class DeviceOUNoise:
    """utils/noise.py OUNoise(action_dimension, scale=0.1, mu=0, theta=0.15, sigma=0.2) for [E,N,A]."""

    def __init__(self, n_envs, n_agents, n_actions, device, scale=0.1, mu=0.0, theta=0.15, sigma=0.2, seed=0):
        self.state = torch.full((n_envs, n_agents, n_actions), float(mu), dtype=torch.float32, device=device)
        self.scale, self.mu, self.theta, self.sigma, self.seed, self.step = scale, mu, theta, sigma, int(seed), 0

    def reset(self):
        self.state.fill_(float(self.mu))

    def as_struct(self, normals=None, reset_mask=None):
        ptr = lambda t: None if t is None else C.c_void_p(t.data_ptr())
        return bs_ou_noise(ptr(self.state), ptr(normals), ptr(reset_mask), self.mu, self.theta, self.sigma, self.scale,
                           self.seed, self.step)
