"""Host-side mirror of the reference's ``parallel_env`` (envs/battle_env.py:61-580) on top of the
CUDA library.  Two faces:

* the PettingZoo-1.19 Parallel API of the reference, dict in / dict out, for ``n_envs == 1`` --
  same constructor kwargs, attributes and return shapes, so ``main.py`` / ``evaluate.py`` /
  ``maddpg.Team`` / ``instinct.Team`` of the reference drive it unchanged;
* a batched tensor API (``reset_batched`` / ``step_batched`` / ``step_team_batched`` /
  ``rollout`` ...) over ``n_envs`` independent envs resident on one B200.

PyTorch is used only for device memory and streams; all env arithmetic is in
``libbattlespace.so`` (deep-rl-battlespace_b200/csrc).  There is no CPU fallback.
"""
from __future__ import annotations

import ctypes as C
import math
import sys
import os

import numpy as np
import torch

from . import _lib
from ._lib import (ACT_F32, ACT_F64, ACT_I32, ACT_NONE, POLICIES, WINNER_NAMES, bs_config, bs_replay,
                   bs_state_view, check, lib)

DISP_WIDTH, DISP_HEIGHT = 1200, 800  # envs/sprites.py:9-10


class _Box:
    """gym.spaces.Box stand-in: only ``.shape/.low/.high/.dtype`` are read by callers (main.py:115)."""

    def __init__(self, low, high, shape=None, dtype=np.float32):
        shape = np.shape(low) if shape is None else tuple(shape)
        self.shape, self.dtype = tuple(shape), np.dtype(dtype)
        self.low = np.broadcast_to(np.asarray(low, dtype=dtype), self.shape).copy()
        self.high = np.broadcast_to(np.asarray(high, dtype=dtype), self.shape).copy()


class _Discrete:
    def __init__(self, n):
        self.n, self.shape, self.dtype = int(n), (), np.dtype(np.int64)


def _ptr(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def _np_ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class parallel_env:
    """Drop-in for ``envs.battle_env.parallel_env`` (battle_env.py:61) plus a batched view."""

    metadata = {"render_modes": ["human"], "name": "battle_env_v1"}

    def __init__(self, n_agents=1, show=False, hit_base_reward=100, hit_plane_reward=10, miss_punishment=-1,
                 die_punishment=-5, lose_punishment=-20, fps=20, continuous_actions=False, *, n_envs=1,
                 device=None, seed=None, env_offset=0):
        if not torch.cuda.is_available():
            raise RuntimeError("battlespace needs a CUDA device (B200); there is no CPU fallback")
        self.n_agents = int(n_agents)
        N = self.n_agents
        self.n_envs = int(n_envs)
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        if self.device.index is None:
            self.device = torch.device("cuda", torch.cuda.current_device())
        # seed=None: a fresh seed per env object, like the reference's unseeded `random` module (independent runs
        # must not replay the same spawn / spread stream); pass an int for reproducible runs
        self.seed = int(seed) if seed is not None else int.from_bytes(os.urandom(8), "little") >> 1
        self.env_offset = int(env_offset)

        # --- attributes of the reference object, battle_env.py:86-184 --------------------------
        self.base_hp, self.plane_hp = 5 * N, 4
        self.possible_agents = [f"plane{r}" for r in range(2 * N)]
        self.possible_red = self.possible_agents[:N]
        self.possible_blue = self.possible_agents[N:]
        self.agents = self.possible_agents[:]
        self.team_map = {a: ("red" if i < N else "blue") for i, a in enumerate(self.possible_agents)}
        self.team = {"red": {"wins": 0}, "blue": {"wins": 0}}
        self.obs_size = 3 * N + 2
        high = np.ones(self.obs_size, dtype=np.float32)
        obs_space = _Box(high, -high)  # sic: the reference builds Box(high, -high), battle_env.py:134
        self.observation_spaces = {a: obs_space for a in self.possible_agents}
        self.continuous_actions = bool(continuous_actions)
        if self.continuous_actions:
            self.n_actions, self.max_turn, self.max_speed, self.min_speed = 3, 35, 275, 200
            action_space = _Box(-1, 1, shape=(3,), dtype=np.float32)
        else:
            self.n_actions, self.step_turn, self.speed = 4, 15, 215
            action_space = _Discrete(4)
        self.action_spaces = {a: action_space for a in self.possible_agents}
        self.width, self.height = DISP_WIDTH, DISP_HEIGHT
        self.max_time = 10 + 2 * N
        self.total_games = self.ties = 0
        self.bullet_speed, self.shot_dist, self.time_step = 450, 500, 0.1
        self.show, self.fps, self.recording = show, fps, False
        self.explosions = []                    # battle_env.py:172: [x, y, frame] of the running animations
        self.hit_base_reward, self.hit_plane_reward = hit_base_reward, hit_plane_reward
        self.miss_punishment, self.die_punishment, self.lose_punishment = miss_punishment, die_punishment, lose_punishment
        self.winner, self.env_done = "none", True
        self.dones = {a: False for a in self.possible_agents}

        # --- native handle -------------------------------------------------------------------
        cfg = bs_config(N, int(self.continuous_actions), float(hit_base_reward), float(hit_plane_reward),
                        float(miss_punishment), float(die_punishment), float(lose_punishment))
        h = C.c_void_p()
        check(lib.bs_create(C.byref(cfg), self.n_envs, self.env_offset, self.device.index, self.seed, C.byref(h)))
        self._h = h
        E, P, L = self.n_envs, 2 * N, self.obs_size
        self._E, self._P, self._L = E, P, L
        dev = self.device
        self.obs_buf = torch.empty((E, P, L), dtype=torch.float32, device=dev)
        self.rew_buf = torch.empty((E, P), dtype=torch.float32, device=dev)
        self.done_buf = torch.empty((E, P), dtype=torch.uint8, device=dev)
        self.env_done_buf = torch.ones((E,), dtype=torch.uint8, device=dev)
        self.winner_buf = torch.zeros((E,), dtype=torch.int8, device=dev)
        self._cursor = 0

    # ------------------------------------------------------------------------------------------
    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h and not sys.is_finalizing():   # at interpreter shutdown the CUDA context may already be gone: leave it to the driver
            try:
                lib.bs_destroy(h)
            except Exception:
                pass

    def close_handle(self):
        """Explicit, deterministic release of the native handle (bs_destroy synchronises the device)."""
        h, self._h = getattr(self, "_h", None), None
        if h:
            check(lib.bs_destroy(h))

    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def _host_call_barrier(self):
        """The *_host entry points run on the handle's own streams: everything queued on the caller's torch
        stream for this env (batched API calls, replay-ring fills) must have completed before they start."""
        torch.cuda.current_stream(self.device).synchronize()

    def observation_space(self, agent):
        return self.observation_spaces[agent]

    def action_space(self, agent):
        return self.action_spaces[agent]

    # ================================ dict API (n_envs == 1) ====================================
    def _need_single(self):
        if self.n_envs != 1:
            raise ValueError("the dict API (reset/step) is for n_envs == 1; use reset_batched/step_batched")

    def _obs_dict(self, obs):
        return {a: obs[0, i].copy() for i, a in enumerate(self.possible_agents)}

    def _sync_counters(self):
        st = self.stats()
        self.team["red"]["wins"], self.team["blue"]["wins"] = int(st[0]), int(st[1])
        self.ties, self.total_games = int(st[2]), int(st[3])

    def reset(self, seed=None, return_info=False, options=None, *, draws=None):
        """battle_env.py:246-279.  ``draws`` (int32 [4+6N]) injects the spawn values (parity tests)."""
        self._need_single()
        obs = np.empty((1, self._P, self._L), np.float32)
        d = None if draws is None else np.ascontiguousarray(draws, np.int32).reshape(1, 4 + 6 * self.n_agents)
        self._host_call_barrier()
        check(lib.bs_reset_host(self._h, None, _np_ptr(d), _np_ptr(obs)))
        self.winner, self.env_done = "none", False
        self.agents = self.possible_agents[:]
        self.dones = {a: False for a in self.possible_agents}
        return self._obs_dict(obs)

    def step(self, actions, *, draws=None):
        """battle_env.py:281-381: returns (observations, rewards, dones, infos) keyed by every
        possible agent.  Continuous actions are clipped in the caller's dict (battle_env.py:295-297);
        a missing key for a live agent raises KeyError (battle_env.py:326)."""
        self._need_single()
        N, P, L = self.n_agents, self._P, self._L
        if self.continuous_actions:
            for key, value in actions.items():
                sp = self.action_space(key)
                actions[key] = np.clip(value, sp.low, sp.high)
        dtype, arr = ACT_NONE, None
        if len(actions) != 0:
            live = [] if self.env_done else self.agents
            if self.continuous_actions:
                arr = np.zeros((1, P, 3), np.float64)
                for a in live:
                    arr[0, self.possible_agents.index(a)] = np.asarray(actions[a], np.float64).reshape(3)
                dtype = ACT_F64
            else:
                arr = np.zeros((1, P), np.int32)
                for a in live:
                    v = actions[a]
                    arr[0, self.possible_agents.index(a)] = int(np.argmax(v)) if isinstance(v, np.ndarray) else int(v)
                dtype = ACT_I32
        obs = np.empty((1, P, L), np.float32); rew = np.empty((1, P), np.float32)
        done = np.empty((1, P), np.uint8); env_done = np.empty(1, np.uint8); winner = np.empty(1, np.int8)
        d = None if draws is None else np.ascontiguousarray(draws, np.float64).reshape(1, P, 5)
        self._host_call_barrier()
        check(lib.bs_step_host(self._h, _np_ptr(arr), dtype, _np_ptr(d), _np_ptr(obs), _np_ptr(rew), _np_ptr(done),
                               _np_ptr(env_done), _np_ptr(winner)))
        was_done = self.env_done
        self.env_done = bool(env_done[0])
        self.winner = WINNER_NAMES[int(winner[0])]
        if self.env_done and not was_done:
            self.dones = {a: True for a in self.possible_agents}  # rebound like tie()/win(), battle_env.py:478,494
            self._sync_counters()
        else:
            for i, a in enumerate(self.possible_agents):
                self.dones[a] = bool(done[0, i])
        if not self.env_done:
            self.agents = [a for i, a in enumerate(self.possible_agents) if not done[0, i]]
        rewards = {a: (float(rew[0, i]) if rew[0, i] != 0 else 0) for i, a in enumerate(self.possible_agents)}
        infos = {a: {} for a in self.possible_agents}
        if self.show:  # battle_env.py:375-376
            self.render()
        return self._obs_dict(obs), rewards, self.dones, infos

    def wins(self):
        """battle_env.py:449-455"""
        return "Wins by red: {}\nWins by blue: {}\nTied games: {}\nWin rate: {}".format(
            self.team["red"]["wins"], self.team["blue"]["wins"], self.ties, self.team["red"]["wins"] / self.total_games)

    def make_discrete(self, actions_dict):
        return {a: np.argmax(v) for a, v in actions_dict.items()}

    # ----- rendering (battle_env.py:498-581): headless CUDA frames, no window ---------
    def render_frame(self, env=0, out=None, sprites=None, explosions=None):
        """RGB uint8 [800,1200,3] device tensor of env ``env`` as ``render()`` would draw it.
        ``explosions``: (x, y, frame) triples, the caller's ``self.explosions`` (battle_env.py:356,534-536)."""
        from .render import Sprites
        if sprites is not None:
            self._sprites = sprites
        if getattr(self, "_sprites", None) is None:
            self._sprites = Sprites(self.device, os.environ.get("BATTLESPACE_ASSETS"))
        if out is None:
            out = torch.empty((self.height, self.width, 3), dtype=torch.uint8, device=self.device)
        check(lib.bs_render(self._h, int(env), C.byref(self._sprites.with_explosions(explosions)), _ptr(out), self._stream()))
        return out

    def _track_explosions(self, env=0):
        """self.explosions of the reference for env ``env``: a plane whose hp fell to 0 since the last
        rendered frame of the same episode starts an animation at its centre (battle_env.py:351-356)."""
        st = self.get_state()
        hp, ep = st["plane_hp"][env].astype(np.int64), int(st["episode"][env])
        prev = getattr(self, "_expl_prev", None)
        if prev is not None and prev[0] == ep:
            for i in np.nonzero((prev[1] > 0) & (hp <= 0))[0]:
                self.explosions.append([int(st["plane_x"][env, i]), int(st["plane_y"][env, i]), 0])
        self._expl_prev = (ep, hp)

    def render(self, mode="human"):
        """The reference opens a pygame window (unavailable here); this keeps its contract for the
        callers that matter: nothing unless ``show``; when recording, one frame of env 0 per call."""
        if not self.show:
            return None
        self._track_explosions(0)
        frame = self.render_frame(0, explosions=self.explosions)
        for ex in self.explosions:          # Explosion.draw advances its frame on every draw, sprites.py:405-419
            ex[2] += 1
        self.explosions = [ex for ex in self.explosions if ex[2] < 9]
        if self.recording and getattr(self, "video", None) is not None:
            self.video.update(frame.cpu().numpy())
        return frame

    def winner_screen(self):
        return None

    def start_recording(self, path):
        from .render import VideoSink
        self.recording = True
        self.video = VideoSink(path, self.fps)

    def export_video(self):
        if self.recording:
            self.recording = False
            self.video.export()
        else:
            print("Not recording!")

    def close(self):
        return None

    # ===================================== batched API ==========================================
    def reset_batched(self, mask=None, draws=None):
        """Reset the envs with ``mask[e] != 0`` (None: all).  Returns obs [E,2N,L] (device tensor)."""
        if mask is not None:
            mask = mask.to(device=self.device, dtype=torch.uint8).contiguous()
        if draws is not None:
            draws = torch.as_tensor(draws, dtype=torch.int32, device=self.device).contiguous()
            assert draws.numel() == self._E * (4 + 6 * self.n_agents)
        check(lib.bs_reset(self._h, _ptr(mask), _ptr(draws), _ptr(self.obs_buf), self._stream()))
        if mask is None:
            self.env_done_buf.zero_(); self.winner_buf.zero_()
        else:
            self.env_done_buf.masked_fill_(mask.bool(), 0); self.winner_buf.masked_fill_(mask.bool(), 0)
        return self.obs_buf

    def reset_done(self):
        """Vectorised ``if env.env_done: env.reset()``."""
        return self.reset_batched(self.env_done_buf.clone())

    def _action_dtype(self, actions, agents):
        if actions is None:
            return ACT_NONE
        E, A = self._E, self.n_actions
        if actions.device != self.device or not actions.is_contiguous():
            raise ValueError("actions must be a contiguous tensor on the env's device")
        if actions.dtype == torch.int32 and not self.continuous_actions and actions.numel() == E * agents:
            return ACT_I32
        if actions.dtype == torch.float32 and actions.numel() == E * agents * A:
            return ACT_F32
        if actions.dtype == torch.float64 and self.continuous_actions and actions.numel() == E * agents * 3:
            return ACT_F64
        raise ValueError(f"bad action tensor {tuple(actions.shape)} {actions.dtype} for this env")

    def _draws(self, draws):
        if draws is None:
            return None
        draws = torch.as_tensor(draws, dtype=torch.float64, device=self.device).contiguous()
        assert draws.numel() == self._E * self._P * 5
        return draws

    def step_batched(self, actions, draws=None):
        """One tick of every env.  actions: int32 [E,2N] | float32 [E,2N,A] | float64 [E,2N,3].
        Returns (obs, rew, done, env_done, winner): persistent device tensors, overwritten each call."""
        dt = self._action_dtype(actions, self._P)
        draws = self._draws(draws)
        check(lib.bs_step(self._h, _ptr(actions), dt, _ptr(draws), _ptr(self.obs_buf), _ptr(self.rew_buf),
                          _ptr(self.done_buf), _ptr(self.env_done_buf), _ptr(self.winner_buf), self._stream()))
        return self.obs_buf, self.rew_buf, self.done_buf, self.env_done_buf, self.winner_buf

    def step_team_batched(self, team_actions, learner_team=0, other_policy="instinct", draws=None, replay=None,
                          cursor=None, auto_reset=False):
        """One tick with ``learner_team``'s actions ([E,N,...]) from the caller and the other team
        played in-kernel (main.py:179-181 with the instinct team fused into the step).  With
        ``replay`` (a DeviceReplay) the same launch also stores the learner team's transition rows
        (main.py:205).  Either call ``reset_done()`` between ticks like the reference's game loop, or
        pass ``auto_reset=True``: envs that finish are reset inside the same launch and ``obs`` is then
        the new game's first observation (rew/done/env_done/winner still describe the finished one)."""
        dt = self._action_dtype(team_actions, self.n_agents)
        draws = self._draws(draws)
        rp = None
        if replay is not None:
            rp = replay.as_struct()
            if cursor is None:
                cursor = replay.mem_cntr
        check(lib.bs_step_team_replay(self._h, int(learner_team), _ptr(team_actions), dt, POLICIES[other_policy],
                                      _ptr(draws), C.byref(rp) if rp is not None else None, int(cursor or 0),
                                      int(bool(auto_reset)), _ptr(self.obs_buf), _ptr(self.rew_buf), _ptr(self.done_buf),
                                      _ptr(self.env_done_buf), _ptr(self.winner_buf), self._stream()))
        if replay is not None:
            replay.mem_cntr += self._E
        return self.obs_buf, self.rew_buf, self.done_buf, self.env_done_buf, self.winner_buf

    def instinct_actions(self, team, obs=None, draws=None):
        """instinct Team.choose_actions for every env: [E,N,3] float64 or [E,N] int32 (device)."""
        obs = self.obs_buf if obs is None else torch.as_tensor(obs, dtype=torch.float32, device=self.device).contiguous()
        draws = self._draws(draws)
        out = (torch.empty((self._E, self.n_agents, 3), dtype=torch.float64, device=self.device)
               if self.continuous_actions else torch.empty((self._E, self.n_agents), dtype=torch.int32, device=self.device))
        check(lib.bs_instinct_actions(self._h, int(team), _ptr(obs), _ptr(draws), _ptr(out), self._stream()))
        return out

    def _check_ext_actions(self, T, ext_actions, on_device):
        """The learner's action block [T,E,N,A] float32 (BS_POL_EXTERNAL): the library reads it as a raw pointer."""
        if ext_actions is None:
            return
        want = int(T) * self._E * self.n_agents * self.n_actions
        if ext_actions.dtype != torch.float32 or ext_actions.numel() != want or not ext_actions.is_contiguous():
            raise ValueError(f"ext_actions must be a contiguous float32 tensor [T,E,N,A] = {want} values, got "
                             f"{tuple(ext_actions.shape)} {ext_actions.dtype}")
        if on_device and ext_actions.device != self.device:
            raise ValueError("ext_actions must live on the env's device")
        if not on_device and ext_actions.device.type != "cpu":
            raise ValueError("ext_actions_host must be a CPU (preferably pinned) tensor")

    def _check_replay(self, replay, T):
        """Ring rules of the C ABI (include/battlespace.h, bs_replay): rows of different envs never collide."""
        if replay is None:
            return
        if replay.mem_size < self._E or (int(T) * self._E > replay.mem_size and replay.mem_size % self._E != 0):
            raise ValueError(f"replay ring of {replay.mem_size} rows cannot take {T} x {self._E} rows per call: it needs "
                             "capacity >= n_envs, and capacity % n_envs == 0 when one call wraps the ring")

    def rollout(self, T, red="random", blue="instinct", learner_team=0, ext_actions=None, replay=None, cursor=None):
        """T fused ticks with in-kernel policies, auto-reset and replay rows (main.py:178-213)."""
        self._check_ext_actions(T, ext_actions, True)
        self._check_replay(replay, T)
        rp = None
        if replay is not None:
            rp = replay.as_struct()
            if cursor is None:
                cursor = replay.mem_cntr
        check(lib.bs_rollout_fused(self._h, int(T), int(learner_team), POLICIES[red], POLICIES[blue], _ptr(ext_actions),
                                   C.byref(rp) if rp is not None else None, int(cursor or 0), self._stream()))
        if replay is not None:
            replay.mem_cntr += int(T) * self._E

    def rollout_host(self, T, red="random", blue="instinct", learner_team=0, ext_actions_host=None, replay=None, cursor=None):
        """Same through the host-buffer entry point: ``ext_actions_host`` is a (pinned) CPU tensor
        [T,E,N,A]; returns the 8 counters as numpy int64 after the device work completed."""
        self._check_ext_actions(T, ext_actions_host, False)
        self._check_replay(replay, T)
        self._host_call_barrier()
        rp = None
        if replay is not None:
            rp = replay.as_struct()
            if cursor is None:
                cursor = replay.mem_cntr
        out = np.zeros(8, np.int64)
        check(lib.bs_rollout_fused_host(self._h, int(T), int(learner_team), POLICIES[red], POLICIES[blue],
                                        _ptr(ext_actions_host), C.byref(rp) if rp is not None else None,
                                        int(cursor or 0), _np_ptr(out)))
        if replay is not None:
            replay.mem_cntr += int(T) * self._E
        return out

    def rollout_host_begin(self, T, red="random", blue="instinct", learner_team=0, ext_actions_host=None, replay=None,
                           cursor=None):
        """First phase of ``rollout_host``: returns once ``ext_actions_host`` has been copied to the device (the
        caller may refill it for the next call while the kernels run).  Pair with ``rollout_host_end``; at most
        two calls may be in flight."""
        self._check_ext_actions(T, ext_actions_host, False)
        self._check_replay(replay, T)
        if getattr(self, "_host_inflight", 0) == 0:
            self._host_call_barrier()
        rp = None
        if replay is not None:
            rp = replay.as_struct()
            if cursor is None:
                cursor = replay.mem_cntr
        check(lib.bs_rollout_fused_host_begin(self._h, int(T), int(learner_team), POLICIES[red], POLICIES[blue],
                                              _ptr(ext_actions_host), C.byref(rp) if rp is not None else None,
                                              int(cursor or 0)))
        self._host_inflight = getattr(self, "_host_inflight", 0) + 1
        if replay is not None:
            replay.mem_cntr += int(T) * self._E

    def rollout_host_end(self):
        """Second phase: waits for the oldest call in flight, returns the 8 counters as they stood after it."""
        out = np.zeros(8, np.int64)
        check(lib.bs_rollout_fused_host_end(self._h, _np_ptr(out)))
        self._host_inflight -= 1
        return out

    # ----- learner team's actors in CUDA -----------------------------------------------
    def actor_actions(self, actors, team=0, obs=None, noise=None, normals=None, reset_mask=None, out=None):
        """Batched ``Team.choose_actions`` (maddpg/team.py:21-28): float32 [E,N,A] actions of ``team``'s
        planes from ``actors`` (a TeamActors) on ``obs`` [E,2N,L] (default: the env's last observation).
        ``noise``: a DeviceOUNoise to advance and add (agent.py:30-32); ``normals`` injects its draws."""
        obs = self.obs_buf if obs is None else obs
        if out is None:
            out = torch.empty((self._E, self.n_agents, self.n_actions), dtype=torch.float32, device=self.device)
        nz = None
        if noise is not None:
            nz = noise.as_struct(normals, reset_mask)
            noise.step += 1
        check(lib.bs_actor_forward(self._h, int(team), actors.as_array(), int(actors.hidden), _ptr(obs),
                                   C.byref(nz) if nz is not None else None, _ptr(out), self._stream()))
        return out

    def rollout_actor(self, T, actors, learner_team=0, other_policy="instinct", noise=None, replay=None, cursor=None):
        """T iterations of the training game loop (main.py:178-213) with ``learner_team`` played by
        ``actors``: per tick actor forward -> step + store_transition + auto-reset, all on the device,
        one native call.  Continues from the env's last observation (call ``reset_batched()`` first)."""
        if not hasattr(self, "_act_buf"):
            self._act_buf = torch.empty((self._E, self.n_agents, self.n_actions), dtype=torch.float32, device=self.device)
        rp = None
        if replay is not None:
            rp = replay.as_struct()
            if cursor is None:
                cursor = replay.mem_cntr
        nz = None
        if noise is not None:
            nz = noise.as_struct(None, self.env_done_buf)
            noise.step += int(T)
        check(lib.bs_rollout_actor(self._h, int(T), int(learner_team), actors.as_array(), int(actors.hidden),
                                   POLICIES[other_policy], C.byref(nz) if nz is not None else None,
                                   C.byref(rp) if rp is not None else None, int(cursor or 0), _ptr(self.obs_buf),
                                   _ptr(self._act_buf), _ptr(self.rew_buf), _ptr(self.done_buf), _ptr(self.env_done_buf),
                                   _ptr(self.winner_buf), self._stream()))
        if replay is not None:
            replay.mem_cntr += int(T) * self._E
        return self.obs_buf, self.rew_buf, self.done_buf, self.env_done_buf, self.winner_buf

    # ----- statistics (battle_env.py:102-103,169-170; main.py:218-253) ---------------------------
    def stats(self):
        """int64[8]: red wins, blue wins, ties, total_games, sum red return*2^20, sum blue return*2^20,
        env ticks, agent steps."""
        out = np.zeros(8, np.int64)
        check(lib.bs_stats(self._h, _np_ptr(out), self._stream()))
        return out

    def stats_tensor(self):
        """Zero-copy int64[8] device view of the same counters (input to the NCCL all-reduce)."""
        p = C.c_void_p()
        check(lib.bs_stats_device(self._h, C.byref(p)))

        class _View:
            __cuda_array_interface__ = {"shape": (8,), "typestr": "<i8", "data": (p.value, False), "version": 3}
        return torch.as_tensor(_View(), device=self.device)

    # ----- state export / import -------------------------------------------------------------------
    def _state_arrays(self):
        E, P, S = self._E, self._P, _lib.BULLET_SLOTS
        return dict(plane_x=np.zeros((E, P), np.int16), plane_y=np.zeros((E, P), np.int16),
                    plane_dir=np.zeros((E, P), np.float64), plane_hp=np.zeros((E, P), np.int8),
                    base_pos=np.zeros((E, 2, 2), np.int16), base_hp=np.zeros((E, 2), np.int16),
                    bullet_pos=np.zeros((E, P, S, 2), np.int16), bullet_vel=np.zeros((E, P, S, 3), np.int8),
                    tick=np.zeros(E, np.int32), env_done=np.zeros(E, np.uint8), winner=np.zeros(E, np.int8),
                    episode=np.zeros(E, np.uint32), team_return=np.zeros((E, 2), np.float64))

    def get_state(self):
        st = self._state_arrays()
        view = bs_state_view(*[st[k].ctypes.data for k in _lib.STATE_FIELDS])
        check(lib.bs_get_state(self._h, C.byref(view)))
        return st

    def set_state(self, st):
        ref = self._state_arrays()
        arrs = {k: np.ascontiguousarray(st[k], ref[k].dtype).reshape(ref[k].shape) for k in _lib.STATE_FIELDS}
        view = bs_state_view(*[arrs[k].ctypes.data for k in _lib.STATE_FIELDS])
        check(lib.bs_set_state(self._h, C.byref(view)))
        self.env_done_buf.copy_(torch.from_numpy(arrs["env_done"]))
        self.winner_buf.copy_(torch.from_numpy(arrs["winner"]))
        if self.n_envs == 1:
            self.env_done = bool(arrs["env_done"][0]); self.winner = WINNER_NAMES[int(arrs["winner"][0])]


BattleEnv = parallel_env


def tie_tick(n_agents: int) -> int:
    """Step call on which the time-out tie fires: 0.1 accumulated in double (battle_env.py:316-319)."""
    t, k = 0.0, 0
    while True:
        k += 1
        t += 0.1
        if t >= 10 + 2 * n_agents:
            return k


DIAG = math.sqrt(math.pow(DISP_WIDTH, 2) + math.pow(DISP_HEIGHT, 2))
