"""TEST INFRASTRUCTURE ONLY -- runs the UNMODIFIED reference env as the parity oracle.

Imports ``/root/reference/envs/battle_env.py``, ``envs/sprites.py`` and
``instinct/*.py`` exactly as shipped, with the four third-party modules they
need but this image lacks (pygame, gym, pettingzoo, vidmaker) replaced by the
stand-ins in ``oracle/stubs``.  Every random draw the reference consumes is
tapped and recorded together with the state the reference reaches, so the same
draws and actions can be replayed into the C restatement
(``oracle/battle_oracle.c``) and into the CUDA path.

Only usable where ``/root/reference`` exists (the build container).  The GPU box
never imports this file; it reads the committed fixtures under
``tests/golden/`` that ``oracle/gen_golden.py`` produced with it.

Pinned third-party semantics (reference requirements.txt):
  * pygame==2.1.2 Rect truncation / colliderect -> ``oracle/stubs/pygame``;
  * numpy==1.23.1 scalar promotion (float32 scalar op python number -> float64):
    this image has numpy 2.x, which keeps float32, so observations handed to the
    instinct policy (instinct/agent.py:15-24) and float32 actions handed to
    ``step`` (battle_env.py:419-421) are widened to float64 first -- an exact
    conversion that reproduces what 1.23.1 computes.
Draw taps:
  * ``envs.sprites.random`` (randint in Plane/Base.reset, sprites.py:82-91,
    246-251; random() in Bullet.__init__, sprites.py:314);
  * ``instinct.agent.np.random.rand`` / ``uniform`` (instinct/agent.py:44,53);
    the tapped ``uniform(lo, hi, n)`` returns ``lo + (hi - lo) * u`` with u the
    recorded unit draw, numpy's own definition of the legacy uniform.
"""
from __future__ import annotations

import contextlib
import math
import os
import random as _pyrandom
import sys

import numpy as np

REF_ROOT = os.environ.get("BATTLESPACE_REFERENCE", "/root/reference")
STUBS = os.path.join(os.path.dirname(os.path.abspath(__file__)), "stubs")

WINNER_CODE = {"none": 0, "red": 1, "blue": 2, "tie": 3}
BULLET_SLOTS = 12  # a bullet is removed on its 12th update (sprites.py:334-337)


def reference_available() -> bool:
    return os.path.isfile(os.path.join(REF_ROOT, "envs", "battle_env.py"))


_mods = None


def _load():
    """Import the reference modules once (stubs first on sys.path)."""
    global _mods
    if _mods is not None:
        return _mods
    if not reference_available():
        raise RuntimeError(f"reference tree not found at {REF_ROOT}")
    for p in (REF_ROOT, STUBS):
        if p not in sys.path:
            sys.path.insert(0, p)
    with _cwd(REF_ROOT):
        import envs.battle_env as be
        import envs.sprites as sprites
        import instinct.agent as iagent
        import instinct.team as iteam
    _mods = (be, sprites, iagent, iteam)
    return _mods


@contextlib.contextmanager
def _cwd(path):
    old = os.getcwd()
    os.chdir(path)  # sprites.py:61,226 open "assets/..." by relative path
    try:
        yield
    finally:
        os.chdir(old)


class _TapRandom:
    """Replaces the ``random`` module inside envs.sprites."""

    def __init__(self, seed):
        self._r = _pyrandom.Random(seed)
        self.ints = []       # randint results, in call order
        self.spread = []     # (agent_id, u) per Bullet.__init__
        self.cur_shooter = None

    def randint(self, a, b):
        v = self._r.randint(a, b)
        self.ints.append(v)
        return v

    def random(self):
        u = self._r.random()
        self.spread.append((self.cur_shooter, u))
        return u


class _TapNpRandom:
    def __init__(self, seed):
        self._rs = np.random.RandomState(seed)
        self.order = []      # agent ids in the order choose_action will be called
        self.records = {}    # agent_id -> [shoot_u, n0, n1, n2]
        self._shoot = math.nan

    def rand(self):
        u = float(self._rs.random_sample())
        self._shoot = u
        return u

    def uniform(self, low, high, size=None):
        u = self._rs.random_sample(size)
        agent = self.order.pop(0)
        self.records[agent] = [self._shoot] + [float(x) for x in np.atleast_1d(u)]
        self._shoot = math.nan
        return low + (high - low) * u


class _NpProxy:
    """numpy with ``random`` swapped for the tap (only instinct.agent sees it)."""

    def __init__(self, tap):
        self.random = tap

    def __getattr__(self, k):
        return getattr(np, k)


def _load_actors(agent_ids, obs_len, model="completed_model"):
    """ActorNetwork instances of the reference with the shipped weights (CPU, eval)."""
    import torch
    import maddpg.networks as nets
    out = {}
    for a in agent_ids:
        net = nets.ActorNetwork(obs_len, 4, 64, 64, 0.001, os.path.join(REF_ROOT, "models", model), f"actor_{a}")
        net.device = torch.device("cpu"); net.to("cpu")
        net.load_state_dict(torch.load(net.checkpoint_file, map_location="cpu"))
        net.eval()
        out[a] = net
    return out


def tie_tick(n_agents: int) -> int:
    """Step-call index on which the time-out tie fires (battle_env.py:316-319)."""
    t, k = 0.0, 0
    while True:
        k += 1
        t += 0.1
        if t >= 10 + 2 * n_agents:
            return k


def run_reference(n_agents=1, continuous=False, rewards=None, red="random", blue="random",
                  seed=0, n_episodes=1, extra_done_steps=0, max_steps=None):
    """Play ``n_episodes`` games on the reference env and return a trace dict.

    Policies: ``random`` (discrete ints / continuous float32 in [-1.3,1.3]),
    ``random_logits`` (discrete, float32 (1,4) arrays -> argmax path,
    battle_env.py:327-328), ``instinct`` (reference instinct.Team),
    ``instinct_ff`` (instinct with a2 forced to +1: full-bullet-pool stress),
    ``noop_shoot`` (discrete action 1 / continuous [1,0,1] every tick),
    ``model`` (red only, 2v2 discrete: the shipped ``models/completed_model`` actors,
    maddpg/networks.py:54-85, evaluated as maddpg/agent.py:25-33 does, noise off).
    """
    be, sprites, iagent, iteam = _load()
    rewards = dict(rewards or {})
    N, P = n_agents, 2 * n_agents
    A = 3 if continuous else 4
    L = 3 * N + 2

    tap_r = _TapRandom(seed * 7919 + 13)
    tap_np = _TapNpRandom(seed * 104729 + 7)
    pol_rng = np.random.RandomState(seed * 31337 + 101)

    class TapBullet(sprites.Bullet):
        def __init__(self, x, y, angle, speed, fcolor, oteam, agent_id, shot_dist):
            tap_r.cur_shooter = agent_id
            super().__init__(x, y, angle, speed, fcolor, oteam, agent_id, shot_dist)

    old = (sprites.random, iagent.np, be.Bullet)
    sprites.random, iagent.np, be.Bullet = tap_r, _NpProxy(tap_np), TapBullet
    try:
        with _cwd(REF_ROOT):
            return _play(be, iteam, tap_r, tap_np, pol_rng, N, P, A, L, continuous, rewards,
                         red, blue, seed, n_episodes, extra_done_steps, max_steps)
    finally:
        sprites.random, iagent.np, be.Bullet = old


def _play(be, iteam, tap_r, tap_np, pol_rng, N, P, A, L, continuous, rewards, red, blue,
          seed, n_episodes, extra_done_steps, max_steps):
    env = be.parallel_env(n_agents=N, continuous_actions=continuous, **rewards)
    tap_r.ints.clear()  # constructor spawns are overwritten by the first reset()
    ids = env.possible_agents
    idx = {a: i for i, a in enumerate(ids)}
    teams = {"red": (env.possible_red, env.possible_blue), "blue": (env.possible_blue, env.possible_red)}
    pol = {"red": red, "blue": blue}
    inst = {c: iteam.Team(teams[c][0], teams[c][1], env) for c in ("red", "blue")
            if pol[c] in ("instinct", "instinct_ff")}
    logits = any(p in ("random_logits", "model") for p in pol.values())
    if logits and continuous:
        raise ValueError("logit policies are discrete-mode only")
    actors = _load_actors(teams["red"][0], L) if pol["red"] == "model" else None

    T = {k: [] for k in ("act", "draws", "obs", "rew", "done", "env_done", "winner", "px", "py",
                          "pdir", "php", "bhp", "bullets")}
    reset_draws, obs0, bpos, ep_start = [], [], [], [0]

    def snapshot():
        px = np.zeros(P, np.int16); py = np.zeros(P, np.int16)
        pdir = np.zeros(P, np.float64); php = np.zeros(P, np.int8)
        for a in ids:
            c = env.team_map[a]
            pl = env.team[c]["planes"].get(a)
            if pl is not None:
                px[idx[a]], py[idx[a]] = pl.rect.center
                pdir[idx[a]] = pl.direction
                php[idx[a]] = pl.hp
        bhp = np.array([env.team["red"]["base"].hp, env.team["blue"]["base"].hp], np.int16)
        bul = -np.ones((P, BULLET_SLOTS, 2), np.int16)
        for b in env.bullets:
            age = int(round(b.dist_travelled / 45.0))
            assert 1 <= age <= BULLET_SLOTS - 1, age
            s = bul[idx[b.agent_id], age - 1]
            assert s[0] == -1
            s[0], s[1] = b.rect.center
        return px, py, pdir, php, bhp, bul

    for _ep in range(n_episodes):
        tap_r.ints.clear()
        obs = env.reset()
        assert len(tap_r.ints) == 4 + 6 * N
        reset_draws.append(np.array(tap_r.ints, np.int32))
        obs0.append(np.stack([obs[a] for a in ids]).astype(np.float32))
        bpos.append(np.array([env.team["red"]["base"].rect.center,
                              env.team["blue"]["base"].rect.center], np.int16))
        after_done = 0
        while True:
            if env.env_done:
                if after_done >= extra_done_steps:
                    break
                after_done += 1
            if max_steps is not None and len(T["act"]) - ep_start[-1] >= max_steps:
                break
            # ---- policies -------------------------------------------------
            actions, drw = {}, np.full((P, 5), np.nan)
            for c in ("red", "blue"):
                mine = teams[c][0]
                p = pol[c]
                if p in ("instinct", "instinct_ff"):
                    tap_np.order = list(mine) if continuous else []
                    tap_np.records = {}
                    # numpy-1.23.1 promotion: widen float32 obs exactly (see module docstring)
                    o64 = {a: obs[a].astype(np.float64) for a in mine}
                    acts = inst[c].choose_actions(o64)
                    for a in mine:
                        v = acts[a]
                        if continuous:
                            v = np.array(v, np.float64)
                            if p == "instinct_ff":
                                v[2] = 1.0
                            drw[idx[a], 0:4] = tap_np.records[a]
                        actions[a] = v
                elif p == "random":
                    for a in mine:
                        if continuous:
                            v32 = pol_rng.uniform(-1.3, 1.3, 3).astype(np.float32)
                            actions[a] = v32.astype(np.float64)  # exact widening (numpy-1.23.1 meaning)
                        else:
                            actions[a] = int(pol_rng.randint(0, 4))
                elif p == "random_logits":
                    for a in mine:
                        actions[a] = pol_rng.uniform(-1, 1, (1, 4)).astype(np.float32)
                elif p == "model":
                    import torch
                    for a in mine:
                        with torch.no_grad():
                            st = torch.tensor(np.array([obs[a]]), dtype=torch.float)
                            actions[a] = actors[a](st).clamp(-1, 1).numpy()
                elif p == "noop_shoot":
                    for a in mine:
                        actions[a] = np.array([1.0, 0.0, 1.0]) if continuous else 1
                else:
                    raise ValueError(p)
            if continuous:
                act_rec = np.stack([np.asarray(actions[a], np.float64) for a in ids])
            elif logits:
                # int actions of a non-logit team are recorded as one-hot rows (same argmax)
                act_rec = np.stack([np.asarray(actions[a], np.float32).reshape(4) if isinstance(actions[a], np.ndarray)
                                    else np.eye(4, dtype=np.float32)[actions[a]] for a in ids])
            else:
                act_rec = np.array([actions[a] for a in ids], np.int32)
            # ---- step -----------------------------------------------------
            tap_r.spread.clear()
            obs, rew, dones, _ = env.step(actions)
            for a, u in tap_r.spread:
                assert np.isnan(drw[idx[a], 4])
                drw[idx[a], 4] = u
            px, py, pdir, php, bhp, bul = snapshot()
            T["act"].append(act_rec); T["draws"].append(drw)
            T["obs"].append(np.stack([obs[a] for a in ids]).astype(np.float32))
            T["rew"].append(np.array([float(rew[a]) for a in ids], np.float64))
            T["done"].append(np.array([bool(dones[a]) for a in ids]))
            T["env_done"].append(bool(env.env_done)); T["winner"].append(WINNER_CODE[env.winner])
            T["px"].append(px); T["py"].append(py); T["pdir"].append(pdir); T["php"].append(php)
            T["bhp"].append(bhp); T["bullets"].append(bul)
        ep_start.append(len(T["act"]))

    out = {k: np.stack(v) for k, v in T.items()}
    out["env_done"] = out["env_done"].astype(np.bool_)
    out["winner"] = out["winner"].astype(np.int8)
    out.update(
        reset_draws=np.stack(reset_draws), obs0=np.stack(obs0), base_pos=np.stack(bpos),
        ep_start=np.array(ep_start, np.int64),
        n_agents=np.int32(N), continuous=np.bool_(continuous), logits=np.bool_(logits),
        rewards=np.array([getattr(env, k) for k in ("hit_base_reward", "hit_plane_reward",
                                                     "miss_punishment", "die_punishment",
                                                     "lose_punishment")], np.float64),
        counters=np.array([env.team["red"]["wins"], env.team["blue"]["wins"], env.ties,
                           env.total_games], np.int64),
        red_policy=np.str_(red), blue_policy=np.str_(blue), seed=np.int64(seed),
    )
    return out
