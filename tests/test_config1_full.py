"""GPU: SURVEY 8(d) config 1 at its full size -- 1v1, 300 full episodes x seeds 1..8, discrete and continuous
(actions uniform(-1.3, 1.3): the clip of battle_env.py:295-297), ~36 k ticks per run, 576 k reference ticks in all.

tests/golden/config1_digest.npz (oracle/gen_config1_digest.py, from the unmodified reference) holds per-episode
digests only; the inputs are regenerated here exactly as the reference consumed them:
  * actions: numpy RandomState(seed * 31337 + 101), one draw per plane and step call (oracle/ref_harness.py);
  * spawns and bullet spreads: ONE Python random.Random(seed * 7919 + 13) stream, consumed in the reference's
    order -- the constructor's spawns, then per game 4 + 6N randint (sprites.py:246-251 bases, :82-91 planes) and
    one random() per bullet in agent order (sprites.py:314) -- which this test can only follow if the CUDA env's
    alive flags and step outcomes are right at every tick; the regenerated spawn draws are checked against the
    recorded ones for all 300 games, so a single mis-ordered draw anywhere fails the run.
Pass 1 drives the CUDA env one env / one tick at a time through bs_reset_host / bs_step_host (the plugin calls
that stand where parallel_env.reset / step stand); pass 2 replays the same 300 games as 300 envs of one batch.
Bar: observations, float32-rounded rewards, dones, winners, lengths and final plane / base / bullet state
bit-equal (CRC-32); float64 reward sums within 1e-5.
"""
import os
import random
import zlib

import numpy as np
import pytest

from conftest import GOLDEN
from port import TRAINED_REWARDS

DEFAULT_REWARDS = (100.0, 10.0, -1.0, -5.0, -20.0)   # battle_env.py:73
TIE_TICK_1V1 = 121


def crc(*arrays):
    c = 0
    for a in arrays:
        c = zlib.crc32(np.ascontiguousarray(a).tobytes(), c)
    return c


def spawn_draws(r):
    """4 + 6N randint calls of one reset, N = 1, in the reference's order and ranges."""
    return [r.randint(62, 379), r.randint(62, 738), r.randint(758, 1138), r.randint(62, 738),    # bases, sprites.py:246-251
            r.randint(50, 383), r.randint(48, 752), r.randint(270, 450),                          # red plane, :82-85
            r.randint(766, 1150), r.randint(48, 752), r.randint(90, 270)]                         # blue plane, :88-91


def state_crc(st, e):
    alive = st["php"][e] > 0
    return crc(np.where(alive, st["px"][e], 0).astype(np.int16), np.where(alive, st["py"][e], 0).astype(np.int16),
               np.maximum(st["php"][e], 0).astype(np.int8), st["bhp"][e].astype(np.int16), st["bullets"][e].astype(np.int16))


def run_case(cont, seed, dig, make_env):
    """make_env(n_envs, host_mode) -> an env with the OracleEnv / CudaEnv call surface"""
    tag = f"{'cont' if cont else 'disc'}_{seed}"
    want = {k: dig[f"{tag}_{k}"] for k in ("steps", "winner", "rew_sum", "crc_obs", "crc_rew", "crc_done", "crc_state", "reset_draws")}
    n_ep = len(want["steps"])
    rewards = TRAINED_REWARDS if cont else DEFAULT_REWARDS
    r = random.Random(seed * 7919 + 13)
    pol = np.random.RandomState(seed * 31337 + 101)
    spawn_draws(r)                                     # parallel_env.__init__ spawns once (battle_env.py:96-118)
    env = make_env(cont, rewards, 1, True)
    games, n_spread = [], 0
    for ep in range(n_ep):
        sd = np.array(spawn_draws(r), np.int32)
        assert np.array_equal(sd, want["reset_draws"][ep]), (tag, ep, "random stream out of step with the reference")
        obs = env.reset(None, sd[None])
        obs_l, rew_l, done_l, act_l, spr_l = [obs[0].copy()], [], [], [], []
        alive, k, env_done, winner = [True, True], 0, False, 0
        while not env_done:
            k += 1
            if cont:
                act = np.stack([pol.uniform(-1.3, 1.3, 3).astype(np.float32) for _ in range(2)])
                shoots = [act[i, 2] > 0 for i in range(2)]
            else:
                act = np.array([int(pol.randint(0, 4)) for _ in range(2)], np.int32)
                shoots = [act[i] == 1 for i in range(2)]
            drw = np.full((1, 2, 5), np.nan)
            if k < TIE_TICK_1V1 and any(alive):        # the tie calls process no action (battle_env.py:309-323)
                for i in range(2):
                    if alive[i] and shoots[i]:
                        drw[0, i, 4] = r.random(); n_spread += 1
            o, rw, dn, ed, win = env.step(act[None], drw)
            obs_l.append(o[0].copy()); rew_l.append(rw[0].copy()); done_l.append((dn[0].copy(), ed[0]))
            act_l.append(act); spr_l.append(drw[0, :, 4].copy())
            env_done, winner = bool(ed[0]), int(win[0])
            alive = [not bool(dn[0, i]) for i in range(2)]
        st = env.get_state()
        rews = np.stack(rew_l)
        got = dict(steps=k, winner=winner, crc_obs=crc(np.stack(obs_l).astype(np.float32)), crc_rew=crc(rews.astype(np.float32)),
                   crc_done=crc(np.stack([d for d, _ in done_l]).astype(np.uint8), np.array([e for _, e in done_l], np.uint8)),
                   crc_state=state_crc(st, 0))
        for key, v in got.items():
            assert v == want[key][ep], (tag, ep, key, v, want[key][ep])
        assert np.allclose(rews.astype(np.float64).sum(axis=0), want["rew_sum"][ep], rtol=1e-5, atol=1e-4), (tag, ep)
        games.append((sd, np.stack(act_l), np.stack(spr_l)))
    assert n_spread == int(dig[f"{tag}_n_spread"])
    c = env.stats()
    assert c[:4].tolist() == dig[f"{tag}_counters"].tolist(), (tag, c[:4], dig[f"{tag}_counters"])

    # ---- pass 2: the same games as one batch of n_ep envs ---------------------------------------------------
    E = n_ep
    benv = make_env(cont, rewards, E, False)
    obs = benv.reset(None, np.stack([g[0] for g in games]))
    T = int(want["steps"].max())
    A = games[0][1].shape[1:]
    acts = np.zeros((T, E) + A, games[0][1].dtype)
    drws = np.full((T, E, 2, 5), np.nan)
    for e, (_, a, s) in enumerate(games):
        acts[:len(a), e] = a
        drws[:len(a), e, :, 4] = s
    obs_t, rew_t, done_t, ed_t = [obs], [], [], []
    for t in range(T):
        o, rw, dn, ed, win = benv.step(acts[t], drws[t])
        obs_t.append(o); rew_t.append(rw); done_t.append(dn); ed_t.append(ed)
    obs_t, rew_t, done_t, ed_t = np.stack(obs_t), np.stack(rew_t), np.stack(done_t), np.stack(ed_t)
    st = benv.get_state()
    for e in range(E):
        n = int(want["steps"][e])
        assert crc(obs_t[:n + 1, e].astype(np.float32)) == want["crc_obs"][e], (tag, e, "obs (batched)")
        assert crc(rew_t[:n, e].astype(np.float32)) == want["crc_rew"][e], (tag, e, "rew (batched)")
        assert crc(done_t[:n, e].astype(np.uint8), ed_t[:n, e].astype(np.uint8)) == want["crc_done"][e], (tag, e)
        assert int(win[e]) == want["winner"][e] and bool(ed_t[n - 1, e]) and (n == 1 or not ed_t[n - 2, e])
        assert state_crc(st, e) == want["crc_state"][e], (tag, e, "final state (batched)")
    assert benv.stats()[:4].tolist() == dig[f"{tag}_counters"].tolist()


def _digest():
    with np.load(os.path.join(GOLDEN, "config1_digest.npz")) as z:
        return {k: z[k] for k in z.files}


@pytest.mark.gpu
@pytest.mark.parametrize("cont", [False, True])
@pytest.mark.parametrize("seed", range(1, 9))
def test_config1_300_episodes(cuda_device, cont, seed):
    from cuda_adapter import CudaEnv
    run_case(cont, seed, _digest(), lambda c, rw, E, host: CudaEnv(1, c, rw, E, seed=0, mode="host" if host else "device"))


@pytest.mark.parametrize("cont,seed", [(False, 3), (True, 6)])
def test_config1_300_episodes_cpu_oracle(cont, seed):
    """The same check on the CPU restatement (oracle/battle_oracle.c): pins the oracle itself at config 1's size."""
    from port import OracleEnv
    run_case(cont, seed, _digest(), lambda c, rw, E, host: OracleEnv(1, c, rw, E, seed=0))
