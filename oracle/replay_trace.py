"""TEST INFRASTRUCTURE ONLY -- replays a reference trace into an env and compares.

A trace (dict of numpy arrays, see oracle/ref_harness.py:run_reference) holds, per
episode, the spawn draws and, per step, the actions and unit draws the reference
consumed plus everything it produced.  ``replay`` feeds the same inputs to any env
exposing the batched surface (OracleEnv in oracle/port.py, or the CUDA BattleEnv
through tests/ adapters): all episodes of the trace run side by side as one batch,
one env per episode, so finished envs keep being stepped (the reference's
"env_done: return zeros" branch, battle_env.py:303-306) while longer ones go on.

exact=True  : every compared array must be bit-identical (CPU port vs reference).
exact=False : discrete quantities bit-identical; float32 obs / rewards within
              rtol 1e-5 (north_star tolerance), fp64 direction within 1e-9.
"""
from __future__ import annotations

import numpy as np

RTOL = 1e-5


class Mismatch(AssertionError):
    pass


def _cmp(name, got, want, exact, step, mask=None, atol=0.0):
    got = np.asarray(got); want = np.asarray(want)
    if mask is not None:
        got = got[mask]; want = want[mask]
    if exact or got.dtype.kind in "biu":
        ok = np.array_equal(got, want)
    else:
        ok = np.allclose(got.astype(np.float64), want.astype(np.float64), rtol=RTOL, atol=atol)
    if not ok:
        bad = np.argwhere(~np.isclose(got.astype(np.float64), want.astype(np.float64), rtol=0, atol=0))
        raise Mismatch(f"{name} differs at step {step}: first bad index {bad[:4].tolist()}, "
                       f"got {got[tuple(bad[0])]!r} want {want[tuple(bad[0])]!r}")


def replay(env, tr, exact=False, check_instinct=True, counts=None):
    """``env`` must have E == number of episodes in ``tr`` and matching config."""
    N = int(tr["n_agents"]); P = 2 * N
    cont = bool(tr["continuous"]); logits = bool(tr["logits"])
    ep = tr["ep_start"]; n_ep = len(ep) - 1
    lens = np.diff(ep)
    assert env.E == n_ep, (env.E, n_ep)
    pol = {0: str(tr["red_policy"]), 1: str(tr["blue_policy"])}
    counts = counts if counts is not None else {}

    obs = env.reset(None, tr["reset_draws"])
    _cmp("reset obs", obs, tr["obs0"], exact, -1, atol=1e-7)
    st = env.get_state()
    _cmp("base_pos", st["base_pos"], tr["base_pos"], True, -1)
    prev_obs = obs

    for t in range(int(lens.max())):
        live = lens > t                       # envs whose trace still has a step t
        gi = np.where(live, ep[:-1] + t, 0)   # global step index per env
        if cont:
            act = np.where(live[:, None, None], tr["act"][gi], 0.0).astype(np.float64)
        elif logits:
            act = np.where(live[:, None, None], tr["act"][gi], 0.0).astype(np.float32)
        else:
            act = np.where(live[:, None], tr["act"][gi], 0).astype(np.int32)
        draws = np.where(live[:, None, None], tr["draws"][gi], np.nan)

        if check_instinct:
            for team in (0, 1):
                if pol[team] not in ("instinct", "instinct_ff"):
                    continue
                got = env.instinct(team, prev_obs, draws)
                want = act[:, team * N:(team + 1) * N]
                if logits:
                    want = want.argmax(-1).astype(np.int32)
                if cont and pol[team] == "instinct_ff":
                    got = got.copy(); got[..., 2] = 1.0
                if cont:
                    # the shoot decision (sign of a2) is a discrete event: exact
                    _cmp(f"instinct team{team} shoot", got[..., 2] > 0, want[..., 2] > 0, True, t, live)
                    _cmp(f"instinct team{team} action", got, want, exact, t, live, atol=1e-12)
                    counts["instinct_exact"] = counts.get("instinct_exact", 0) + int(np.sum(got[live] == want[live]))
                    counts["instinct_total"] = counts.get("instinct_total", 0) + int(got[live].size)
                else:
                    _cmp(f"instinct team{team} action", got, want, True, t, live)

        obs, rew, done, env_done, winner = env.step(act, draws)
        st = env.get_state()
        _cmp("obs", obs, tr["obs"][gi], exact, t, live, atol=1e-7)
        counts["obs_exact"] = counts.get("obs_exact", 0) + int(np.sum(obs[live] == tr["obs"][gi][live]))
        counts["obs_total"] = counts.get("obs_total", 0) + int(obs[live].size)
        _cmp("rew", rew, tr["rew"][gi], exact, t, live, atol=1e-9)
        _cmp("done", done, tr["done"][gi], True, t, live)
        _cmp("env_done", env_done, tr["env_done"][gi], True, t, live)
        _cmp("winner", winner, tr["winner"][gi], True, t, live)
        _cmp("php", np.maximum(st["php"], 0), tr["php"][gi], True, t, live)
        alive = live[:, None] & (tr["php"][gi] > 0)
        _cmp("px", st["px"], tr["px"][gi], True, t, alive)
        _cmp("py", st["py"], tr["py"][gi], True, t, alive)
        if exact:
            _cmp("pdir", st["pdir"], tr["pdir"][gi], True, t, alive)
        elif not np.allclose(st["pdir"][alive], tr["pdir"][gi][alive], rtol=0, atol=1e-9):
            raise Mismatch(f"pdir differs at step {t}")
        _cmp("bhp", st["bhp"], tr["bhp"][gi], True, t, live)
        _cmp("bullets", st["bullets"], tr["bullets"][gi], True, t, live)
        prev_obs = obs
    counts["steps"] = counts.get("steps", 0) + int(lens.sum())
    return counts
