"""CPU: the C oracle (oracle/battle_oracle.c) must reproduce the reference traces bit for bit.

The fixtures under tests/golden/ were produced by the UNMODIFIED reference source
(oracle/gen_golden.py -> oracle/ref_harness.py); this is what pins the oracle.
"""
import os

import numpy as np
import pytest

from conftest import golden_traces, load_trace
from port import OracleEnv, tie_tick
from replay_trace import replay


@pytest.mark.parametrize("path", golden_traces(), ids=lambda p: os.path.basename(p)[6:-4])
def test_port_matches_reference_trace(path):
    tr = load_trace(path)
    env = OracleEnv(int(tr["n_agents"]), bool(tr["continuous"]), tr["rewards"], n_envs=len(tr["ep_start"]) - 1)
    counts = replay(env, tr, exact=True)
    assert counts["obs_exact"] == counts["obs_total"]
    assert (env.stats()[:4] == tr["counters"]).all()


def test_fixture_inventory():
    names = [os.path.basename(p) for p in golden_traces()]
    assert len(names) >= 15
    for n in (1, 2, 3, 5):
        assert any(f"trace_{n}v{n}_" in x for x in names)


def test_tie_tick_table():
    # battle_env.py:316-319: total_time is an accumulated double, so the table is irregular
    assert [tie_tick(n) for n in range(1, 9)] == [121, 141, 161, 181, 200, 220, 240, 260]


def test_known_answers_from_traces():
    """Derived known answers: clamp box, spawn ranges, bullet lifetime, hp bounds."""
    for path in golden_traces():
        tr = load_trace(path)
        N = int(tr["n_agents"])
        alive = tr["php"] > 0
        assert tr["px"][alive].min() >= 25 and tr["px"][alive].max() <= 1175
        assert tr["py"][alive].min() >= 24 and tr["py"][alive].max() <= 776
        d = tr["reset_draws"]
        assert (d[:, 0] >= 62).all() and (d[:, 0] <= 379).all() and (d[:, 2] >= 758).all() and (d[:, 2] <= 1138).all()
        red = d[:, 4:4 + 3 * N].reshape(-1, N, 3); blue = d[:, 4 + 3 * N:].reshape(-1, N, 3)
        assert red[..., 0].min() >= 50 and red[..., 0].max() <= 383 and blue[..., 0].min() >= 766 and blue[..., 0].max() <= 1150
        assert red[..., 2].min() >= 270 and red[..., 2].max() <= 450 and blue[..., 2].min() >= 90 and blue[..., 2].max() <= 270
        b = tr["bullets"]
        assert (b[:, :, 11] == -1).all()          # no bullet survives its 12th update
        assert tr["php"].max() <= 4 and tr["bhp"].max() <= 5 * N
        live = b[..., 0] >= 0
        assert b[..., 0][live].max() <= 1200 and b[..., 1][live].max() <= 800 and b[..., 1][live].min() >= 0
