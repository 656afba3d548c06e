"""GPU: the episode log against the REFERENCE's own episodes.

main.py:193-221 accumulates red_score / blue_score over the per-step reward dicts of one game and
appends round(score, 3) to score_dict; env.winner names the outcome.  Every golden trace holds the
reference's per-step rewards and episode boundaries, so the expected scores are a reduceat away.  Each
trace is replayed into the CUDA env (one env per episode, injected draws) with an EpisodeLog attached;
the log must hold exactly one entry per game with those scores and the reference's winner."""
import os

import numpy as np
import pytest

from conftest import golden_traces, load_trace
from replay_trace import replay

pytestmark = pytest.mark.gpu
CODE_OF_WINNER = {1: 1, 2: 2, 3: 3}        # bs winner code -> log code (6 = both bases fell in one tick, ends 'blue')


@pytest.mark.parametrize("path", golden_traces(), ids=lambda p: os.path.basename(p)[6:-4])
def test_logged_scores_equal_reference_score_dict(path, cuda_device):
    from battlespace import EpisodeLog
    from cuda_adapter import CudaEnv
    tr = load_trace(path)
    N = int(tr["n_agents"])
    ep = tr["ep_start"]
    n_ep = len(ep) - 1
    env = CudaEnv(N, bool(tr["continuous"]), tr["rewards"], n_ep)
    log = EpisodeLog(env.env, 1024)
    replay(env, tr, exact=False, check_instinct=False)
    games = log.new_games()
    # the reference's score_dict entries (main.py:193-203,219-221): python float sums in step / agent order
    want = []
    for e in range(n_ep):
        rew = tr["rew"][ep[e]:ep[e + 1]].astype(np.float64)          # [steps, 2N]
        red = blue = 0.0
        for t in range(rew.shape[0]):
            for i in range(N):
                red += rew[t, i]
            for i in range(N, 2 * N):
                blue += rew[t, i]
        finished = bool(tr["env_done"][ep[e + 1] - 1])
        want.append((round(red, 3), round(blue, 3), int(tr["winner"][ep[e + 1] - 1]), finished))
    fin = [w for w in want if w[3]]
    assert len(games) == len(fin), (len(games), len(fin))
    # the log is appended in completion order: match by (scores, code), every entry used once
    got = sorted((round(float(g[0]), 3), round(float(g[1]), 3), int(g[2])) for g in games)
    exp = sorted((w[0], w[1], CODE_OF_WINNER[w[2]]) for w in fin)
    for g, w in zip(got, exp):
        if g[2] == 6:                       # double win: the reference ends as 'blue'
            assert w[2] == 2
        else:
            assert g[2] == w[2], (g, w)
        assert abs(g[0] - w[0]) <= 1.5e-3 and abs(g[1] - w[1]) <= 1.5e-3, (g, w)   # float32 log entry, rounded to 3 places
    sd = EpisodeLog.score_dict(games)
    assert len(sd["red"]) == len(fin) and all(isinstance(x, float) for x in sd["red"])
