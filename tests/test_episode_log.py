"""GPU: the device episode log (main.py:218-253) against the env's own counters and against
returns accumulated on the host from per-tick rewards."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
CF = dict(hit_base_reward=1.0, hit_plane_reward=0.9, miss_punishment=-0.02, die_punishment=-0.03, lose_punishment=-0.05)


def test_log_matches_counters_fused_rollout(cuda_device):
    from battlespace import EpisodeLog, parallel_env
    env = parallel_env(n_agents=3, continuous_actions=True, n_envs=4096, seed=5, **CF)
    log = EpisodeLog(env, 1 << 18)
    for _ in range(12):
        env.rollout(32, "instinct_ff", "instinct")
    st = env.stats()
    g = log.new_games()
    code = g[:, 2].astype(np.int64)
    assert len(g) > 5000 and set(np.unique(code)) <= {1, 2, 3, 6}
    both = int((code == 6).sum())
    assert int((code == 1).sum()) + both == st[0] and int((code == 2).sum()) + both == st[1]
    assert int((code == 3).sum()) == st[2] and len(g) + both == st[3]
    # sums of episode returns: the counters hold sum(rint(ret * 2^20)), the log float32(ret)
    assert abs(g[:, 0].astype(np.float64).sum() - st[4] / 2 ** 20) < 1e-3 * max(1.0, abs(st[4] / 2 ** 20))
    assert abs(g[:, 1].astype(np.float64).sum() - st[5] / 2 ** 20) < 1e-3 * max(1.0, abs(st[5] / 2 ** 20))
    assert len(log.new_games()) == 0
    rep = EpisodeLog.report(g, st[3], st[6])
    assert abs(rep["red_winrate"] + rep["blue_winrate"] + rep["tie_rate"] - 1) < 2e-3
    sd = EpisodeLog.score_dict(g[:10])
    assert list(sd) == ["red", "blue"] and len(sd["red"]) == 10


def test_log_scores_equal_host_accumulated_returns(cuda_device):
    """main.py:193-203: red_score / blue_score are the per-tick team reward sums of one game."""
    import torch
    from battlespace import EpisodeLog, parallel_env
    E, N = 256, 2
    env = parallel_env(n_agents=N, n_envs=E, seed=8, **CF)
    log = EpisodeLog(env, 4096)  # power of two
    env.reset_batched()
    acc = torch.zeros(E, 2, dtype=torch.float64, device=cuda_device)
    finished = []
    g = torch.Generator(device=cuda_device).manual_seed(0)
    for t in range(400):
        act = torch.randint(0, 4, (E, 2 * N), device=cuda_device, generator=g, dtype=torch.int32)
        obs, rew, done, env_done, winner = env.step_batched(act)
        r = rew.double()
        acc[:, 0] += r[:, :N].sum(1); acc[:, 1] += r[:, N:].sum(1)
        ended = env_done.bool()
        if ended.any():
            finished.append(acc[ended].cpu().numpy().copy())
            acc[ended] = 0
            env.reset_done()
    want = np.concatenate(finished)
    got = log.new_games()
    assert len(got) == len(want) > E
    key = lambda a: np.lexsort((np.round(a[:, 1], 4), np.round(a[:, 0], 4)))
    assert np.allclose(got[key(got), :2], want[key(want)], atol=1e-5)
