"""GPU: the two thread mappings of the fused rollout give bit-identical results.

bs_rollout_fused runs rollout_kernel (lane-per-plane, csrc/battlespace.cu); a handle created under
BS_ROLLOUT_V2=1 runs rollout2_kernel (team-per-thread, csrc/rollout2.cuh) for team sizes 1..5.  Both restate parallel_env.step (battle_env.py:281-381) + the instinct team
(instinct/agent.py:10-62) + store_transition (maddpg/buffer.py:25-47) with the same device functions, and
both are compared with the oracle elsewhere; here they are compared with each other, bit for bit, on every
counter, the whole state, every bullet and every replay row -- including launches interleaved with the
single-tick kernels, which forces the bullet layout conversions (pools <-> per-team lists) both ways.
"""
import os

import numpy as np
import pytest

from port import TRAINED_REWARDS

pytestmark = pytest.mark.gpu


def _mk(N, cont, E, seed, v1, env_offset=0):
    from cuda_adapter import CudaEnv
    old = os.environ.pop("BS_ROLLOUT_V2", None)
    if not v1:
        os.environ["BS_ROLLOUT_V2"] = "1"
    try:
        return CudaEnv(N, cont, TRAINED_REWARDS, E, env_offset=env_offset, seed=seed)
    finally:
        os.environ.pop("BS_ROLLOUT_V2", None)
        if old is not None:
            os.environ["BS_ROLLOUT_V2"] = old


def _same_state(a, b):
    sa, sb = a.env.get_state(), b.env.get_state()
    for k in sa:
        if k == "plane_hp":
            assert np.array_equal(np.maximum(sa[k], 0), np.maximum(sb[k], 0)), k
        else:
            assert np.array_equal(sa[k], sb[k]), k
    assert np.array_equal(a.stats(), b.stats())


def _same_rows(ra, rb):
    import torch
    for k in ("state_mem", "new_state_mem", "action_mem", "rew_mem", "done_mem"):
        assert torch.equal(getattr(ra, k), getattr(rb, k)), k


CASES = [  # N, continuous, E, red, blue, learner
    (3, True, 4099, "random", "instinct", 0),
    (3, True, 2048, "instinct", "random", 1),
    (3, False, 1500, "random", "instinct", 0),
    (1, True, 777, "random", "instinct", 0),
    (2, False, 1024, "instinct", "instinct", 0),
    (4, True, 513, "instinct_ff", "instinct", 1),
    (5, True, 1000, "instinct_ff", "instinct_ff", 0),
    (5, False, 320, "random", "instinct", 0),
]


@pytest.mark.parametrize("N,cont,E,red,blue,learner", CASES)
def test_team_per_thread_equals_lane_per_plane(cuda_device, N, cont, E, red, blue, learner):
    import torch
    from battlespace import DeviceReplay
    L, A, T = 3 * N + 2, 3 if cont else 4, 24
    a, b = _mk(N, cont, E, 5, v1=False), _mk(N, cont, E, 5, v1=True)
    cap = 3 * T * E + 7                                   # odd capacity, cursor not a multiple of 4: unaligned and wrapping rows
    ra, rb = DeviceReplay(cap, 64, N, L, A, a.dev), DeviceReplay(cap, 64, N, L, A, b.dev)
    for launch in range(8):
        a.env.rollout(T, red, blue, learner, None, ra)
        b.env.rollout(T, red, blue, learner, None, rb)
        torch.cuda.synchronize()
        assert np.array_equal(a.stats(), b.stats()), launch
        _same_rows(ra, rb)
        if launch % 3 == 2:
            _same_state(a, b)                             # lists -> pools on `a`, then back on the next launch
    assert a.stats()[3] > 0


def test_external_actions_and_aligned_ring(cuda_device):
    import torch
    from battlespace import DeviceReplay
    N, E, T = 3, 4096, 16
    a, b = _mk(N, True, E, 11, v1=False), _mk(N, True, E, 11, v1=True)
    ra, rb = DeviceReplay(4 * T * E, 64, N, 11, 3, a.dev), DeviceReplay(4 * T * E, 64, N, 11, 3, b.dev)   # aligned: the 128-bit path
    g = torch.Generator(device="cpu").manual_seed(3)
    for launch in range(6):
        ext = (torch.rand((T, E, N, 3), generator=g) * 2.6 - 1.3).to(a.dev)
        a.env.rollout(T, "external", "instinct", 0, ext, ra)
        b.env.rollout(T, "external", "instinct", 0, ext, rb)
        torch.cuda.synchronize()
        _same_rows(ra, rb)
    _same_state(a, b)


def test_interleaved_with_single_tick_kernels(cuda_device):
    """rollout -> step -> masked reset -> rollout: the bullets change layout each time."""
    import torch
    N, E = 3, 1111
    a, b = _mk(N, True, E, 21, v1=False), _mk(N, True, E, 21, v1=True)
    rng = np.random.default_rng(0)
    for rnd in range(4):
        for env in (a, b):
            env.env.rollout(40, "random", "instinct_ff")
        acts = rng.uniform(-1, 1, (E, 2 * N, 3)).astype(np.float32)
        oa, ob = a.step(acts), b.step(acts)
        for x, y in zip(oa, ob):
            assert np.array_equal(x, y)
        mask = (rng.random(E) < 0.3).astype(np.uint8)
        assert np.array_equal(a.reset(mask), b.reset(mask))
        _same_state(a, b)


def test_episode_log_same_games(cuda_device):
    import torch
    from battlespace import EpisodeLog
    N, E = 2, 512
    logs = []
    for v1 in (False, True):
        env = _mk(N, False, E, 3, v1=v1)
        log = EpisodeLog(env.env, 1 << 14)
        for _ in range(6):
            env.env.rollout(64, "random", "instinct")
        torch.cuda.synchronize()
        rec = log.new_games()
        logs.append(np.asarray(sorted(map(tuple, np.asarray(rec, np.float64).reshape(-1, 3).tolist()))))
    assert logs[0].shape == logs[1].shape and logs[0].shape[0] > 100
    assert np.array_equal(logs[0], logs[1])
