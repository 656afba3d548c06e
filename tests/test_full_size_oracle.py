"""GPU: the CUDA path against the CPU oracle AT THE BASELINE SIZES ("what's missing" #2).

The small-size comparisons (tests/test_cuda_vs_oracle.py) rarely reach the pool occupancies and the
same-tick multi-hit / death orderings that 65,536 envs x 256 ticks produce; here the ordered bullet loop
(battle_env.py:332-360) is compared at full size: every counter, the whole env state (planes, bases,
every bullet slot) and the replay rows the last launch left in the ring.

  config 3: 3v3 continuous, red = uniform random, blue = instinct, 65,536 envs, T = 32 x 8 launches
  config 4: 5v5 continuous, instinct_ff on both sides (forced fire: full 120-slot pools), 131,072 envs
            = one GPU's share of the 1,048,576-env job on 8 GPUs, global env offset of rank 3, T = 32 x 4
"""
import numpy as np
import pytest

from port import OracleEnv, TRAINED_REWARDS, POL_INSTINCT, POL_INSTINCT_FF, POL_RANDOM

pytestmark = pytest.mark.gpu
POLNAME = {POL_RANDOM: "random", POL_INSTINCT: "instinct", POL_INSTINCT_FF: "instinct_ff"}


def _compare(N, E, T, launches, pr, pb, ring_ticks, env_offset, seed):
    import torch
    from battlespace import DeviceReplay
    from cuda_adapter import CudaEnv
    L, A = 3 * N + 2, 3
    cap = ring_ticks * E                    # the ring keeps the last `ring_ticks` ticks of a launch (capacity % E == 0)
    c = CudaEnv(N, True, TRAINED_REWARDS, E, env_offset=env_offset, seed=seed)
    o = OracleEnv(N, True, TRAINED_REWARDS, E, env_offset=env_offset, seed=seed)
    rp = DeviceReplay(cap, 512, N, L, A, c.dev)
    orp, cursor = None, 0
    for _ in range(launches):
        c.env.rollout(T, POLNAME[pr], POLNAME[pb], 0, None, rp)
        orp = o.rollout(T, [pr, pb], 0, None, cap, cursor, orp)
        cursor += T * E
    torch.cuda.synchronize()
    assert np.array_equal(c.stats(), o.stats())                      # wins, ties, games, returns, ticks, agent-steps
    cs, os_ = c.get_state(), o.get_state()
    alive = os_["php"] > 0
    assert np.array_equal(np.maximum(cs["php"], 0), os_["php"])
    assert np.array_equal(cs["px"][alive], os_["px"][alive]) and np.array_equal(cs["py"][alive], os_["py"][alive])
    assert np.allclose(cs["pdir"][alive], os_["pdir"][alive], rtol=0, atol=1e-9)
    for k in ("bhp", "bullets", "base_pos", "tick", "env_done", "winner", "episode"):
        assert np.array_equal(cs[k], os_[k]), k
    occupancy = float((cs["bullets"][..., 0] >= 0).sum()) / E
    got = dict(state=rp.state_mem, new_state=rp.new_state_mem, action=rp.action_mem, reward=rp.rew_mem, done=rp.done_mem)
    for k, v in got.items():
        g = v.cpu().numpy().reshape(orp[k].shape)
        if k == "done":
            assert np.array_equal(g, orp[k]), k
        else:
            assert np.allclose(g, orp[k], rtol=1e-5, atol=1e-7), k   # north_star tolerance ...
            assert (g == orp[k]).mean() > 0.999999, (k, (g == orp[k]).mean())   # ... and in fact bit-equal
    return occupancy, c.stats()


def test_config3_full_size_vs_oracle(cuda_device):
    occ, st = _compare(3, 65536, 32, 8, POL_RANDOM, POL_INSTINCT, ring_ticks=32, env_offset=0, seed=31)
    assert st[3] > 65536 and st[6] == 8 * 32 * 65536                 # every env finished a game on average
    assert occ > 5                                                    # live bullets per env at the end


def test_config4_per_gpu_share_vs_oracle(cuda_device):
    occ, st = _compare(5, 131072, 32, 4, POL_INSTINCT_FF, POL_INSTINCT_FF, ring_ticks=4, env_offset=3 * 131072, seed=9)
    assert st[6] == 4 * 32 * 131072 and occ > 40                      # forced fire: the pools are busy
