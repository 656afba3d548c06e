"""GPU: edge cases of the batched path -- env counts that do not fill a warp (E = 1, E not a
multiple of the K envs a warp carries, a ragged last block), every team size 1..8 through the fused
rollout, and the C ABI's argument errors on a live handle (the reference has no error convention
beyond KeyError; the library answers bad calls with BS_E_INVALID and a message)."""
import ctypes as C

import numpy as np
import pytest

from port import OracleEnv, TRAINED_REWARDS, POL_INSTINCT, POL_INSTINCT_FF, POL_RANDOM

pytestmark = pytest.mark.gpu
POLNAME = {POL_RANDOM: "random", POL_INSTINCT: "instinct", POL_INSTINCT_FF: "instinct_ff"}


def _same_state(cs, os_):
    alive = os_["php"] > 0
    assert np.array_equal(np.maximum(cs["php"], 0), os_["php"])
    assert np.array_equal(cs["px"][alive], os_["px"][alive]) and np.array_equal(cs["py"][alive], os_["py"][alive])
    assert np.allclose(cs["pdir"][alive], os_["pdir"][alive], rtol=0, atol=1e-9)
    for k in ("bhp", "bullets", "tick", "env_done", "winner", "episode", "base_pos"):
        assert np.array_equal(cs[k], os_[k]), k


# K = 32 // (2N) envs per warp: 16, 8, 5, 4, 3, 2, 2, 2 -- E is chosen below, at and just past a warp / block
@pytest.mark.parametrize("N,cont,E", [(1, True, 1), (1, False, 17), (2, True, 9), (3, True, 1), (3, False, 7), (3, True, 21),
                                      (4, False, 5), (5, True, 13), (6, True, 3), (6, False, 9), (7, True, 5), (7, False, 1),
                                      (8, True, 1), (8, False, 11)])
def test_ragged_env_counts_rollout_and_step(N, cont, E, cuda_device):
    import torch
    from battlespace import DeviceReplay
    from cuda_adapter import CudaEnv
    L, A, T = 3 * N + 2, 3 if cont else 4, 24
    c = CudaEnv(N, cont, TRAINED_REWARDS, E, seed=5, env_offset=3)
    o = OracleEnv(N, cont, TRAINED_REWARDS, E, seed=5, env_offset=3)
    cap = 5 * T * E + 3
    rp = DeviceReplay(cap, 8, N, L, A, c.dev)
    orp, cursor = None, 0
    for k in range(12):                                   # 288 ticks: several episodes incl. time-out ties
        pr, pb = (POL_RANDOM, POL_INSTINCT) if k % 2 == 0 else (POL_INSTINCT_FF, POL_RANDOM)
        c.env.rollout(T, POLNAME[pr], POLNAME[pb], 0, None, rp)
        orp = o.rollout(T, [pr, pb], 0, None, cap, cursor, orp)
        cursor += T * E
    _same_state(c.get_state(), o.get_state())
    assert np.array_equal(c.stats(), o.stats())
    for k, v in dict(state=rp.state_mem, new_state=rp.new_state_mem, action=rp.action_mem, reward=rp.rew_mem).items():
        assert np.allclose(v.cpu().numpy().reshape(orp[k].shape), orp[k], rtol=1e-5, atol=1e-7), k
    assert np.array_equal(rp.done_mem.cpu().numpy().reshape(orp["done"].shape), orp["done"])
    # the single-tick entry point on the same ragged handle: reset everything, then external actions
    rng = np.random.RandomState(E)
    oc, oo = c.reset(), o.reset()
    assert np.allclose(oc, oo, rtol=1e-5, atol=1e-7)
    for t in range(40):
        a = (rng.uniform(-1.3, 1.3, (E, 2 * N, 3)).astype(np.float32) if cont
             else rng.randint(0, 4, (E, 2 * N)).astype(np.int32))
        rc, ro = c.step(a), o.step(a)
        assert np.allclose(rc[0], ro[0], rtol=1e-5, atol=1e-7) and np.allclose(rc[1], ro[1], rtol=1e-5, atol=1e-7), t
        for x, y in zip(rc[2:], ro[2:]):
            assert np.array_equal(x, y), t
    _same_state(c.get_state(), o.get_state())


def test_c_abi_argument_errors(cuda_device):
    import torch
    from battlespace import _lib
    from battlespace._lib import lib
    INV = _lib.BS_E_INVALID

    def msg():
        return lib.bs_last_error().decode()

    cfg = _lib.bs_config()
    C.memset(C.byref(cfg), 0, C.sizeof(cfg))
    h = C.c_void_p()
    for n_agents, n_envs, off, what in [(0, 4, 0, "n_agents"), (9, 4, 0, "n_agents"), (3, 0, 0, "n_envs"), (3, 4, -1, "env_offset")]:
        cfg.n_agents = n_agents
        assert lib.bs_create(C.byref(cfg), C.c_int64(n_envs), C.c_int64(off), 0, C.c_uint64(0), C.byref(h)) == INV
        assert what in msg()
    cfg.n_agents = 3
    assert lib.bs_create(C.byref(cfg), C.c_int64(4), C.c_int64(0), 99, C.c_uint64(0), C.byref(h)) == INV   # no such device
    assert lib.bs_create(None, C.c_int64(4), C.c_int64(0), 0, C.c_uint64(0), C.byref(h)) == INV

    from battlespace import parallel_env
    env = parallel_env(n_agents=3, continuous_actions=True, n_envs=8, seed=0)
    hh, nul = env._h, None
    stats = (C.c_int64 * 8)()
    # rollout: T, teams, policies, cursor
    assert lib.bs_rollout_fused(hh, 0, 0, _lib.POL_RANDOM, _lib.POL_INSTINCT, nul, nul, C.c_int64(0), nul) == INV and "T must" in msg()
    assert lib.bs_rollout_fused(hh, 4, 2, _lib.POL_RANDOM, _lib.POL_INSTINCT, nul, nul, C.c_int64(0), nul) == INV
    assert lib.bs_rollout_fused(hh, 4, 0, 17, _lib.POL_INSTINCT, nul, nul, C.c_int64(0), nul) == INV and "policy" in msg()
    assert lib.bs_rollout_fused(hh, 4, 0, _lib.POL_EXTERNAL, _lib.POL_INSTINCT, nul, nul, C.c_int64(0), nul) == INV   # no ext_actions
    assert lib.bs_rollout_fused(hh, 4, 0, _lib.POL_RANDOM, _lib.POL_INSTINCT, nul, nul, C.c_int64(-1), nul) == INV and "cursor" in msg()
    # the host entry point checks the same thing (it used to launch with a null action pointer)
    assert lib.bs_rollout_fused_host(hh, 4, 0, _lib.POL_EXTERNAL, _lib.POL_INSTINCT, nul, nul, C.c_int64(0), stats) == INV
    assert "ext_actions" in msg()
    # replay ring: rows of different envs must never collide inside one launch
    from battlespace import DeviceReplay
    small = DeviceReplay(4, 2, 3, 11, 3, env.device)                  # fewer rows than envs (8)
    rs = small.as_struct()
    assert lib.bs_rollout_fused(hh, 4, 0, _lib.POL_RANDOM, _lib.POL_INSTINCT, nul, C.byref(rs), C.c_int64(0), nul) == INV
    assert "smaller than n_envs" in msg()
    odd = DeviceReplay(20, 2, 3, 11, 3, env.device)                   # 4 x 8 rows wrap a 20-row ring, 20 % 8 != 0
    ro = odd.as_struct()
    assert lib.bs_rollout_fused(hh, 4, 0, _lib.POL_RANDOM, _lib.POL_INSTINCT, nul, C.byref(ro), C.c_int64(0), nul) == INV
    assert "capacity % n_envs" in msg()
    assert lib.bs_rollout_fused(hh, 2, 0, _lib.POL_RANDOM, _lib.POL_INSTINCT, nul, C.byref(ro), C.c_int64(0), nul) == 0   # 16 rows fit
    with pytest.raises(ValueError):
        env.rollout(4, "random", "instinct", 0, None, odd)
    with pytest.raises(ValueError):                                   # wrong dtype / shape of the learner's action block
        env.rollout(4, "external", "instinct", 0, torch.zeros(4, 8, 3, 3, dtype=torch.float64, device=env.device))
    with pytest.raises(ValueError):
        env.rollout(4, "external", "instinct", 0, torch.zeros(3, 8, 3, 3, device=env.device))
    # single tick: action dtype must fit the action space; null handle
    a = torch.zeros(8, 6, 3, device=env.device)
    assert lib.bs_step(hh, C.c_void_p(a.data_ptr()), _lib.ACT_I32, nul, nul, nul, nul, nul, nul, nul) == INV and "continuous" in msg()
    assert lib.bs_step(None, C.c_void_p(a.data_ptr()), _lib.ACT_F32, nul, nul, nul, nul, nul, nul, nul) == INV
    assert lib.bs_stats(hh, None, nul) == INV and lib.bs_stats(None, stats, nul) == INV
    # frames and the episode log
    assert lib.bs_render(hh, C.c_int64(8), None, None, nul) == INV
    buf = torch.zeros(6 * 100, device=env.device)
    assert lib.bs_episode_log(hh, C.c_void_p(buf.data_ptr()), C.c_int64(100)) == INV and "power of two" in msg()
    # a failed call leaves the handle usable
    env.rollout(4, "random", "instinct")
    assert lib.bs_stats(hh, stats, nul) == 0 and stats[6] == (4 + 2) * 8
    env.close()
