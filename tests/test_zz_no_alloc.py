"""GPU: bs_rollout_actor_static -- the allocation-free variant of bs_rollout_actor (include/battlespace.h) whose work
buffers live in a file-scope `static __device__` array of the library -- against bs_rollout_actor itself.

No counterpart in the reference (added for callers that cannot call cudaMalloc); the file name keeps it LAST in the
collection order so that a failure here cannot hide another test under `pytest -x`.

Both entry points launch the same kernels in the same order, so the bar is the one test_actor_and_sample.py uses
for bs_rollout_actor against the Python loop: bit-equal observations, rewards, dones, winners, replay rows, OU noise
state and counters.  Checked on two calls in a row (the array keeps the previous call's contents), with another
handle of a different shape scribbling over the array in between, with env_done given and left NULL, and with a
handle that is too large for the array (BS_E_INVALID, nothing launched).  The calls go through the C ABI (ctypes).
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

KW = dict(hit_base_reward=1.0, hit_plane_reward=0.9, miss_punishment=-0.02, die_punishment=-0.03, lose_punishment=-0.05)


def _params(N, L, A, dev, seed):
    import torch
    g = torch.Generator(device="cpu").manual_seed(seed)
    u = lambda *s, a=1.0: ((torch.rand(*s, generator=g) * 2 - 1) * a).to(dev)
    return [{"fc1.weight": u(64, L, a=0.125), "fc1.bias": u(64, a=0.125), "bn1.weight": 1 + u(64, a=0.3), "bn1.bias": u(64, a=0.3),
             "fc2.weight": u(64, 64, a=0.125), "fc2.bias": u(64, a=0.125), "bn2.weight": 1 + u(64, a=0.3), "bn2.bias": u(64, a=0.3),
             "pi.weight": u(A, 64, a=0.5), "pi.bias": u(A, a=0.1)} for _ in range(N)]


def _static_call(env, T, actors, team, policy, noise, replay, env_done=True):
    """bs_rollout_actor_static through the C ABI, on the env's own output buffers."""
    from battlespace import _lib
    from battlespace.env import _ptr
    rp = replay.as_struct() if replay is not None else None
    nz = None
    if noise is not None:
        nz = noise.as_struct(None, env.env_done_buf)
        noise.step += int(T)
    rc = _lib.lib.bs_rollout_actor_static(
        env._h, int(T), int(team), actors.as_array(), int(actors.hidden), _lib.POLICIES[policy],
        C.byref(nz) if nz is not None else None, C.byref(rp) if rp is not None else None,
        int(replay.mem_cntr if replay is not None else 0), _ptr(env.obs_buf), _ptr(env.rew_buf), _ptr(env.done_buf),
        _ptr(env.env_done_buf) if env_done else None, _ptr(env.winner_buf), env._stream())
    if rc == 0 and replay is not None:
        replay.mem_cntr += int(T) * env.n_envs
    return rc


@pytest.mark.parametrize("N,cont,E", [(3, True, 4099), (2, False, 1500)])
def test_static_scratch_equals_caller_buffers_two_calls(cuda_device, N, cont, E):
    import torch
    from battlespace import DeviceOUNoise, DeviceReplay, TeamActors, parallel_env
    dev = cuda_device
    L, A, T = 3 * N + 2, 3 if cont else 4, 24
    actors = TeamActors(_params(N, L, A, dev, 31 + N), dev)
    kw = dict(n_agents=N, continuous_actions=cont, n_envs=E, seed=6, **KW)
    a, b, c = parallel_env(**kw), parallel_env(**kw), parallel_env(**kw)
    ra, rb, rc_ = (DeviceReplay(2 * T * E, 64, N, L, A, dev) for _ in range(3))
    na, nb, nc = (DeviceOUNoise(E, N, A, dev, seed=17) for _ in range(3))
    other = parallel_env(n_agents=1, continuous_actions=not cont, n_envs=9001, seed=1, **KW)   # a different shape in the same array
    other_actors = TeamActors(_params(1, 5, 4 if cont else 3, dev, 3), dev)
    for env in (a, b, c, other):
        env.reset_batched()
    for call in range(2):
        a.rollout_actor(T, actors, 0, "instinct", noise=na, replay=ra)              # existing entry point, caller buffers
        torch.cuda.synchronize()
        free0 = torch.cuda.mem_get_info()[0]
        assert _static_call(b, T, actors, 0, "instinct", nb, rb, env_done=True) == 0
        assert _static_call(c, T, actors, 0, "instinct", nc, rc_, env_done=False) == 0
        torch.cuda.synchronize()
        if call == 1:   # (the first call may load kernel images lazily; from the second on the free memory must not move)
            assert torch.cuda.mem_get_info()[0] == free0, "device memory was allocated or freed inside bs_rollout_actor_static"
        for x, y, z in ((a.obs_buf, b.obs_buf, c.obs_buf), (a.rew_buf, b.rew_buf, c.rew_buf), (a.done_buf, b.done_buf, c.done_buf),
                        (a.winner_buf, b.winner_buf, c.winner_buf), (na.state, nb.state, nc.state),
                        (ra.state_mem, rb.state_mem, rc_.state_mem), (ra.new_state_mem, rb.new_state_mem, rc_.new_state_mem),
                        (ra.action_mem, rb.action_mem, rc_.action_mem), (ra.rew_mem, rb.rew_mem, rc_.rew_mem),
                        (ra.done_mem, rb.done_mem, rc_.done_mem)):
            assert torch.equal(x, y), call
            assert torch.equal(x, z), call
        assert torch.equal(a.env_done_buf, b.env_done_buf), call
        assert np.array_equal(a.stats(), b.stats()) and np.array_equal(a.stats(), c.stats()), call
        # the next call must not depend on what the array holds: let another handle overwrite it
        assert _static_call(other, 5, other_actors, 1, "random", None, None, env_done=False) == 0
        torch.cuda.synchronize()
        # c keeps no env_done of its own between calls: hand it a's, as a caller without the output would have to
        c.env_done_buf.copy_(a.env_done_buf)
    assert a.stats()[3] > 0 and ra.mem_cntr == rb.mem_cntr == rc_.mem_cntr == 2 * T * E


def test_static_scratch_rejects_oversized_handle(cuda_device):
    from battlespace import TeamActors, _lib, parallel_env
    dev = cuda_device
    N, E = 3, 240_000                                     # 240,000 * 3 * 3 * 4 = 8,640,000 B > 8 MiB
    assert E * N * 3 * 4 > _lib.STATIC_SCRATCH_BYTES
    env = parallel_env(n_agents=N, continuous_actions=True, n_envs=E, seed=2, **KW)
    env.reset_batched()
    actors = TeamActors(_params(N, 11, 3, dev, 5), dev)
    before = env.stats().copy()
    rc = _static_call(env, 1, actors, 0, "instinct", None, None)
    assert rc == _lib.BS_E_INVALID
    msg = (_lib.lib.bs_last_error() or b"").decode()
    assert "BS_STATIC_SCRATCH_BYTES" in msg and str(_lib.STATIC_SCRATCH_BYTES) in msg, msg
    assert np.array_equal(env.stats(), before)            # nothing ran
    env.rollout_actor(1, actors, 0, "instinct")           # the caller-buffer entry point still takes this size
    assert env.stats()[6] == before[6] + E
